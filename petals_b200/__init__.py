"""Importable alias of the framework package.

The source tree lives in the directory named after the project
(``global_capstone_design_distributed-inference-of-llms-over-the-internet_b200``);
a hyphenated directory cannot be spelled in an ``import`` statement, so this
alias points its ``__path__`` at that directory and executes its ``__init__``.
``import petals_b200.ops.gemm`` therefore loads
``<project dir>/ops/gemm.py``.
"""
import os as _os

_ROOT = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
PACKAGE_DIR = _os.path.join(
    _ROOT, "global_capstone_design_distributed-inference-of-llms-over-the-internet_b200"
)
__path__ = [PACKAGE_DIR]
with open(_os.path.join(PACKAGE_DIR, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(PACKAGE_DIR, "__init__.py"), "exec"))
del _os, _f
