"""`bench.py --impl reference`: the reference arm.

Install outcome (recorded in DESIGN.md §Reference arm): ``pip install --no-index
--no-build-isolation --find-links /opt/wheelhouse --target baseline/_ref /root/reference`` fails
("Neither 'setup.py' nor 'pyproject.toml' found": the reference is a bare source tree), also
with ``--no-deps`` from a /tmp copy.  Its public entry point ``python -m src.main`` imports
hivemind 1.1.11 (+ the Go p2pd daemon) at module top, which is not installable without network.
"""
from __future__ import annotations

import json
import os

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def unavailable(reason: str) -> None:
    print(json.dumps({"impl": "reference", "unavailable": reason}), flush=True)


def run_reference(args) -> None:
    try:
        from .reference_harness import run as run_harness
    except Exception as e:  # noqa: BLE001
        unavailable(f"reference not installable (no setup.py/pyproject.toml; hivemind + p2pd need network): {e}")
        return
    try:
        run_harness(args)
    except SystemExit:
        raise
    except Exception as e:  # noqa: BLE001
        unavailable(f"reference harness failed: {type(e).__name__}: {e}"[:300])
