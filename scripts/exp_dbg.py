import sys, os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.argv = [sys.argv[0]]
import importlib.util, types
src = open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "exp_gemm_overhead.py")).read()
src = src[:src.index("for dbg in (0, 3):")]
exec(compile(src, "exp", "exec"))
for dbg in (0, 1, 2, 3):
    run("o_proj", 4096, 4096, 32, debug=dbg)
for dbg in (0, 1, 2, 3):
    run("o_proj", 4096, 4096, 32, T=16, debug=dbg)
