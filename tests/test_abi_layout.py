"""The ctypes mirrors of the C ABI structs must have the layout the C compiler gives the header (CPU test:
compiles a tiny C program against include/b200_petals.h with the host compiler, no CUDA needed)."""
import ctypes
import os
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

PROG = r"""
#include <stdio.h>
#include <stddef.h>
#include "b200_petals.h"
int main(void) {
    printf("gemm %zu %zu %zu %zu %zu\n", sizeof(b200_gemm_args), offsetof(b200_gemm_args, w_scale),
           offsetof(b200_gemm_args, ll_buf), offsetof(b200_gemm_args, rstd), offsetof(b200_gemm_args, chain));
    printf("chain %zu %zu %zu\n", sizeof(b200_ll_chain), offsetof(b200_ll_chain, ldx_ll), offsetof(b200_ll_chain, half));
    printf("tp %zu %zu %zu\n", sizeof(b200_tp_args), offsetof(b200_tp_args, seq), offsetof(b200_tp_args, mode));
    printf("attn %zu %zu %zu %zu\n", sizeof(b200_attn_decode_args), offsetof(b200_attn_decode_args, scale),
           offsetof(b200_attn_decode_args, ll_scratch), offsetof(b200_attn_decode_args, chain));
    printf("hop %zu %zu %zu\n", sizeof(b200_hop_args), offsetof(b200_hop_args, counter), offsetof(b200_hop_args, pdl));
    return 0;
}
"""


@pytest.mark.skipif(shutil.which("gcc") is None, reason="needs a host C compiler")
def test_ctypes_structs_match_the_header(tmp_path):
    from petals_b200.ops._prototypes import AttnDecodeArgs, HopArgs
    from petals_b200.ops.native import GemmArgs, LLChainArgs
    from petals_b200.parallel.tensor_parallel import TpArgs

    src = tmp_path / "abi.c"
    src.write_text(PROG)
    exe = tmp_path / "abi"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    out = dict((ln.split()[0], [int(v) for v in ln.split()[1:]])
               for ln in subprocess.check_output([str(exe)], text=True).strip().splitlines())
    assert out["gemm"] == [ctypes.sizeof(GemmArgs), GemmArgs.w_scale.offset, GemmArgs.ll_buf.offset, GemmArgs.rstd.offset,
                           GemmArgs.chain.offset]
    assert out["chain"] == [ctypes.sizeof(LLChainArgs), LLChainArgs.ldx_ll.offset, LLChainArgs.half.offset]
    assert out["tp"] == [ctypes.sizeof(TpArgs), TpArgs.seq.offset, TpArgs.mode.offset]
    assert out["attn"] == [ctypes.sizeof(AttnDecodeArgs), AttnDecodeArgs.scale.offset, AttnDecodeArgs.ll_scratch.offset,
                           AttnDecodeArgs.chain.offset]
    assert out["hop"] == [ctypes.sizeof(HopArgs), HopArgs.counter.offset, HopArgs.pdl.offset]


def test_library_loads_with_every_symbol_resolved():
    """The in-tree library must dlopen with immediate binding (a dangling reference between translation units only
    shows at load time) and export every entry point the Python layer registers; no GPU needed for either."""
    import ctypes
    import os

    from petals_b200.ops import native

    if not os.path.exists(native.LIB_PATH):
        import pytest

        pytest.skip("library not built (python csrc/build.py)")
    ctypes.CDLL(native.LIB_PATH, mode=os.RTLD_NOW)
    L = native.lib()                       # registers all prototypes: a missing export raises AttributeError
    assert L.b200_version() > 0
