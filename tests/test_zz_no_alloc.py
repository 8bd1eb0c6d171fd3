"""b200_gemm_static_scratch — the GEMM entry point for hosts that cannot allocate device memory.

Some users of libb200petals.so only hand it pointers: a plugin inside a server whose allocator
already owns the whole GPU, or a token loop that must not hit a device allocation.  They cannot
provide the stream-K workspace b200_gemm wants, so this entry point keeps a fixed workspace inside
the library.  The test drives it purely through the C ABI (ctypes, raw pointers) and checks that it
returns what the ordinary entry point returns — twice in a row, because the second call runs on the
workspace contents the first one left behind.

This file sorts last on purpose: with ``pytest -x`` every other test has already run if it fails.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _args(native, x, pw, out, scratch=None, counters=None, residual=None):
    a = native.GemmArgs()
    a.w_packed = pw.data.data_ptr()
    a.x = x.data_ptr()
    a.out = out.data_ptr()
    a.residual = residual.data_ptr() if residual is not None else None
    a.bias = pw.bias.data_ptr() if pw.bias is not None else None
    a.scratch = scratch.data_ptr() if scratch is not None else None
    a.counters = counters.data_ptr() if counters is not None else None
    a.T, a.N, a.K = x.shape[0], pw.n, x.shape[1]
    a.ldx, a.ldo = x.stride(0), out.stride(0)
    a.ldr = residual.stride(0) if residual is not None else 0
    a.n_tiles = pw.n_tiles
    a.dual = 1 if pw.dual else 0
    a.epilogue = 2 if pw.dual else 0
    a.dtype = 1 if x.dtype == torch.bfloat16 else 0
    return a


@pytest.mark.parametrize("T,N,K,dual", [(1, 512, 2048, False), (16, 4096, 1024, False), (5, 1408, 1024, True)])
def test_static_scratch_matches_regular_entry_point(T, N, K, dual):
    from petals_b200.ops import native, packing

    L = native.lib()
    fn = L.b200_gemm_static_scratch
    fn.argtypes = [ctypes.POINTER(native.GemmArgs), ctypes.c_void_p]
    fn.restype = ctypes.c_int
    dtype = torch.bfloat16
    torch.manual_seed(T + N)
    x = (torch.randn(T, K, device="cuda") * 0.5).to(dtype)
    mk = lambda: (torch.randn(N, K, device="cuda") / K ** 0.5).to(dtype)  # noqa: E731
    pw = packing.pack_dual(mk(), mk()) if dual else packing.pack_weight(mk())
    res = None if dual else torch.randn(T, N, device="cuda").to(dtype)
    stream = torch.cuda.current_stream().cuda_stream

    # the ordinary entry point with a caller-owned workspace is the oracle
    scratch = torch.empty(L.b200_gemm_scratch_bytes(16) // 4, dtype=torch.float32, device="cuda")
    counters = torch.zeros(4096, dtype=torch.int32, device="cuda")
    ref = torch.empty(T, N, device="cuda", dtype=dtype)
    native.check(L.b200_gemm(ctypes.byref(_args(native, x, pw, ref, scratch, counters, res)), stream), "b200_gemm")

    outs = []
    for _ in range(2):  # second call starts from the first call's leftovers in the shared array
        out = torch.full((T, N), float("nan"), device="cuda", dtype=dtype)
        native.check(fn(ctypes.byref(_args(native, x, pw, out, None, None, res)), stream), "b200_gemm_static_scratch")
        outs.append(out)
    torch.cuda.synchronize()
    scale = ref.float().abs().max().item() + 1e-6
    for out in outs:
        err = (out.float() - ref.float()).abs().max().item()
        assert err / scale < 2e-2, f"rel err {err / scale:.3g}"
    if not dual:   # and against plain torch in fp32, so that the oracle is not only this library's other entry point
        w = packing.unpack_weight(pw).float()
        want = (x.float() @ w.t()).to(dtype).float() + res.float()
        err = (outs[1].float() - want).abs().max().item()
        assert err / (want.abs().max().item() + 1e-6) < 2e-2


def test_static_scratch_rejects_oversize():
    from petals_b200.ops import native, packing

    L = native.lib()
    fn = L.b200_gemm_static_scratch
    fn.argtypes = [ctypes.POINTER(native.GemmArgs), ctypes.c_void_p]
    fn.restype = ctypes.c_int
    x = torch.randn(17, 128, device="cuda").to(torch.float16)
    pw = packing.pack_weight(torch.randn(128, 128, device="cuda").to(torch.float16))
    out = torch.empty(17, 128, device="cuda", dtype=torch.float16)
    rc = fn(ctypes.byref(_args(native, x, pw, out)), torch.cuda.current_stream().cuda_stream)
    assert rc == -3
    assert b"exceeds the built-in workspace limit" in L.b200_last_error()
