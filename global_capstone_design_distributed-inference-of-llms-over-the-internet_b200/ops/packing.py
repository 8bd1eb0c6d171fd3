"""Weight layout for the swap-AB tcgen05 GEMM (csrc/gemm.cu).

A torch ``Linear`` weight ``W[N, K]`` (row = output feature, K contiguous) is
re-tiled once at load time into the exact shared-memory image the UMMA
descriptor expects, so that a k-block of a tile is ONE contiguous 16 KiB run in
HBM and can be fetched with a single ``cp.async.bulk``:

    packed[n_tile][k_block][a][128 rows][8 chunks][8 elems]      (16-bit elems)

* ``n_tile``  — 128 output rows, ``k_block`` — 64 input features (128 bytes);
* inside a block, row ``r`` keeps its 8 sixteen-byte chunks at position
  ``chunk ^ (r % 8)`` — the hardware 128-byte swizzle (bank-conflict-free for
  both the bulk copy and the tensor core reads);
* ``a`` is 1, or 2 when two matrices are interleaved per tile (``gate``/``up``
  for the fused SwiGLU epilogue: both accumulators of an output feature end up
  in the same thread).

N is zero-padded to a multiple of 128 and K to a multiple of 64.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch

TILE_N = 128
BLOCK_K = 64


@dataclass
class PackedWeight:
    data: torch.Tensor          # flat 16-bit tensor in the layout above
    n: int                      # logical output features
    k: int                      # logical input features
    k_padded: int
    n_tiles: int
    dual: bool
    bias: Optional[torch.Tensor] = None
    fp8: bool = False
    scale: Optional[torch.Tensor] = None   # fp8: fp32 [n_tiles*128] (dual: [2 * n_tiles*128], gate then up)
    mx: bool = False                       # fp8 with OCP-MX block scales (UE8M0 per 32 elements) packed behind each k-block

    @property
    def kb_per_tile(self) -> int:
        return self.k_padded // (2 * BLOCK_K if self.fp8 else BLOCK_K)

    @property
    def nbytes(self) -> int:
        return self.data.numel() * self.data.element_size()


def _tile(w: torch.Tensor) -> torch.Tensor:
    """[N, K] -> [n_tiles, kb, 128, 64] swizzled; pads N and K."""
    n, k = w.shape
    n_pad = (n + TILE_N - 1) // TILE_N * TILE_N
    k_pad = (k + BLOCK_K - 1) // BLOCK_K * BLOCK_K
    if n_pad != n or k_pad != k:
        wp = w.new_zeros((n_pad, k_pad))
        wp[:n, :k] = w
        w = wp
    nt, kb = n_pad // TILE_N, k_pad // BLOCK_K
    blk = w.view(nt, TILE_N, kb, 8, 8).permute(0, 2, 1, 3, 4)  # [nt, kb, 128, 8 chunks, 8]
    r = torch.arange(TILE_N, device=w.device)
    c = torch.arange(8, device=w.device)
    src_chunk = c[None, :] ^ (r[:, None] % 8)                   # position p holds chunk p ^ (r%8)
    blk = blk[:, :, r[:, None], src_chunk, :]                  # [nt, kb, 128, 8, 8]
    return blk.contiguous().view(nt, kb, TILE_N, BLOCK_K)


def pack_weight(w: torch.Tensor, bias: Optional[torch.Tensor] = None) -> PackedWeight:
    """Pack one ``[N, K]`` weight (fp16/bf16)."""
    assert w.dim() == 2 and w.dtype in (torch.float16, torch.bfloat16)
    n, k = w.shape
    t = _tile(w)
    return PackedWeight(t.reshape(-1), n, k, t.shape[1] * BLOCK_K, t.shape[0], False,
                        None if bias is None else bias.contiguous())


def pack_dual(w_gate: torch.Tensor, w_up: torch.Tensor) -> PackedWeight:
    """Interleave two same-shaped weights tile by tile (SwiGLU gate/up)."""
    assert w_gate.shape == w_up.shape and w_gate.dtype == w_up.dtype
    n, k = w_gate.shape
    g, u = _tile(w_gate), _tile(w_up)
    t = torch.stack((g, u), dim=2)                             # [nt, kb, 2, 128, 64]
    return PackedWeight(t.contiguous().reshape(-1), n, k, g.shape[1] * BLOCK_K, g.shape[0], True)


def unpack_weight(p: PackedWeight) -> torch.Tensor:
    """Inverse of :func:`pack_weight` / :func:`pack_dual` (tests, re-sharding)."""
    a = 2 if p.dual else 1
    kb = p.kb_per_tile
    t = p.data.view(p.n_tiles, kb, a, TILE_N, 8, 8)
    r = torch.arange(TILE_N, device=t.device)
    c = torch.arange(8, device=t.device)
    src = c[None, :] ^ (r[:, None] % 8)
    t = t[:, :, :, r[:, None], src, :]                         # swizzle is an involution
    t = t.reshape(p.n_tiles, kb, a, TILE_N, BLOCK_K).permute(2, 0, 3, 1, 4)
    full = t.reshape(a, p.n_tiles * TILE_N, kb * BLOCK_K)[:, : p.n, : p.k]
    return full if p.dual else full[0]


# ----------------------------------------------------------------------------- fp8 (opt-in)
FP8_MAX = 448.0
BLOCK_K8 = 128          # one 128-byte swizzle row holds 128 e4m3 elements


def _quantize_rows(w: torch.Tensor):
    amax = w.float().abs().amax(dim=1).clamp_min(1e-12)
    scale = amax / FP8_MAX
    q = (w.float() / scale[:, None]).to(torch.float8_e4m3fn)
    return q, scale


def _tile8(q: torch.Tensor) -> torch.Tensor:
    """e4m3 [N, K] -> [n_tiles, kb, 128, 128] bytes, 16-byte chunks swizzled like the 16-bit layout."""
    n, k = q.shape
    n_pad = (n + TILE_N - 1) // TILE_N * TILE_N
    k_pad = (k + BLOCK_K8 - 1) // BLOCK_K8 * BLOCK_K8
    b = q.view(torch.uint8)
    if n_pad != n or k_pad != k:
        bp = b.new_zeros((n_pad, k_pad))
        bp[:n, :k] = b
        b = bp
    nt, kb = n_pad // TILE_N, k_pad // BLOCK_K8
    blk = b.view(nt, TILE_N, kb, 8, 16).permute(0, 2, 1, 3, 4)
    r = torch.arange(TILE_N, device=q.device)
    c = torch.arange(8, device=q.device)
    src = c[None, :] ^ (r[:, None] % 8)
    return blk[:, :, r[:, None], src, :].contiguous().view(nt, kb, TILE_N, BLOCK_K8)


def pack_weight_fp8(w: torch.Tensor, bias: Optional[torch.Tensor] = None) -> PackedWeight:
    """Per-output-row scaled e4m3 copy of ``w`` in the GEMM's tiled layout."""
    n, k = w.shape
    q, scale = _quantize_rows(w)
    t = _tile8(q)
    sc = torch.zeros(t.shape[0] * TILE_N, dtype=torch.float32, device=w.device)
    sc[:n] = scale
    return PackedWeight(t.reshape(-1), n, k, t.shape[1] * BLOCK_K8, t.shape[0], False,
                        None if bias is None else bias.contiguous(), True, sc)


def pack_dual_fp8(w_gate: torch.Tensor, w_up: torch.Tensor) -> PackedWeight:
    n, k = w_gate.shape
    qg, sg = _quantize_rows(w_gate)
    qu, su = _quantize_rows(w_up)
    g, u = _tile8(qg), _tile8(qu)
    t = torch.stack((g, u), dim=2)
    npad = g.shape[0] * TILE_N
    sc = torch.zeros(2 * npad, dtype=torch.float32, device=w_gate.device)
    sc[:n] = sg
    sc[npad:npad + n] = su
    return PackedWeight(t.contiguous().reshape(-1), n, k, g.shape[1] * BLOCK_K8, g.shape[0], True, None, True, sc)


def dequantize_fp8(w: torch.Tensor) -> torch.Tensor:
    """What the fp8 GEMM effectively multiplies by (tests): per-row quantise -> dequantise."""
    q, scale = _quantize_rows(w)
    return q.float() * scale[:, None]


# ----------------------------------------------------------------------------- MX fp8 (opt-in, block-scaled)
MX_BLOCK = 32            # elements that share one UE8M0 scale
MX_SF_PAD = 1024         # bytes of scale factors (+ padding) behind the weights of a k-block


def _mx_quantize(w: torch.Tensor):
    """[N, K] (K % 32 == 0) -> (e4m3 values, UE8M0 scale bytes [N, K/32]).  scale = 2^ceil(log2(amax / 448))."""
    n, k = w.shape
    blk = w.float().view(n, k // MX_BLOCK, MX_BLOCK)
    amax = blk.abs().amax(dim=2)
    m, ex = torch.frexp(amax / FP8_MAX)                     # amax/448 = m * 2^ex, m in [0.5, 1)
    ex = torch.where(m == 0.5, ex - 1, ex)
    ex = torch.where(amax > 0, ex, torch.full_like(ex, -127)).clamp(-127, 126)
    q = (blk * torch.exp2(-ex.float())[:, :, None]).to(torch.float8_e4m3fn).view(n, k)
    return q, (ex + 127).to(torch.uint8)


def mx_dequantize(w: torch.Tensor) -> torch.Tensor:
    """What the MX GEMM effectively multiplies by (tests): block-quantise -> dequantise, fp32."""
    n, k = w.shape
    k_pad = (k + BLOCK_K8 - 1) // BLOCK_K8 * BLOCK_K8
    wp = w.new_zeros((n, k_pad))
    wp[:, :k] = w
    q, sf = _mx_quantize(wp)
    return (q.float().view(n, k_pad // MX_BLOCK, MX_BLOCK) * torch.exp2(sf.float() - 127)[:, :, None]).view(n, k_pad)[:, :k]


def _mx_tiles(w: torch.Tensor):
    """-> (weights [nt, kb, 128, 128] uint8 swizzled, scale atoms [nt, kb, 512] uint8)."""
    n, k = w.shape
    n_pad = (n + TILE_N - 1) // TILE_N * TILE_N
    k_pad = (k + BLOCK_K8 - 1) // BLOCK_K8 * BLOCK_K8
    wp = w.new_zeros((n_pad, k_pad))
    wp[:n, :k] = w
    q, sf = _mx_quantize(wp)
    t = _tile8(q)                                                   # [nt, kb, 128, 128]
    nt, kb = t.shape[0], t.shape[1]
    # scale atom of a (tile, k-block): byte (row % 32) * 16 + (row / 32) * 4 + group  <- sf[row, kb*4 + group]
    a = sf.view(nt, 4, 32, kb, 4)                                   # [tile, row/32, row%32, kb, group]
    atom = torch.zeros(nt, kb, 32, 16, dtype=torch.uint8, device=w.device)
    atom.view(nt, kb, 32, 4, 4).copy_(a.permute(0, 3, 2, 1, 4))     # [tile, kb, row%32, row/32, group]
    return t, atom.view(nt, kb, 512)


def _mx_assemble(mats, atoms) -> torch.Tensor:
    """k-block stream: [weights of each matrix (16 KiB)] [512 B of scales per matrix] [pad to 1 KiB]."""
    nt, kb = mats[0].shape[0], mats[0].shape[1]
    w = torch.stack([m.reshape(nt, kb, TILE_N * BLOCK_K8) for m in mats], dim=2).reshape(nt, kb, -1)
    sf = torch.zeros(nt, kb, MX_SF_PAD, dtype=torch.uint8, device=w.device)
    for i, a in enumerate(atoms):
        sf[:, :, i * 512:(i + 1) * 512] = a
    return torch.cat((w, sf), dim=2).contiguous().reshape(-1)


def pack_weight_mx(w: torch.Tensor, bias: Optional[torch.Tensor] = None) -> PackedWeight:
    """Block-scaled (OCP MX) e4m3 copy of ``w`` for ``tcgen05.mma.kind::mxf8f6f4.block_scale``."""
    n, k = w.shape
    t, atom = _mx_tiles(w)
    return PackedWeight(_mx_assemble([t], [atom]), n, k, t.shape[1] * BLOCK_K8, t.shape[0], False,
                        None if bias is None else bias.contiguous(), True, None, True)


def pack_dual_mx(w_gate: torch.Tensor, w_up: torch.Tensor) -> PackedWeight:
    n, k = w_gate.shape
    g, ag = _mx_tiles(w_gate)
    u, au = _mx_tiles(w_up)
    return PackedWeight(_mx_assemble([g, u], [ag, au]), n, k, g.shape[1] * BLOCK_K8, g.shape[0], True, None, True, None, True)
