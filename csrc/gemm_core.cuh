// Shared device-side pieces of the tcgen05 GEMM: tile configuration, kernel parameters and the
// fused epilogue of the stream-K kernel (gemm.cu).
#pragma once
#include <cuda_fp8.h>

#include "common.h"
#include "ptx.cuh"
#include "llchain.cuh"

#ifndef B200_SA16
#define B200_SA16 3        /* 32 KiB weight stages for the decode tile (BN = 16) */
#endif
#ifndef B200_GROUP16
#define B200_GROUP16 2     /* k-blocks per activation stage for the decode tile */
#endif
#ifndef B200_MINBLK16
#define B200_MINBLK16 2   /* CTAs per SM the decode tile is compiled for */
#endif

namespace b200 {

constexpr int kTileM = 128;            // weight rows per tile (UMMA M)
constexpr int kBK = 64;                // k elements per block = one 128 B swizzle row
constexpr int kABytes = kTileM * 128;  // bytes of one packed weight block
constexpr int kThreads = 192;
constexpr int kEpiThreads = 128;

// MX = 1: block-scaled e4m3 weights (OCP MX: one UE8M0 scale per 32 elements along k).  A k-block is then 128
// elements (one 128-byte swizzle row) and carries its scale factors behind the weights: 512 bytes per matrix in the
// order tcgen05.cp expects ((row % 32) * 16 + (row / 32) * 4 + 32-element group), padded to 1 KiB per k-block so that
// every block stays 1024-byte aligned - weights and scales of a k-block still arrive in ONE bulk copy.
constexpr int kMxSfBytes = 512;        // scale factors of one 128-row matrix for one k-block
constexpr int kMxSfPad = 1024;         // scale-factor area of a k-block (1 or 2 matrices)
constexpr int kMxSfSlots = 4;          // k-blocks whose scale factors can sit in tensor memory at once

template <int BN, int DUAL, int MX = 0>
struct GemmCfg {
    static constexpr int kAPer = DUAL ? 2 : 1;
    static constexpr int kBlockA = kABytes * kAPer + (MX ? kMxSfPad : 0);   // bytes of one k-block of the weight stream
    // One weight-ring stage = one 32 KiB bulk copy.  The single-thread loader and
    // MMA loops pay ~100+ ns of barrier latency per stage, so stages must be big.
    static constexpr int kKps = DUAL ? 1 : 2;              // k-blocks per weight stage
    static constexpr int kStageA = kBlockA * kKps;         // 32768
    static constexpr int kSA = (BN <= 16) ? B200_SA16 : (BN == 256) ? 3 : 4;
    static constexpr int kGroup = MX ? 1 : (BN == 16) ? B200_GROUP16 : (BN == 32) ? 2 : 1;  // k-blocks per activation stage
    static constexpr int kSubB = BN * 128;                 // one k-block of activations
    static constexpr int kStageB = kSubB * kGroup;
    static constexpr int kSB = MX ? 3 : (BN == 16) ? 3 : (BN == 32) ? 4 : (BN == 64) ? 6 : (BN == 128) ? 4 : 3;
    static constexpr bool kPrefetch = BN <= 32;            // register double-buffering of the next B stage
    // Back-to-back tcgen05.mma into ONE accumulator serialise on the tensor-pipe latency
    // (~180 cycles each when N is tiny), so narrow tiles rotate the four k-steps of a
    // k-block over independent accumulators that the epilogue adds up.
    static constexpr int kNAcc = (BN <= 32) ? 4 : (BN == 64) ? 2 : 1;
    // MX: behind the accumulators, kMxSfSlots sets of scale-factor columns (4 per weight matrix + 4 for the activations)
    static constexpr int kSfCols = MX ? kMxSfSlots * (kAPer * 4 + 4) : 0;
    static constexpr int kTmemColsRaw = BN * kAPer * kNAcc + kSfCols;
    static constexpr int kTmemCols = kTmemColsRaw <= 32 ? 32 : kTmemColsRaw <= 64 ? 64 : kTmemColsRaw <= 128 ? 128 : kTmemColsRaw <= 256 ? 256 : 512;
    static constexpr int kMinBlocks = (BN <= 16) ? B200_MINBLK16 : 1;
    static constexpr size_t kSmemData = (size_t)kSA * kStageA + (size_t)kSB * kStageB;
    // MX: + the activations' scale factors of every activation stage (one 512-byte atom per k-block in flight)
    static constexpr size_t kSfbBytes = MX ? (size_t)kSB * kGroup * kMxSfBytes : 0;
    static constexpr size_t kSmemBytes = kSmemData + 1024 /*align*/ + 1024 /*barriers, stats*/ + kSfbBytes;
};

struct GemmParams {
    const uint8_t* __restrict__ w;
    const uint16_t* __restrict__ x;
    void* __restrict__ out;
    const uint16_t* __restrict__ residual;
    const uint16_t* __restrict__ bias;
    const uint16_t* __restrict__ norm_w;
    const uint16_t* __restrict__ norm_b;
    float* __restrict__ scratch;
    int* __restrict__ counters;
    int T, N, K;
    int ldx, ldo, ldr;
    int n_tiles, t_tiles, kbpt;
    long long Q;  // total k-blocks = n_tiles * t_tiles * kbpt
    int epilogue, norm_mode;
    float norm_eps;
    int is_bf16, out_fp32;
    // tensor-parallel all-reduce fused into the tile epilogue (tp_world <= 1: off)
    int tp_world, tp_rank, tp_n_max, tp_tiles_max;
    float* tp_stage[8];   // peer-mapped staging buffers: [2 parity][world][16 tokens][tp_n_max] fp32 on each rank
    int* tp_flags[8];     // peer-mapped arrival flags:   [2 parity][world][tp_tiles_max]
    int* tp_seq;          // local: number of fused launches completed on this workspace
    int* tp_exit;         // local: CTAs that have left the current launch
    int tp_defer;         // 1: push partial tiles only; the CONSUMER kernel waits for them and adds them up
    // consumer side ("gather" prologue): rebuild x = residual + bias + sum_r partial_r before staging it
    int gather_world, gather_rank, gather_n_max, gather_tiles_max;
    const float* gather_stage;        // this rank's staging buffer
    const int* gather_flags;          // this rank's arrival flags
    const int* gather_seq;            // launch number of the producer (read after it has completed)
    const uint16_t* gather_residual;  // [T, K] 16-bit
    const uint16_t* gather_bias;      // [K] 16-bit or null
    int gather_ldr;
    int is_fp8;                       // e4m3 weights + on-the-fly e4m3 activations (kind::f8f6f4)
    const float* __restrict__ w_scale;  // fp8: per-output-row dequantisation scale
    // Fix-up of split tiles without tickets (decode, T <= 4): the grid gives every tile exactly `ll_parts` CTAs;
    // the first ll_parts-1 of them publish their partial sums as (value, tag) 8-byte words and the last one
    // polls for them and clears them again (no fence, no atomic, no counter) - see b200_gemm_args.ll_buf.
    unsigned long long* ll_words;
    int ll_parts;
    const float* rstd_in;       // precomputed per-token 1/rms (b200_rms_rstd) or null
    const uint8_t* pf_ptr[2];   // L2 warm-up ranges for the kernels that follow (see b200_gemm_args.prefetch)
    long long pf_bytes[2];
    int debug;  // experiment switches (scripts/exp_*.py): 1 = no MMA issue, 2 = no B stage waits
    LLChain ch; // flag-chained decode step (ch.seq != null): x / residual polled as LL words, output published as LL words
};

__device__ __forceinline__ long long range_begin(const GemmParams& p, int c) {
    return (long long)c * p.Q / (long long)gridDim.x;
}

__device__ __forceinline__ float gelu_tanh(float x) {
    const float k0 = 0.7978845608028654f, k1 = 0.044715f;
    return 0.5f * x * (1.0f + tanhf(k0 * (x + k1 * x * x * x)));
}

// ---- fused all-reduce of one finished 128-row tile across the tensor-parallel group ------------
// Every rank holds a partial sum (its K shard) of the same output tile.  The finisher of the tile
//   1. pushes its fp32 partial into EVERY rank's staging buffer with peer stores over NVLink, each value
//      paired with the launch sequence number in one 8-byte word (no fences, no separate flags);
//   2. polls the other ranks' words of this tile in LOCAL memory until they carry the same number;
//   3. adds the partials in rank order — every rank computes bit-identical sums — into vg.
// Staging buffers and flags alternate between two halves by launch parity, so a rank that is one
// launch ahead never overwrites data a slower peer is still reading (a rank cannot get two ahead:
// finishing a launch requires every peer's pushes for it).
__device__ __forceinline__ float round16(float v, int is_bf16);

// "LL" exchange words (as in NCCL's low-latency protocol): a 4-byte value and the 4-byte launch sequence
// number travel in ONE 8-byte store, so the receiver needs no fence and no separate flag — it polls the
// word until the sequence number matches.  Sequence numbers never repeat for a slot.
__device__ __forceinline__ void ll_store(uint2* dst, float v, int seq) {
    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};\n" ::"l"(dst), "r"(__float_as_uint(v)), "r"(seq) : "memory");
}
__device__ __forceinline__ float ll_load(const uint2* src, int seq) {
    uint32_t a, b;
    long long t0 = 0;
    for (;;) {
        asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];\n" : "=r"(a), "=r"(b) : "l"(src) : "memory");
        if ((int)b == seq) break;
        if (t0 == 0) t0 = clock64();
        else if (clock64() - t0 > 8000000000LL) {
            printf("b200: tensor-parallel exchange timed out (block %d thread %d, seq %d, saw %d)\n", blockIdx.x,
                   threadIdx.x, seq, (int)b);
            __trap();
        }
    }
    return __uint_as_float(a);
}

// Consumer half of the deferred exchange (all-gather -> GEMM fusion): executed by the 128 worker threads
// of EVERY CTA of the next GEMM before they stage activations.  Waits until all ranks' partial tiles of
// the producer launch have landed in local memory, then writes x[t, :] = residual + bias + sum over
// ranks (rank order) as 16-bit values.  All CTAs write identical values, so the redundancy is benign;
// it costs two L2 round trips instead of an exposed NVLink round trip at the producer's tail.
__device__ __forceinline__ void tp_gather_prologue(const GemmParams& p, int e) {
    const int R = p.gather_world, seq = __ldcv(p.gather_seq), par = seq & 1;
    const int H = p.K;
    const uint2* stage = reinterpret_cast<const uint2*>(p.gather_stage);
    uint16_t* x = const_cast<uint16_t*>(p.x);
    for (int t = 0; t < p.T; ++t) {
        for (int n = e * 2; n < H; n += 256) {
            float a = 0.f, b = 0.f;
            for (int r = 0; r < R; ++r) {
                const uint2* src = stage + ((size_t)(par * R + r) * 16 + t) * p.gather_n_max + n;
                a += ll_load(src, seq);
                b += ll_load(src + 1, seq);
            }
            if (p.gather_bias) {
                a += h16_to_float(p.gather_bias[n], p.is_bf16);
                b += h16_to_float(p.gather_bias[n + 1], p.is_bf16);
            }
            const uint32_t rw = *reinterpret_cast<const uint32_t*>(p.gather_residual + (size_t)t * p.gather_ldr + n);
            // same rounding sequence as the in-producer epilogue: round the linear's output, then add the residual
            a = round16(a, p.is_bf16) + h16_to_float((uint16_t)(rw & 0xffff), p.is_bf16);
            b = round16(b, p.is_bf16) + h16_to_float((uint16_t)(rw >> 16), p.is_bf16);
            *reinterpret_cast<uint32_t*>(x + (size_t)t * p.ldx + n) =
                (uint32_t)float_to_h16(a, p.is_bf16) | ((uint32_t)float_to_h16(b, p.is_bf16) << 16);
        }
    }
    named_bar_sync(1, 128);
}

__device__ __forceinline__ void tp_allreduce_tile(const GemmParams& p, int tile, int row, int ncol, int e, int seq,
                                                  float (&vg)[16]) {
    const int R = p.tp_world, me = p.tp_rank, par = seq & 1;
    const size_t slot = (size_t)(par * R + me) * 16;
    const int n = tile * kTileM + row;
    for (int r = 0; r < R; ++r) {
        uint2* dst = reinterpret_cast<uint2*>(p.tp_stage[r]) + slot * p.tp_n_max + n;
#pragma unroll
        for (int j = 0; j < 16; ++j)
            if (j < ncol) ll_store(dst + (size_t)j * p.tp_n_max, vg[j], seq);
    }
    if (p.tp_defer) return;   // the consumer kernel waits and adds (tp_gather_prologue)
    const uint2* mine = reinterpret_cast<const uint2*>(p.tp_stage[me]);
    float own[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) { own[j] = vg[j]; vg[j] = 0.f; }
    for (int r = 0; r < R; ++r) {
        const uint2* src = mine + (size_t)(par * R + r) * 16 * p.tp_n_max + n;
#pragma unroll
        for (int j = 0; j < 16; ++j)
            if (j < ncol) vg[j] += (r == me) ? own[j] : ll_load(src + (size_t)j * p.tp_n_max, seq);
    }
}

// Per-token 1/rms for the deferred RMSNorm: all 128 worker threads cooperate, result in rstd[0..T).
// `red` is 4*32 floats of shared scratch; callers separate uses with their own barriers.
template <int LLIO = 0>
__device__ __forceinline__ void rms_stats_128(const GemmParams& p, int T, int e, float* red, float* rstd, uint32_t xtag = 0) {
    // All 128 threads on one token after the other, two tokens in flight (their loads are independent).  The
    // reduction tree of a token is the same for every T, so a token's 1/rms does not depend on how many sessions
    // share the step.  Measured alternatives: eight tokens in flight spills registers into the GEMM main loops
    // (T=1 chain 77 -> 87 us/layer); one warp per token needs 4+ dependent L2 round trips per token (77 -> 81.5).
    const int warp = e >> 5, lane = e & 31;
    for (int t0 = 0; t0 < T; t0 += 2) {
        float sa = 0.f, sb = 0.f;
        const bool two = t0 + 1 < T;
        const uint16_t* xa = p.x + (size_t)t0 * p.ldx;
        const uint16_t* xb = xa + (two ? p.ldx : 0);
        for (int c = e; c < p.K / 8; c += 128) {
            uint4 va, vb = make_uint4(0, 0, 0, 0);
            if (LLIO) {   // the caller has validated the whole vector (ll_validate_range)
                va = ll_load8_valid(p.ch.x + (size_t)t0 * p.ch.ldx + c * 4);
                if (two) vb = ll_load8_valid(p.ch.x + (size_t)(t0 + 1) * p.ch.ldx + c * 4);
            } else {
                va = ld_cg_v4(xa + c * 8);
                if (two) vb = ld_cg_v4(xb + c * 8);
            }
            const uint32_t wa[4] = {va.x, va.y, va.z, va.w};
            const uint32_t wb[4] = {vb.x, vb.y, vb.z, vb.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float a0 = h16_to_float((uint16_t)(wa[i] & 0xffff), p.is_bf16);
                const float a1 = h16_to_float((uint16_t)(wa[i] >> 16), p.is_bf16);
                sa += a0 * a0 + a1 * a1;
                const float b0 = h16_to_float((uint16_t)(wb[i] & 0xffff), p.is_bf16);
                const float b1 = h16_to_float((uint16_t)(wb[i] >> 16), p.is_bf16);
                sb += b0 * b0 + b1 * b1;
            }
        }
        sa = warp_sum(sa);
        sb = warp_sum(sb);
        if (lane == 0) {
            red[warp * 32 + t0] = sa;
            if (two) red[warp * 32 + t0 + 1] = sb;
        }
    }
    named_bar_sync(1, 128);
    if (e < T) {
        const float s = red[e] + red[32 + e] + red[64 + e] + red[96 + e];
        rstd[e] = rsqrtf(s / (float)p.K + p.norm_eps);
    }
    named_bar_sync(1, 128);
}

// Coalesced variant for tile-filling inputs: the warp's 32 rows x 16 token columns go through a 2 KiB
// shared-memory transpose so that every global access is a 16-byte vector along n (one 64-byte segment per
// token row per warp) instead of 2-byte accesses 2*ld bytes apart.  Same arithmetic as above.
// Requirements (checked by the caller): 16-bit output, N, ldo, ldr multiples of 8.
template <int BN, int DUAL>
__device__ __forceinline__ void epilogue_store_chunk_vec(const GemmParams& p, int n_tile, int t_tile, int lane, int lg,
                                                         int chunk, const float (&vg)[16], const float (&vu)[16],
                                                         float* warp_smem /* [16][32] */) {
    const int n_row = n_tile * kTileM + lg * 32 + lane;
    const float b = (p.bias && n_row < p.N) ? h16_to_float(p.bias[n_row], p.is_bf16) : 0.0f;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        float v = vg[j] + b;
        if (DUAL) {
            const float g = round16(v, p.is_bf16);
            const float u = round16(vu[j], p.is_bf16);
            const float s = round16(g / (1.0f + __expf(-g)), p.is_bf16);
            v = s * u;
        } else if (p.epilogue == B200_EPI_GELU_TANH) {
            v = gelu_tanh(round16(v, p.is_bf16));
        }
        warp_smem[j * 32 + lane] = v;
    }
    __syncwarp();
    const int q = lane & 3;                               // 8-wide n segment inside the warp's 32 rows
    const int n0 = n_tile * kTileM + lg * 32 + q * 8;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int j = (lane >> 2) + 8 * i;                 // token column inside the chunk
        const int t = t_tile * BN + chunk * 16 + j;
        if (t < p.T && n0 < p.N) {
            const float4 lo = *reinterpret_cast<const float4*>(warp_smem + j * 32 + q * 8);
            const float4 hi = *reinterpret_cast<const float4*>(warp_smem + j * 32 + q * 8 + 4);
            float f[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
            if (p.residual) {
                const uint4 r4 = ld_cg_v4(p.residual + (size_t)t * p.ldr + n0);
                const uint32_t rw[4] = {r4.x, r4.y, r4.z, r4.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    f[2 * k] = round16(f[2 * k], p.is_bf16) + h16_to_float((uint16_t)(rw[k] & 0xffff), p.is_bf16);
                    f[2 * k + 1] = round16(f[2 * k + 1], p.is_bf16) + h16_to_float((uint16_t)(rw[k] >> 16), p.is_bf16);
                }
            }
            uint32_t w[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)
                w[k] = (uint32_t)float_to_h16(f[2 * k], p.is_bf16) | ((uint32_t)float_to_h16(f[2 * k + 1], p.is_bf16) << 16);
            *reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.out) + (size_t)t * p.ldo + n0) =
                make_uint4(w[0], w[1], w[2], w[3]);
        }
    }
    __syncwarp();
}

// fp8 activations: per-token quantisation statistics in one pass over the row.
//   qmul[t]  = 448 / max|x*g|                 (multiplier applied before the e4m3 conversion)
//   scale[t] = max|x*g| / 448 * (RMSNorm ? 1/rms : 1)   (applied to the accumulator in the epilogue)
__device__ __forceinline__ void fp8_stats_128(const GemmParams& p, int T, int e, float* red, float* qmul, float* scale) {
    const int warp = e >> 5, lane = e & 31;
    for (int t = 0; t < T; ++t) {
        float s2 = 0.f, am = 0.f;
        const uint16_t* xr = p.x + (size_t)t * p.ldx;
        for (int c = e; c < p.K / 8; c += 128) {
            const uint4 v = ld_cg_v4(xr + c * 8);
            uint4 g4 = make_uint4(0, 0, 0, 0);
            if (p.norm_mode != B200_NORM_NONE) g4 = ld_nc_v4(p.norm_w + c * 8);
            const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
            const uint32_t gg[4] = {g4.x, g4.y, g4.z, g4.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float a = h16_to_float((uint16_t)(w4[i] & 0xffff), p.is_bf16);
                float b = h16_to_float((uint16_t)(w4[i] >> 16), p.is_bf16);
                s2 += a * a + b * b;
                if (p.norm_mode != B200_NORM_NONE) {
                    a *= h16_to_float((uint16_t)(gg[i] & 0xffff), p.is_bf16);
                    b *= h16_to_float((uint16_t)(gg[i] >> 16), p.is_bf16);
                }
                am = fmaxf(am, fmaxf(fabsf(a), fabsf(b)));
            }
        }
        s2 = warp_sum(s2);
        am = warp_max(am);
        if (lane == 0) { red[warp * 32 + t] = s2; red[128 + warp * 32 + t] = am; }
    }
    named_bar_sync(1, 128);
    if (e < T) {
        const float s = red[e] + red[32 + e] + red[64 + e] + red[96 + e];
        const float am = fmaxf(fmaxf(fmaxf(red[128 + e], red[160 + e]), fmaxf(red[192 + e], red[224 + e])), 1e-20f);
        const float rstd = (p.norm_mode == B200_NORM_RMS) ? rsqrtf(s / (float)p.K + p.norm_eps) : 1.0f;
        qmul[e] = 448.0f / am;
        scale[e] = am / 448.0f * rstd;
    }
    named_bar_sync(1, 128);
}

__device__ __forceinline__ uint32_t pack_e4m3x4(float a, float b, float c, float d) {
    const uint32_t lo = __nv_cvt_float2_to_fp8x2(make_float2(a, b), __NV_SATFINITE, __NV_E4M3);
    const uint32_t hi = __nv_cvt_float2_to_fp8x2(make_float2(c, d), __NV_SATFINITE, __NV_E4M3);
    return lo | (hi << 16);
}

// Rounds through the 16-bit storage type so fused epilogues see the same values
// the unfused torch sequence would have stored between ops.
__device__ __forceinline__ float round16(float v, int is_bf16) {
    return h16_to_float(float_to_h16(v, is_bf16), is_bf16);
}

// LLIO: the residual is polled from LL words (tag rtag) and the 16-bit result is ALSO published as LL words (tag otag);
// every lane of the warp must call (the lanes of a pair of rows combine their halves with a shuffle).
template <int BN, int DUAL, int LLIO = 0>
__device__ __forceinline__ void epilogue_store_chunk(const GemmParams& p, int n_tile, int t_tile,
                                                     int row, int chunk, const float (&vg)[16],
                                                     const float (&vu)[16], const float* tok_scale = nullptr,
                                                     uint32_t rtag = 0, uint32_t otag = 0) {
    const int n = n_tile * kTileM + row;
    if (!LLIO && n >= p.N) return;
    const bool live = n < p.N;
    const float b = (p.bias && live) ? h16_to_float(p.bias[n], p.is_bf16) : 0.0f;
    const float ws_g = (p.w_scale && live) ? p.w_scale[n] : 1.0f;
    const float ws_u = (DUAL && p.w_scale && live) ? p.w_scale[(size_t)p.n_tiles * kTileM + n] : 1.0f;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        const int t = t_tile * BN + chunk * 16 + j;
        if (t >= p.T) break;
        // deferred RMSNorm: the GEMM ran on x*g, the per-token 1/rms is applied here
        const float ts = tok_scale ? tok_scale[chunk * 16 + j] : 1.0f;
        float v = vg[j] * ts * ws_g + b;
        if (DUAL) {
            const float g = round16(v, p.is_bf16);
            const float u = round16(vu[j] * ts * ws_u, p.is_bf16);
            const float s = round16(g / (1.0f + __expf(-g)), p.is_bf16);
            v = s * u;
        } else if (p.epilogue == B200_EPI_GELU_TANH) {
            v = gelu_tanh(round16(v, p.is_bf16));
        }
        if (LLIO && p.ch.res) {
            if (live) v = round16(v, p.is_bf16) + h16_to_float(ll_load1(p.ch.res + (size_t)t * p.ch.ldr, n, rtag), p.is_bf16);
        } else if (p.residual) {
            v = round16(v, p.is_bf16) + h16_to_float(__ldcg(p.residual + (size_t)t * p.ldr + n), p.is_bf16);
        }
        if (p.out_fp32) {
            if (live) reinterpret_cast<float*>(p.out)[(size_t)t * p.ldo + n] = v;
        } else {
            const uint16_t hv = float_to_h16(v, p.is_bf16);
            if (live && p.out) reinterpret_cast<uint16_t*>(p.out)[(size_t)t * p.ldo + n] = hv;
            if (LLIO && p.ch.out) ll_store_pairwise(p.ch.out + (size_t)t * p.ch.ldo, n, hv, live, otag);
        }
    }
}

// The same epilogue for ONE element (output feature n, token t): used by the wide ticket-free fix-up, where a thread
// finishes elements of rows it did not accumulate.  Expression order as in epilogue_store_chunk (bit-identical).
// LLIO: as above; `valid` = this lane has an element at all (every lane of the warp calls, consecutive lanes hold
// consecutive n of one token).
template <int DUAL, int LLIO = 0>
__device__ __forceinline__ void epilogue_store_scalar(const GemmParams& p, int n, int t, float g_sum, float u_sum, float ts,
                                                      bool valid = true, uint32_t rtag = 0, uint32_t otag = 0) {
    const bool live = valid && n < p.N && t < p.T;
    if (!LLIO && !live) return;
    const float b = (p.bias && live) ? h16_to_float(p.bias[n], p.is_bf16) : 0.0f;
    const float ws_g = (p.w_scale && live) ? p.w_scale[n] : 1.0f;
    const float ws_u = (DUAL && p.w_scale && live) ? p.w_scale[(size_t)p.n_tiles * kTileM + n] : 1.0f;
    float v = g_sum * ts * ws_g + b;
    if (DUAL) {
        const float g = round16(v, p.is_bf16);
        const float u = round16(u_sum * ts * ws_u, p.is_bf16);
        const float s = round16(g / (1.0f + __expf(-g)), p.is_bf16);
        v = s * u;
    } else if (p.epilogue == B200_EPI_GELU_TANH) {
        v = gelu_tanh(round16(v, p.is_bf16));
    }
    if (LLIO && p.ch.res) {
        if (live) v = round16(v, p.is_bf16) + h16_to_float(ll_load1(p.ch.res + (size_t)t * p.ch.ldr, n, rtag), p.is_bf16);
    } else if (p.residual) {
        v = round16(v, p.is_bf16) + h16_to_float(__ldcg(p.residual + (size_t)t * p.ldr + n), p.is_bf16);
    }
    if (p.out_fp32) {
        if (live) reinterpret_cast<float*>(p.out)[(size_t)t * p.ldo + n] = v;
        return;
    }
    const uint16_t hv = float_to_h16(v, p.is_bf16);
    if (live && p.out) reinterpret_cast<uint16_t*>(p.out)[(size_t)t * p.ldo + n] = hv;
    if (LLIO && p.ch.out) ll_store_pairwise(p.ch.out + (size_t)(live ? t : 0) * p.ch.ldo, n, hv, live, otag);
}

}  // namespace b200
