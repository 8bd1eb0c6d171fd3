// One persistent launch for all the blocks of a decode step, with dataflow flags instead of launch boundaries.
//
// The per-launch timeline of a decode step (scripts/trace_decode.py, profiles/decode_timeline_*.txt) shows what a
// launch boundary costs even with programmatic dependent launch: the CTAs of kernel N+1 cannot become resident before
// CTAs of kernel N leave (registers and shared memory are full), and a fresh CTA needs 2-3 us (barrier set-up, tensor
// memory allocation, first dependent loads) before its first MMA - five times per decoder block, with the HBM idle.
// Here the grid (two CTAs per SM) stays resident and walks a list of ops:
//
//   warp 0      weight loader: streams the packed weights of GEMM after GEMM through one shared-memory ring and never
//               waits for activations - while the other warps finish op i the ring fills with op i+1's weights;
//   warp 1      MMA issuer (tcgen05, fp32 accumulators in tensor memory);
//   warps 2..5  workers: poll the tagged input words (b200_ll_chain), stage them, drain the accumulators, exchange
//               stream-K partials, publish the output as tagged words; RoPE + KV append + split-KV attention for
//               attention ops.
//
// There is NO grid barrier: an op starts on a CTA as soon as the words that CTA needs carry the op's input tag, so
// the only waits are true data dependencies (about one L2 round trip each).  Every CTA walks the ops in order and every
// dependency points to an earlier op (or, inside an op, to a lower item), so with all CTAs resident nothing can
// deadlock; every spin is bounded and traps with a message.
//
// Same arithmetic, summation order and rounding as the per-launch kernels: bit-identical results
// (tests/test_native_span_gpu.py).  Replaces the reference's eager per-op loop (src/llama_partition.py:233-297).
#include <stdlib.h>
#include <string.h>

#include "attn_core.cuh"
#include "gemm_core.cuh"
#include "trace.cuh"

namespace b200 {

struct ChainOp {
    int type;    // 0 = GEMM, 1 = attention
    int dual;    // GEMM: gate/up interleaved weight (SwiGLU epilogue)
    int parts;   // GEMM: CTAs per output tile (2..9), or 1 = whole tiles, tile u on CTA u, u + grid, ...
    int attn_d, attn_group;
    int pad;
    GemmParams g;
    AttnDecodeParams a;
};

constexpr int kCBN = 16;                       // token tile
constexpr int kCSA = 3;                        // 32 KiB weight stages
constexpr int kCStageA = 32768;
constexpr int kCGroup = 2;                     // k-blocks per activation stage
constexpr int kCSubB = kCBN * 128;             // 2 KiB
constexpr int kCStageB = kCSubB * kCGroup;
constexpr int kCSB = 3;
constexpr int kCNAcc = 4;
constexpr int kCTmemCols = 128;                // 16 columns x 2 matrices x 4 rotating accumulators
constexpr size_t kCSmemData = (size_t)kCSA * kCStageA + (size_t)kCSB * kCStageB;
constexpr size_t kCSmemBytes = kCSmemData + 1024 /*align*/ + 1024 /*barriers, statistics*/;
constexpr int kCMaxAttnGroup = 4;
static_assert(sizeof(AttnDecodeSmem<128, kCMaxAttnGroup>) <= (size_t)kCSB * kCStageB, "attention scratch aliases the activation ring");
constexpr int kCLLHalfWords = 512 * 2 * 16 * kTileM;   // words in one half of the exchange buffer (b200_gemm_ll_bytes)

// Unit j of a GEMM op on this CTA: k-blocks [kb0, kb1) of tile u.  Returns false when there is no such unit.
__device__ __forceinline__ bool chain_unit(const ChainOp& op, int cta, int G, int j, int& u, int& kb0, int& kb1) {
    const int kbpt = op.g.kbpt, nt = op.g.n_tiles;
    if (op.parts >= 2) {
        if (j > 0 || cta >= op.parts * nt) return false;
        u = cta / op.parts;
        const int part = cta - u * op.parts;
        kb0 = part * kbpt / op.parts;
        kb1 = (part + 1) * kbpt / op.parts;
        return true;
    }
    u = cta + j * G;
    kb0 = 0;
    kb1 = kbpt;
    return u < nt;
}

template <int D, int GROUP>
__device__ __forceinline__ void chain_attn(const AttnDecodeParams& A, uint8_t* smem, int e) {
    auto& sm = *reinterpret_cast<AttnDecodeSmem<D, GROUP>*>(smem);
    const int items = A.T * A.n_kv * A.splits;
    for (int item = blockIdx.x; item < items; item += gridDim.x) {   // increasing order: see the deadlock note above
        int b = item;
        const int s = b % A.splits; b /= A.splits;
        const int g = b % A.n_kv;
        const int t = b / A.n_kv;
        attn_decode_body<D, GROUP>(A, t, g, s, e, sm, [] { named_bar_sync(1, kEpiThreads); });
        named_bar_sync(1, kEpiThreads);   // the scratch is reused by the next item
    }
}

template <int DUAL>
__device__ __forceinline__ void chain_gemm_unit_epilogue(const GemmParams& p, int parts, int cta, int u, bool split,
                                                         const float (&vg)[16], const float (&vu)[16],
                                                         const float* tok_scale, int row, int lane, int e,
                                                         uint32_t rtag, uint32_t otag) {
    const int t_live = min(kCBN, p.T);
    if (!split) {
        epilogue_store_chunk<kCBN, DUAL, 1>(p, u, 0, row, 0, vg, vu, tok_scale, rtag, otag);
        return;
    }
    // wide exchange (see gemm.cu): publish the whole partial, then reduce this CTA's share of the tile's elements
    constexpr int APER = DUAL ? 2 : 1;
    constexpr int kStride = APER * 16 * kTileM;
    uint2* words = reinterpret_cast<uint2*>(p.ll_words) + (size_t)p.ch.half * kCLLHalfWords;
    uint2* mine = words + (size_t)cta * kStride + row;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        if (j < t_live) {
            ll_store(mine + j * kTileM, vg[j], (int)otag);
            if (DUAL) ll_store(mine + (16 + j) * kTileM, vu[j], (int)otag);
        }
    }
    const int P = parts, first = u * P;
    const int E = kTileM * t_live;
    const int per = (((E + P - 1) / P) + 1) & ~1;
    const int e_lo = min(E, (cta - first) * per), e_hi = min(E, e_lo + per);
    uint2* base = words + (size_t)first * kStride;
    for (int idx0 = e_lo + (e & ~31); idx0 < e_hi; idx0 += kEpiThreads) {
        const int idx = idx0 + lane;
        const bool valid = idx < e_hi;
        const int j = idx >> 7, r = idx & (kTileM - 1);
        float sg = 0.f, su = 0.f;
        if (valid) {
            uint2* w0 = base + j * kTileM + r;
            uint32_t vgw[9], tgw[9], vuw[9], tuw[9];
            long long t0 = 0;
            for (;;) {
                bool ok = true;
#pragma unroll
                for (int c = 0; c < 9; ++c) {
                    if (c < P) {
                        asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];\n"
                                     : "=r"(vgw[c]), "=r"(tgw[c]) : "l"(w0 + (size_t)c * kStride) : "memory");
                        if (DUAL)
                            asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];\n"
                                         : "=r"(vuw[c]), "=r"(tuw[c]) : "l"(w0 + (size_t)c * kStride + 16 * kTileM) : "memory");
                    }
                }
#pragma unroll
                for (int c = 0; c < 9; ++c)
                    if (c < P) ok = ok && tgw[c] == otag && (!DUAL || tuw[c] == otag);
                if (ok) break;
                ll_timeout(t0, "stream-K exchange of the chain kernel", otag);
            }
#pragma unroll
            for (int c = 0; c < 9; ++c) {
                if (c < P) {
                    sg += __uint_as_float(vgw[c]);
                    if (DUAL) su += __uint_as_float(vuw[c]);
                }
            }
        }
        epilogue_store_scalar<DUAL, 1>(p, u * kTileM + r, valid ? j : 0, sg, su, (tok_scale && valid) ? tok_scale[j] : 1.0f,
                                       valid, rtag, otag);
    }
}

#ifndef B200_CHAIN_BOUND_THREADS
#define B200_CHAIN_BOUND_THREADS 192   /* register budget of the chain kernel: 65536 / (2 x this) per thread */
#endif
// One instantiation per attention shape (head_dim, GQA group): every op of a plan has the same one, and a single
// inlined attention body keeps the kernel's live state in registers (six bodies behind calls put it in local memory).
template <int D, int GROUP>
__global__ void __launch_bounds__(B200_CHAIN_BOUND_THREADS, 2)
decode_chain_kernel(const ChainOp* __restrict__ ops, int n_ops, int debug) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + (size_t)kCSA * kCStageA;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kCSmemData);
    uint64_t* a_full = bars;
    uint64_t* a_empty = a_full + kCSA;
    uint64_t* b_full = a_empty + kCSA;
    uint64_t* b_empty = b_full + kCSB;
    uint64_t* acc_full = b_empty + kCSB;
    uint64_t* acc_empty = acc_full + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 1);
    float* stat_mean = reinterpret_cast<float*>(tmem_slot + 4);   // [32]
    float* stat_rstd = stat_mean + 32;                            // [32]
    float* stat_red = stat_rstd + 32;                             // [4][32]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cta = blockIdx.x, G = gridDim.x;

    if (threadIdx.x == 0) {
        for (int i = 0; i < kCSA; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 1); }
        for (int i = 0; i < kCSB; ++i) { mbar_init(&b_full[i], 4); mbar_init(&b_empty[i], 1); }
        mbar_init(acc_full, 1);
        mbar_init(acc_empty, 4);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<kCTmemCols>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ------------------------------------------------------------ weight loader (free-running over the op list)
        const uint64_t policy = l2_policy_evict_first();
        int s = 0;
        uint32_t ph = 0;
        for (int oi = 0; oi < n_ops; ++oi) {
            const ChainOp& op = ops[oi];
            if (op.type != 0) continue;
            const int block_a = op.dual ? 2 * kABytes : kABytes;
            const int kps = op.dual ? 1 : 2;
            int u, kb0, kb1;
            if (lane == 0) trace_event(3, 4, oi);
            for (int j = 0; chain_unit(op, cta, G, j, u, kb0, kb1); ++j) {
                const uint8_t* tile_base = op.g.w + (size_t)u * op.g.kbpt * block_a;
                for (int kb = kb0; kb < kb1;) {
                    const int n = min(kps, kb1 - kb);
                    mbar_wait(&a_empty[s], ph ^ 1);
                    const uint32_t bytes = (uint32_t)n * block_a;
                    if (elect_one()) {
                        if (debug & 4) mbar_arrive(&a_full[s]);       // experiment: no weight traffic at all
                        else {
                            mbar_arrive_expect_tx(&a_full[s], bytes);
                            bulk_g2s_hint(smem_a + (size_t)s * kCStageA, tile_base + (size_t)kb * block_a, bytes, &a_full[s], policy);
                        }
                    }
                    __syncwarp();
                    if (++s == kCSA) { s = 0; ph ^= 1; }
                    kb += n;
                }
            }
            if (lane == 0) trace_event(3, 5, oi);
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer (whole warp, elected lane issues)
        int sa = 0, sb = 0;
        uint32_t pha = 0, phb = 0, php = 0;
        const uint32_t a_base = smem_u32(smem_a), b_base = smem_u32(smem_b);
        for (int oi = 0; oi < n_ops; ++oi) {
            const ChainOp& op = ops[oi];
            if (op.type != 0) continue;
            const int aper = op.dual ? 2 : 1;
            const int block_a = aper * kABytes;
            const int kps = op.dual ? 1 : 2;
            const uint32_t idesc = umma_idesc(kTileM, kCBN, op.g.is_bf16, op.g.is_bf16, 0, 0);
            int u, kb0, kb1;
            for (int j = 0; chain_unit(op, cta, G, j, u, kb0, kb1); ++j) {
                const int nk = kb1 - kb0;
                if (lane == 0) trace_event(3, 6, oi);
                mbar_wait(acc_empty, php ^ 1);
                tc_fence_after();
                uint32_t acc = 0;
                int done = 0, bofs = 0;
                while (done < nk) {
                    const int na = min(kps, nk - done);
                    mbar_wait(&a_full[sa], pha);
                    for (int jj = 0; jj < na; ++jj) {
                        if (bofs == 0) mbar_wait(&b_full[sb], phb);
                        tc_fence_after();
                        const uint64_t bdesc = umma_desc_kmajor_sw128(b_base + sb * kCStageB + bofs * kCSubB, 1024);
                        if (!(debug & 2) && elect_one()) {
                            for (int a = 0; a < aper; ++a) {
                                const uint64_t adesc =
                                    umma_desc_kmajor_sw128(a_base + sa * kCStageA + jj * block_a + a * kABytes, 1024);
#pragma unroll
                                for (int k = 0; k < kBK / 16; ++k)
                                    umma_f16(tmem_base + (a * kCNAcc + k) * kCBN, adesc + (uint64_t)(k * 2),
                                             bdesc + (uint64_t)(k * 2), idesc, acc);
                            }
                        }
                        __syncwarp();
                        acc = 1;
                        ++bofs;
                        if (bofs == kCGroup || done + jj + 1 == nk) {
                            if (elect_one()) umma_commit(&b_empty[sb]);
                            __syncwarp();
                            bofs = 0;
                            if (++sb == kCSB) { sb = 0; phb ^= 1; }
                        }
                    }
                    if (elect_one()) umma_commit(&a_empty[sa]);
                    __syncwarp();
                    if (++sa == kCSA) { sa = 0; pha ^= 1; }
                    done += na;
                }
                if (elect_one()) umma_commit(acc_full);
                __syncwarp();
                if (lane == 0) trace_event(3, 7, oi);
                php ^= 1;
            }
        }
        __syncwarp();
    } else {
        // ------------------------------------------------------------ workers
        const int e = threadIdx.x - 64;
        const int lg = warp & 3;
        const int row = lg * 32 + lane;
        const int ch = e & 7;
        const int t = e >> 3;   // BN == 16: the token row this thread stages
        uint32_t part = 0, phb = 0;
        int sb = 0;
        int seq = 0;

        for (int oi = 0; oi < n_ops; ++oi) {
            const ChainOp& op = ops[oi];
            if (oi == 0) seq = ll_read_seq(op.type == 0 ? op.g.ch.seq : op.a.ch.seq);
            if (e == 0) trace_event(3, 0, oi);
            if (op.type == 1) {
                chain_attn<D, GROUP>(op.a, smem_b, e);
                if (e == 0) trace_event(3, 3, oi);
                continue;
            }
            const GemmParams& p = op.g;
            const uint32_t xtag = ll_tag(seq, p.ch.x_op), rtag = ll_tag(seq, p.ch.res_op), otag = ll_tag(seq, p.ch.out_op);
            const int T = min(kCBN, p.T);
            bool stats_ready = p.norm_mode == B200_NORM_NONE;
            int u, kb0, kb1;
            bool validated = false;
            for (int j = 0; chain_unit(op, cta, G, j, u, kb0, kb1); ++j) {
                if (!validated) {
                    // wait until the input words exist - the whole vector when a norm needs its statistics (or this CTA
                    // works on whole tiles), else the k-range of this CTA's unit; afterwards plain pipelined loads
                    int w0 = 0, w1 = p.K >> 1;
                    if (p.norm_mode == B200_NORM_NONE && op.parts >= 2) { w0 = kb0 * (kBK / 2); w1 = kb1 * (kBK / 2); }
                    ll_validate_range(p.ch.x, p.ch.ldx, T, w0, w1, xtag, e, kEpiThreads, 1);
                    validated = true;
                }
                if (p.norm_mode == B200_NORM_LAYER && !stats_ready) {
                    // LayerNorm needs its statistics before the staging (RMSNorm defers 1/rms to the epilogue)
                    const int pw = warp - 2;
                    for (int tt = pw; tt < T; tt += 4) {
                        float s1 = 0.f, s2 = 0.f;
                        for (int c = lane; c < p.K / 8; c += 32) {
                            const uint4 v = ll_load8_valid(p.ch.x + (size_t)tt * p.ch.ldx + c * 4);
                            const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const float a = h16_to_float((uint16_t)(w4[i] & 0xffff), p.is_bf16);
                                const float b = h16_to_float((uint16_t)(w4[i] >> 16), p.is_bf16);
                                s1 += a + b;
                                s2 += a * a + b * b;
                            }
                        }
                        s1 = warp_sum(s1);
                        s2 = warp_sum(s2);
                        if (lane == 0) {
                            const float inv = 1.0f / (float)p.K;
                            const float m = s1 * inv;
                            stat_mean[tt] = m;
                            stat_rstd[tt] = rsqrtf(fmaxf(s2 * inv - m * m, 0.f) + p.norm_eps);
                        }
                    }
                    named_bar_sync(1, kEpiThreads);
                    stats_ready = true;
                }
                // ---- stage activations, kCGroup k-blocks per stage, the next stage prefetched in registers
                {
                    uint4 cur[kCGroup], cgw[kCGroup], cgb[kCGroup], nxt[kCGroup], ngw[kCGroup], ngb[kCGroup];
                    auto load_stage = [&](int kb, uint4(&v)[kCGroup], uint4(&gw)[kCGroup], uint4(&gb)[kCGroup]) {
#pragma unroll
                        for (int g = 0; g < kCGroup; ++g) {
                            if (kb + g < kb1) {
                                const int kcol = (kb + g) * kBK + ch * 8;
                                v[g] = (t < p.T && !(debug & 1)) ? ll_load8_valid(p.ch.x + (size_t)t * p.ch.ldx + (kcol >> 1))
                                                                 : make_uint4(0, 0, 0, 0);
                                if (p.norm_mode != B200_NORM_NONE && !(debug & 1)) {
                                    gw[g] = ld_nc_v4(p.norm_w + kcol);
                                    gb[g] = p.norm_b ? ld_nc_v4(p.norm_b + kcol) : make_uint4(0, 0, 0, 0);
                                }
                            }
                        }
                    };
                    load_stage(kb0, cur, cgw, cgb);
                    for (int kb = kb0; kb < kb1; kb += kCGroup) {
                        if (kb + kCGroup < kb1) load_stage(kb + kCGroup, nxt, ngw, ngb);
                        mbar_wait(&b_empty[sb], phb ^ 1);
#pragma unroll
                        for (int g = 0; g < kCGroup; ++g) {
                            if (kb + g < kb1) {
                                uint4 val = cur[g];
                                if (p.norm_mode != B200_NORM_NONE && t < p.T) {
                                    const float mean = stat_mean[t], rstd = stat_rstd[t];
                                    uint32_t xin[4] = {val.x, val.y, val.z, val.w};
                                    const uint32_t win[4] = {cgw[g].x, cgw[g].y, cgw[g].z, cgw[g].w};
                                    const uint32_t bin[4] = {cgb[g].x, cgb[g].y, cgb[g].z, cgb[g].w};
#pragma unroll
                                    for (int w = 0; w < 4; ++w) {
                                        float x0 = h16_to_float((uint16_t)(xin[w] & 0xffff), p.is_bf16);
                                        float x1 = h16_to_float((uint16_t)(xin[w] >> 16), p.is_bf16);
                                        const float g0 = h16_to_float((uint16_t)(win[w] & 0xffff), p.is_bf16);
                                        const float g1 = h16_to_float((uint16_t)(win[w] >> 16), p.is_bf16);
                                        if (p.norm_mode == B200_NORM_RMS) {
                                            x0 = g0 * x0;   // 1/rms deferred to the epilogue
                                            x1 = g1 * x1;
                                        } else {
                                            const float b0 = h16_to_float((uint16_t)(bin[w] & 0xffff), p.is_bf16);
                                            const float b1 = h16_to_float((uint16_t)(bin[w] >> 16), p.is_bf16);
                                            x0 = (x0 - mean) * rstd * g0 + b0;
                                            x1 = (x1 - mean) * rstd * g1 + b1;
                                        }
                                        xin[w] = (uint32_t)float_to_h16(x0, p.is_bf16) | ((uint32_t)float_to_h16(x1, p.is_bf16) << 16);
                                    }
                                    val = make_uint4(xin[0], xin[1], xin[2], xin[3]);
                                }
                                const uint32_t base = smem_u32(smem_b + (size_t)sb * kCStageB + (size_t)g * kCSubB);
                                st_shared_v4(base + t * 128 + ((ch ^ (t & 7)) << 4), val);
                            }
                        }
                        fence_proxy_async_smem();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&b_full[sb]);
                        if (++sb == kCSB) { sb = 0; phb ^= 1; }
#pragma unroll
                        for (int g = 0; g < kCGroup; ++g) { cur[g] = nxt[g]; cgw[g] = ngw[g]; cgb[g] = ngb[g]; }
                    }
                }
                if (p.norm_mode == B200_NORM_RMS && !stats_ready) {
                    rms_stats_128<1>(p, T, e, stat_red, stat_rstd, xtag);   // while the tensor core works
                    stats_ready = true;
                }
                const float* tok_scale = (p.norm_mode == B200_NORM_RMS) ? stat_rstd : nullptr;
                // ---- drain the accumulators of this unit
                if (e == 0) trace_event(3, 1, oi);
                mbar_wait(acc_full, part & 1);
                tc_fence_after();
                if (e == 0) trace_event(3, 2, oi);
                float vg[16], vu[16];
#pragma unroll
                for (int jj = 0; jj < 16; ++jj) { vg[jj] = 0.f; vu[jj] = 0.f; }
                {
                    const uint32_t taddr = tmem_base + ((uint32_t)(lg * 32) << 16);
                    uint32_t rg[16], ru[16];
#pragma unroll
                    for (int ac = 0; ac < kCNAcc; ++ac) {
                        tmem_ld16(taddr + ac * kCBN, rg);
                        if (op.dual) tmem_ld16(taddr + (kCNAcc + ac) * kCBN, ru);
                        tmem_ld_wait();
#pragma unroll
                        for (int jj = 0; jj < 16; ++jj) {
                            vg[jj] += __uint_as_float(rg[jj]);
                            if (op.dual) vu[jj] += __uint_as_float(ru[jj]);
                        }
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(acc_empty);
                ++part;
                const bool split = op.parts >= 2;
                if (op.dual) chain_gemm_unit_epilogue<1>(p, op.parts, cta, u, split, vg, vu, tok_scale, row, lane, e, rtag, otag);
                else chain_gemm_unit_epilogue<0>(p, op.parts, cta, u, split, vg, vu, tok_scale, row, lane, e, rtag, otag);
            }
            if (e == 0) trace_event(3, 3, oi);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<kCTmemCols>(tmem_base);
}

}  // namespace b200

using namespace b200;

B200_DEFINE_TRACE_SETTER(b200_trace_set_chain)

extern "C" int64_t b200_decode_chain_plan_bytes(int32_t n_ops) { return (int64_t)n_ops * (int64_t)sizeof(ChainOp); }
extern "C" int32_t b200_decode_chain_max_attn_group(void) { return kCMaxAttnGroup; }

static int fill_gemm_op(const b200_gemm_args* a, ChainOp& op, int grid) {
    if (!a->w_packed || !a->chain.seq || !a->chain.x_ll || (!a->out_fp32 && !a->chain.out_ll) || (a->residual && !a->chain.res_ll))
        return fail(B200_EINVAL, "b200_decode_chain_build: GEMM op without weights / chain words");
    if (a->T <= 0 || a->T > kCBN || (a->K % kBK) != 0 || a->fp8 || a->rstd || a->n_tiles * kTileM < a->N)
        return fail(B200_ELIMIT, "b200_decode_chain_build: GEMM ops take 1..16 tokens, 16-bit weights (T=%d K=%d)", a->T, a->K);
    if (!a->ll_buf) return fail(B200_EINVAL, "b200_decode_chain_build: GEMM ops need ll_buf");
    if ((a->chain.ldx_ll % 4) != 0 || (reinterpret_cast<uintptr_t>(a->chain.x_ll) & 31) != 0)
        return fail(B200_EINVAL, "b200_decode_chain_build: chain.x_ll must be 32-byte aligned with ldx_ll %% 4 == 0");
    if (a->dual != (a->epilogue == B200_EPI_SWIGLU ? 1 : 0))
        return fail(B200_EINVAL, "b200_decode_chain_build: dual-packed weights <=> SwiGLU epilogue");
    memset(&op, 0, sizeof(op));
    op.type = 0;
    op.dual = a->dual ? 1 : 0;
    GemmParams& p = op.g;
    p.w = static_cast<const uint8_t*>(a->w_packed);
    p.x = static_cast<const uint16_t*>(a->x);
    p.out = a->out;
    p.residual = static_cast<const uint16_t*>(a->residual);
    p.bias = static_cast<const uint16_t*>(a->bias);
    p.norm_w = static_cast<const uint16_t*>(a->norm_w);
    p.norm_b = static_cast<const uint16_t*>(a->norm_b);
    p.T = a->T; p.N = a->N; p.K = a->K;
    p.ldx = a->ldx; p.ldo = a->ldo; p.ldr = a->ldr;
    p.n_tiles = a->n_tiles; p.t_tiles = 1; p.kbpt = a->K / kBK;
    p.Q = (long long)a->n_tiles * p.kbpt;
    p.epilogue = a->epilogue; p.norm_mode = a->norm_mode; p.norm_eps = a->norm_eps;
    p.is_bf16 = (a->dtype == B200_BF16) ? 1 : 0;
    p.out_fp32 = a->out_fp32;
    if (a->norm_mode != B200_NORM_NONE && !a->norm_w) return fail(B200_EINVAL, "b200_decode_chain_build: norm without norm_w");
    p.ll_words = reinterpret_cast<unsigned long long*>(static_cast<uint8_t*>(a->ll_buf) + 256);
    const b200_ll_chain& c = a->chain;
    p.ch.seq = c.seq;
    p.ch.x = static_cast<const uint2*>(c.x_ll);
    p.ch.res = a->residual ? static_cast<const uint2*>(c.res_ll) : nullptr;
    p.ch.out = a->out_fp32 ? nullptr : static_cast<uint2*>(c.out_ll);
    p.ch.ldx = c.ldx_ll; p.ch.ldr = c.ldr_ll; p.ch.ldo = c.ldo_ll;
    p.ch.x_op = c.x_op; p.ch.res_op = c.res_op; p.ch.out_op = c.out_op; p.ch.half = c.half & 1;
    int parts = grid / a->n_tiles;
    if (parts > 9) parts = 9;
    if (parts > p.kbpt) parts = p.kbpt;
    if (parts < 2 || parts * a->n_tiles > 512) parts = 1;
    op.parts = parts;
    return B200_OK;
}

// Validates the op list and writes the device-side plan (synchronous copy into plan_device, which must hold
// b200_decode_chain_plan_bytes(n_ops)).  GEMM ops: b200_gemm_args with .chain filled (T <= 16, 16-bit weights, ll_buf);
// attention ops: b200_attn_decode_args with .chain filled.  grid = the grid the plan will be launched with.
extern "C" int b200_decode_chain_build(const b200_chain_op* ops, int32_t n_ops, int32_t grid, void* plan_device) {
    if (!ops || n_ops <= 0 || !plan_device) return fail(B200_EINVAL, "b200_decode_chain_build: null argument");
    const int max_grid = 2 * sm_count_cached();
    if (grid <= 0 || grid > max_grid) return fail(B200_EINVAL, "b200_decode_chain_build: grid %d outside 1..%d", grid, max_grid);
    ChainOp* host = static_cast<ChainOp*>(malloc(sizeof(ChainOp) * (size_t)n_ops));
    if (!host) return fail(B200_ECUDA, "b200_decode_chain_build: out of host memory");
    const int32_t* seq = nullptr;
    for (int i = 0; i < n_ops; ++i) {
        int rc = B200_OK;
        if (ops[i].type == 0) {
            rc = fill_gemm_op(&ops[i].gemm, host[i], grid);
            if (rc == B200_OK) seq = seq ? seq : ops[i].gemm.chain.seq;
        } else {
            const b200_attn_decode_args* a = &ops[i].attn;
            const int group = a->n_kv_heads > 0 ? a->n_heads / a->n_kv_heads : 0;
            if (!a->chain.seq || !a->chain.x_ll || !a->chain.out_ll || !a->ll_scratch || !a->out || !a->kv_ptrs || !a->pos)
                rc = fail(B200_EINVAL, "b200_decode_chain_build: attention op without chain words / ll_scratch");
            else if (!(a->head_dim == 128 || a->head_dim == 64) || !(group == 1 || group == 2 || group == 4) || a->T > kCBN)
                rc = fail(B200_ELIMIT, "b200_decode_chain_build: attention ops take head_dim 64/128, GQA group 1/2/4 (got %d/%d)",
                          a->head_dim, group);
            else {
                memset(&host[i], 0, sizeof(ChainOp));
                host[i].type = 1;
                host[i].attn_d = a->head_dim;
                host[i].attn_group = group;
                AttnDecodeParams& p = host[i].a;
                p.qkv = static_cast<const uint16_t*>(a->qkv);
                p.out = static_cast<uint16_t*>(a->out);
                p.kv_ptrs = reinterpret_cast<const unsigned long long*>(a->kv_ptrs);
                p.pos = a->pos;
                p.inv_freq = a->inv_freq;
                p.scratch = a->scratch;
                p.counters = a->counters;
                p.ll = static_cast<unsigned long long*>(a->ll_scratch);
                p.T = a->T; p.n_heads = a->n_heads; p.n_kv = a->n_kv_heads; p.s_max = a->s_max;
                p.splits = a->splits > 0 ? a->splits : 1;
                p.ldqkv = a->ldqkv; p.ldo = a->ldo;
                p.scale = a->scale;
                p.is_bf16 = (a->dtype == B200_BF16) ? 1 : 0;
                p.ch.seq = a->chain.seq;
                p.ch.x = static_cast<const uint2*>(a->chain.x_ll);
                p.ch.out = static_cast<uint2*>(a->chain.out_ll);
                p.ch.ldx = a->chain.ldx_ll; p.ch.ldo = a->chain.ldo_ll;
                p.ch.x_op = a->chain.x_op; p.ch.out_op = a->chain.out_op;
                seq = seq ? seq : a->chain.seq;
            }
        }
        if (rc != B200_OK) { free(host); return rc; }
    }
    const cudaError_t e = cudaMemcpy(plan_device, host, sizeof(ChainOp) * (size_t)n_ops, cudaMemcpyHostToDevice);
    free(host);
    if (e != cudaSuccess) return fail(B200_ECUDA, "b200_decode_chain_build: copy failed: %s", cudaGetErrorString(e));
    return B200_OK;
}

template <int D, int GROUP>
static int launch_chain(const void* plan_device, int n_ops, int grid, bool pdl, int dbg, cudaStream_t stream) {
    static thread_local int configured_dev = -1;
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    if (configured_dev != dev) {
        // (cudaOccupancyMaxActiveBlocksPerMultiprocessor answers 1 for every tcgen05 kernel of this library, whatever its
        // registers and shared memory; the hardware runs two of these CTAs per SM - scripts/exp_residency.py stamps SM ids.)
        B200_CUDA_CHECK(cudaFuncSetAttribute(decode_chain_kernel<D, GROUP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)kCSmemBytes));
        B200_CUDA_CHECK(cudaFuncSetAttribute(decode_chain_kernel<D, GROUP>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                             (int)cudaSharedmemCarveoutMaxShared));
        configured_dev = dev;
    }
    B200_CUDA_CHECK(launch(decode_chain_kernel<D, GROUP>, dim3(grid), dim3(kThreads), kCSmemBytes, stream, pdl,
                           static_cast<const ChainOp*>(plan_device), n_ops, dbg));
    return B200_OK;
}

// head_dim / gqa_group: the attention shape of the plan's ops (all ops of a plan share it).
extern "C" int b200_decode_chain_launch(const void* plan_device, int32_t n_ops, int32_t grid, int32_t head_dim,
                                        int32_t gqa_group, int32_t pdl, void* stream_) {
    if (!plan_device || n_ops <= 0) return fail(B200_EINVAL, "b200_decode_chain_launch: null plan");
    const int max_grid = 2 * sm_count_cached();
    if (grid <= 0 || grid > max_grid) return fail(B200_EINVAL, "b200_decode_chain_launch: grid %d outside 1..%d", grid, max_grid);
    static const int dbg = [] { const char* e = getenv("B200_CHAIN_DEBUG"); return e ? atoi(e) : 0; }();
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
#define B200_CHAIN_CASE(D_, G_) \
    if (head_dim == D_ && gqa_group == G_) return launch_chain<D_, G_>(plan_device, n_ops, grid, pdl != 0, dbg, stream);
    B200_CHAIN_CASE(128, 1) B200_CHAIN_CASE(128, 2) B200_CHAIN_CASE(128, 4)
    B200_CHAIN_CASE(64, 1) B200_CHAIN_CASE(64, 2) B200_CHAIN_CASE(64, 4)
#undef B200_CHAIN_CASE
    return fail(B200_ELIMIT, "b200_decode_chain_launch: unsupported attention shape head_dim=%d group=%d", head_dim, gqa_group);
}
