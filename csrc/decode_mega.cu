// Persistent multi-op decode kernel: one launch runs every GEMM and attention op of a
// stage's decode step (all blocks, plus the LM head on the last stage).
//
// Why: at batch-1 decode each weight-streaming GEMM lasts 5-40 us, and a separate launch per op
// costs 8-12 us of launch gap, TMEM/barrier set-up, ring fill and drain — more than half of the
// HBM roofline (profiles/decode_step_kernel_shares_v1.csv).  Here the grid stays resident
// (one CTA per SM, cooperative launch) and walks an op list:
//
//   warp 0      weight loader: streams the packed weights of op after op through ONE shared-memory
//               ring and never waits for activations — while the other warps finish op i (epilogue,
//               grid barrier) the ring already holds the first ~28 MB of op i+1's weights, so the
//               HBM stream does not stop at op boundaries;
//   warp 1      MMA issuer (tcgen05, fp32 accumulators in TMEM, 4 rotating accumulators);
//   warps 2..5  workers: grid barrier -> norm statistics -> activation staging -> epilogue /
//               stream-K fix-up for GEMM ops; RoPE + KV append + split-KV attention for ATTN ops.
//
// Ops form a chain (op i consumes op i-1's output), so a grid-wide barrier (one global counter,
// release/acquire) separates them; activations are read through L2 (ld.global.cg) because other
// CTAs rewrite them inside the same launch.  T <= 16 tokens (one per session).
#include <string.h>

#include "attn_core.cuh"
#include "gemm_core.cuh"

namespace b200 {

struct MegaOp {
    int type;  // 0 = GEMM, 1 = ATTN
    int dual;
    int attn_d, attn_group;
    GemmParams g;
    AttnDecodeParams a;
};

constexpr int kMegaBN = 16;
constexpr int kMegaSA = 6;                       // 32 KiB weight stages
constexpr int kMegaStageA = 32768;
constexpr int kMegaGroup = 4;                    // k-blocks per activation stage
constexpr int kMegaSubB = kMegaBN * 128;         // 2 KiB
constexpr int kMegaStageB = kMegaSubB * kMegaGroup;
constexpr int kMegaSB = 3;
constexpr int kMegaNAcc = 4;
constexpr int kMegaTmemCols = 128;               // 16 cols x 2 matrices x 4 accumulators
constexpr size_t kMegaSmemData = (size_t)kMegaSA * kMegaStageA + (size_t)kMegaSB * kMegaStageB;
constexpr size_t kMegaSmemBytes = kMegaSmemData + 1024 + 2048;
static_assert(sizeof(AttnDecodeSmem<128, 8>) <= (size_t)kMegaSB * kMegaStageB, "attention scratch must fit the B ring");

__device__ __forceinline__ int ld_acquire_gpu(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_gpu_add(int* p, int v) {
    asm volatile("red.release.gpu.global.add.s32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t));
    return t;
}

__device__ __forceinline__ long long mega_range_begin(long long Q, int c) {
    return (long long)c * Q / (long long)gridDim.x;
}

template <int D, int GROUP>
__device__ __forceinline__ void mega_attn(const AttnDecodeParams& A, uint8_t* smem, int e) {
    auto& sm = *reinterpret_cast<AttnDecodeSmem<D, GROUP>*>(smem);
    const int items = A.T * A.n_kv * A.splits;
    for (int item = blockIdx.x; item < items; item += gridDim.x) {
        int b = item;
        const int s = b % A.splits; b /= A.splits;
        const int g = b % A.n_kv;
        const int t = b / A.n_kv;
        attn_decode_body<D, GROUP>(A, t, g, s, e, sm, [] { named_bar_sync(1, kEpiThreads); });
        named_bar_sync(1, kEpiThreads);   // smem is reused by the next item
    }
}

__global__ void __launch_bounds__(kThreads, 1)
decode_mega_kernel(const MegaOp* __restrict__ ops, int n_ops, int* __restrict__ grid_bar,
                   unsigned long long* __restrict__ trace) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + (size_t)kMegaSA * kMegaStageA;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kMegaSmemData);
    uint64_t* a_full = bars;
    uint64_t* a_empty = a_full + kMegaSA;
    uint64_t* b_full = a_empty + kMegaSA;
    uint64_t* b_empty = b_full + kMegaSB;
    uint64_t* acc_full = b_empty + kMegaSB;
    uint64_t* acc_empty = acc_full + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 1);
    volatile int* fin_flag = reinterpret_cast<volatile int*>(tmem_slot + 1);  // [2]
    float* stat_mean = reinterpret_cast<float*>(tmem_slot + 4);               // [16]
    float* stat_rstd = stat_mean + 32;
    float* stat_red = stat_rstd + 32;   // [4][32]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cta = blockIdx.x, G = gridDim.x;

    if (threadIdx.x == 0) {
        for (int i = 0; i < kMegaSA; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 1); }
        for (int i = 0; i < kMegaSB; ++i) { mbar_init(&b_full[i], 4); mbar_init(&b_empty[i], 1); }
        mbar_init(acc_full, 1);
        mbar_init(acc_empty, 4);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<kMegaTmemCols>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();

    if (warp == 0) {
        // ------------------------------------------------------------ weight loader (free-running)
        // whole warp, one elected lane issues (see gemm.cu)
        {
            const uint64_t policy = l2_policy_evict_first();
            int s = 0;
            uint32_t ph = 0;
            for (int oi = 0; oi < n_ops; ++oi) {
                if (ops[oi].type != 0) continue;
                const GemmParams& P = ops[oi].g;
                const int kbpt = P.kbpt;
                const int block_a = ops[oi].dual ? 32768 : 16384;
                const int kps = ops[oi].dual ? 1 : 2;
                const long long Q = P.Q;
                const long long qb = mega_range_begin(Q, cta), qe = mega_range_begin(Q, cta + 1);
                int u = (int)(qb / kbpt);
                int kb = (int)(qb - (long long)u * kbpt);
                long long left = qe - qb;
                const uint8_t* tile_base = P.w + (size_t)u * kbpt * block_a;
                while (left > 0) {
                    int n = kbpt - kb;
                    if ((long long)n > left) n = (int)left;
                    if (n > kps) n = kps;
                    mbar_wait(&a_empty[s], ph ^ 1);
                    const uint32_t bytes = (uint32_t)n * block_a;
                    if (elect_one()) {
                        mbar_arrive_expect_tx(&a_full[s], bytes);
                        bulk_g2s_hint(smem_a + (size_t)s * kMegaStageA, tile_base + (size_t)kb * block_a, bytes,
                                      &a_full[s], policy);
                    }
                    __syncwarp();
                    if (++s == kMegaSA) { s = 0; ph ^= 1; }
                    left -= n;
                    kb += n;
                    if (kb == kbpt) { kb = 0; tile_base += (size_t)kbpt * block_a; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer (whole warp, elected lane issues)
        {
            int sa = 0, sb = 0;
            uint32_t pha = 0, phb = 0, php = 0;
            const uint32_t a_base = smem_u32(smem_a), b_base = smem_u32(smem_b);
            for (int oi = 0; oi < n_ops; ++oi) {
                if (ops[oi].type != 0) continue;
                const GemmParams& P = ops[oi].g;
                const int kbpt = P.kbpt;
                const int aper = ops[oi].dual ? 2 : 1;
                const int block_a = aper * kABytes;
                const int kps = ops[oi].dual ? 1 : 2;
                const uint32_t idesc = umma_idesc(kTileM, kMegaBN, P.is_bf16, P.is_bf16, 0, 0);
                const long long qb = mega_range_begin(P.Q, cta), qe = mega_range_begin(P.Q, cta + 1);
                int kb0 = (int)(qb % kbpt);
                long long left = qe - qb;
                while (left > 0) {
                    int nk = kbpt - kb0;
                    if ((long long)nk > left) nk = (int)left;
                    mbar_wait(acc_empty, php ^ 1);
                    tc_fence_after();
                    uint32_t acc = 0;
                    int done = 0, bofs = 0;
                    while (done < nk) {
                        int na = nk - done;
                        if (na > kps) na = kps;
                        mbar_wait(&a_full[sa], pha);
                        for (int j = 0; j < na; ++j) {
                            if (bofs == 0) mbar_wait(&b_full[sb], phb);
                            tc_fence_after();
                            const uint64_t bdesc = umma_desc_kmajor_sw128(b_base + sb * kMegaStageB + bofs * kMegaSubB, 1024);
                            if (elect_one())
                            for (int a = 0; a < aper; ++a) {
                                const uint64_t adesc =
                                    umma_desc_kmajor_sw128(a_base + sa * kMegaStageA + j * block_a + a * kABytes, 1024);
#pragma unroll
                                for (int k = 0; k < kBK / 16; ++k) {
                                    umma_f16(tmem_base + (a * kMegaNAcc + k) * kMegaBN, adesc + (uint64_t)(k * 2),
                                             bdesc + (uint64_t)(k * 2), idesc, acc);
                                }
                            }
                            __syncwarp();
                            acc = 1;
                            ++bofs;
                            if (bofs == kMegaGroup || done + j + 1 == nk) {
                                if (elect_one()) umma_commit(&b_empty[sb]);
                                __syncwarp();
                                bofs = 0;
                                if (++sb == kMegaSB) { sb = 0; phb ^= 1; }
                            }
                        }
                        if (elect_one()) umma_commit(&a_empty[sa]);
                        __syncwarp();
                        if (++sa == kMegaSA) { sa = 0; pha ^= 1; }
                        done += na;
                    }
                    if (elect_one()) umma_commit(acc_full);
                    __syncwarp();
                    php ^= 1;
                    left -= nk;
                    kb0 = 0;
                }
            }
        }
        __syncwarp();
    } else {
        // ------------------------------------------------------------ workers
        const int e = threadIdx.x - 64;
        const int lg = warp & 3;
        const int row = lg * 32 + lane;
        const int ch = e & 7;
        const int r0 = e >> 3;   // 0..15: token row staged by this thread
        uint32_t part = 0, phb = 0;
        int sb = 0;

        for (int oi = 0; oi < n_ops; ++oi) {
            if (oi > 0) {
                // grid barrier: every CTA has finished op oi-1
                if (e == 0) {
                    const int target = G * oi;
                    long long t0 = clock64();
                    while (ld_acquire_gpu(grid_bar) < target) {
                        __nanosleep(40);
                        if (clock64() - t0 > 8000000000LL) {
                            printf("b200: decode_mega grid barrier timed out (op %d, block %d)\n", oi, cta);
                            __trap();
                        }
                    }
                }
                named_bar_sync(1, kEpiThreads);
                __threadfence();
            }
            if (trace && e == 0 && (cta == 0 || cta == G - 1)) trace[((cta ? 1 : 0) * n_ops + oi) * 2] = global_timer_ns();
            if (ops[oi].type == 1) {
                const AttnDecodeParams A = ops[oi].a;
                const int D = ops[oi].attn_d, GR = ops[oi].attn_group;
                if (D == 128 && GR == 1) mega_attn<128, 1>(A, smem_b, e);
                else if (D == 128 && GR == 4) mega_attn<128, 4>(A, smem_b, e);
                else if (D == 128 && GR == 8) mega_attn<128, 8>(A, smem_b, e);
                else if (D == 128 && GR == 2) mega_attn<128, 2>(A, smem_b, e);
                else if (D == 64 && GR == 1) mega_attn<64, 1>(A, smem_b, e);
            } else {
                const GemmParams p = ops[oi].g;
                const int dual = ops[oi].dual;
                const int aper = dual ? 2 : 1;
                if (p.norm_mode == B200_NORM_LAYER) {
                    const int pw = warp - 2;
                    for (int t = pw; t < kMegaBN; t += 4) {
                        float s1 = 0.f, s2 = 0.f;
                        if (t < p.T) {
                            const uint16_t* xr = p.x + (size_t)t * p.ldx;
                            for (int c = lane; c < p.K / 8; c += 32) {
                                const uint4 v = ld_cg_v4(xr + c * 8);
                                const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                                for (int i = 0; i < 4; ++i) {
                                    const float a = h16_to_float((uint16_t)(w4[i] & 0xffff), p.is_bf16);
                                    const float b = h16_to_float((uint16_t)(w4[i] >> 16), p.is_bf16);
                                    s1 += a + b;
                                    s2 += a * a + b * b;
                                }
                            }
                        }
                        s1 = warp_sum(s1);
                        s2 = warp_sum(s2);
                        if (lane == 0) {
                            const float inv = 1.0f / (float)p.K;
                            if (p.norm_mode == B200_NORM_RMS) {
                                stat_mean[t] = 0.f;
                                stat_rstd[t] = rsqrtf(s2 * inv + p.norm_eps);
                            } else {
                                const float m = s1 * inv;
                                stat_mean[t] = m;
                                stat_rstd[t] = rsqrtf(fmaxf(s2 * inv - m * m, 0.f) + p.norm_eps);
                            }
                        }
                    }
                    named_bar_sync(1, kEpiThreads);
                }
                const long long q_begin = mega_range_begin(p.Q, cta), q_end = mega_range_begin(p.Q, cta + 1);
                long long q = q_begin;
                int first_part = 1;
                bool rms_ready = false;
                while (q < q_end) {
                    const int u = (int)(q / p.kbpt);
                    const int kb0 = (int)(q - (long long)u * p.kbpt);
                    const long long rem = q_end - q;
                    const int kb1 = (rem < (long long)(p.kbpt - kb0)) ? kb0 + (int)rem : p.kbpt;
                    const int n_tile = u;
                    const int t = r0;   // BN == 16: one token row per thread
                    // ---- stage activations, kMegaGroup k-blocks per stage, next stage prefetched in registers
                    {
                        uint4 cur[kMegaGroup], cgw[kMegaGroup], cgb[kMegaGroup];
                        uint4 nxt[kMegaGroup], ngw[kMegaGroup], ngb[kMegaGroup];
                        auto load_stage = [&](int kb, uint4(&v)[kMegaGroup], uint4(&gw)[kMegaGroup], uint4(&gb)[kMegaGroup]) {
#pragma unroll
                            for (int g = 0; g < kMegaGroup; ++g) {
                                if (kb + g < kb1) {
                                    const int kcol = (kb + g) * kBK + ch * 8;
                                    v[g] = (t < p.T) ? ld_cg_v4(p.x + (size_t)t * p.ldx + kcol) : make_uint4(0, 0, 0, 0);
                                    if (p.norm_mode != B200_NORM_NONE) {
                                        gw[g] = ld_nc_v4(p.norm_w + kcol);
                                        gb[g] = p.norm_b ? ld_nc_v4(p.norm_b + kcol) : make_uint4(0, 0, 0, 0);
                                    }
                                }
                            }
                        };
                        load_stage(kb0, cur, cgw, cgb);
                        for (int kb = kb0; kb < kb1; kb += kMegaGroup) {
                            if (kb + kMegaGroup < kb1) load_stage(kb + kMegaGroup, nxt, ngw, ngb);
                            mbar_wait(&b_empty[sb], phb ^ 1);
#pragma unroll
                            for (int g = 0; g < kMegaGroup; ++g) {
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
                                            xin[w] = (uint32_t)float_to_h16(x0, p.is_bf16) |
                                                     ((uint32_t)float_to_h16(x1, p.is_bf16) << 16);
                                        }
                                        val = make_uint4(xin[0], xin[1], xin[2], xin[3]);
                                    }
                                    const uint32_t base = smem_u32(smem_b + (size_t)sb * kMegaStageB + (size_t)g * kMegaSubB);
                                    st_shared_v4(base + t * 128 + ((ch ^ (t & 7)) << 4), val);
                                }
                            }
                            fence_proxy_async_smem();
                            __syncwarp();
                            if (lane == 0) mbar_arrive(&b_full[sb]);
                            if (++sb == kMegaSB) { sb = 0; phb ^= 1; }
#pragma unroll
                            for (int g = 0; g < kMegaGroup; ++g) { cur[g] = nxt[g]; cgw[g] = ngw[g]; cgb[g] = ngb[g]; }
                        }
                    }
                    if (p.norm_mode == B200_NORM_RMS && !rms_ready) {
                        rms_stats_128(p, min(kMegaBN, p.T), e, stat_red, stat_rstd);
                        rms_ready = true;
                    }
                    const float* tok_scale = (p.norm_mode == B200_NORM_RMS) ? stat_rstd : nullptr;
                    // ---- drain the accumulators of this part
                    const bool full_tile = (kb0 == 0) && (kb1 == p.kbpt);
                    const int slot = first_part ? 0 : 1;
                    first_part = 0;
                    float* my_partial = p.scratch + ((size_t)cta * 2 + slot) * (size_t)(aper * kMegaBN * kTileM);
                    mbar_wait(acc_full, part & 1);
                    tc_fence_after();
                    const int ncol = min(kMegaBN, p.T);
                    float vg[16], vu[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) { vg[j] = 0.f; vu[j] = 0.f; }
                    {
                        const uint32_t taddr = tmem_base + ((uint32_t)(lg * 32) << 16);
                        uint32_t rg[16], ru[16];
#pragma unroll
                        for (int ac = 0; ac < kMegaNAcc; ++ac) {
                            tmem_ld16(taddr + ac * kMegaBN, rg);
                            if (dual) tmem_ld16(taddr + (kMegaNAcc + ac) * kMegaBN, ru);
                            tmem_ld_wait();
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                vg[j] += __uint_as_float(rg[j]);
                                if (dual) vu[j] += __uint_as_float(ru[j]);
                            }
                        }
                    }
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(acc_empty);
                    if (full_tile) {
                        if (dual) epilogue_store_chunk<kMegaBN, 1>(p, n_tile, 0, row, 0, vg, vu, tok_scale);
                        else epilogue_store_chunk<kMegaBN, 0>(p, n_tile, 0, row, 0, vg, vu, tok_scale);
                    } else {
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            if (j < ncol) {
                                my_partial[(size_t)j * kTileM + row] = vg[j];
                                if (dual) my_partial[(size_t)(kMegaBN + j) * kTileM + row] = vu[j];
                            }
                        }
                        __threadfence();
                        named_bar_sync(1, kEpiThreads);
                        if (e == 0) {
                            const int add = kb1 - kb0;
                            const int old = atomicAdd(&p.counters[u], add);
                            fin_flag[part & 1] = (old + add == p.kbpt) ? 1 : 0;
                        }
                        named_bar_sync(1, kEpiThreads);
                        if (fin_flag[part & 1]) {
                            __threadfence();
                            const long long t_lo = (long long)u * p.kbpt, t_hi = t_lo + p.kbpt;
                            int c_first = (int)((t_lo * (long long)G) / p.Q);
                            while (c_first + 1 < G && mega_range_begin(p.Q, c_first + 1) <= t_lo) ++c_first;
                            while (c_first > 0 && mega_range_begin(p.Q, c_first) > t_lo) --c_first;
#pragma unroll
                            for (int j = 0; j < 16; ++j) { vg[j] = 0.f; vu[j] = 0.f; }
#pragma unroll 4
                            for (int cc = c_first; cc < G; ++cc) {
                                const long long rb = mega_range_begin(p.Q, cc), re = mega_range_begin(p.Q, cc + 1);
                                if (rb >= t_hi) break;
                                if (re <= rb || re <= t_lo) continue;
                                const int cslot = (rb >= t_lo) ? 0 : 1;
                                const float* src = p.scratch + ((size_t)cc * 2 + cslot) * (size_t)(aper * kMegaBN * kTileM);
#pragma unroll
                                for (int j = 0; j < 16; ++j) {
                                    if (j < ncol) {
                                        vg[j] += __ldcg(src + (size_t)j * kTileM + row);
                                        if (dual) vu[j] += __ldcg(src + (size_t)(kMegaBN + j) * kTileM + row);
                                    }
                                }
                            }
                            if (dual) epilogue_store_chunk<kMegaBN, 1>(p, n_tile, 0, row, 0, vg, vu, tok_scale);
                            else epilogue_store_chunk<kMegaBN, 0>(p, n_tile, 0, row, 0, vg, vu, tok_scale);
                            if (e == 0) p.counters[u] = 0;
                        }
                    }
                    q += kb1 - kb0;
                    ++part;
                }
            }
            // ---- this CTA is done with op oi
            if (trace && e == 0 && (cta == 0 || cta == G - 1)) trace[((cta ? 1 : 0) * n_ops + oi) * 2 + 1] = global_timer_ns();
            __threadfence();
            named_bar_sync(1, kEpiThreads);
            if (e == 0) red_release_gpu_add(grid_bar, 1);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<kMegaTmemCols>(tmem_base);
}

}  // namespace b200

using namespace b200;

extern "C" int64_t b200_decode_mega_plan_bytes(int32_t n_ops) { return (int64_t)n_ops * (int64_t)sizeof(MegaOp); }

// Validates and compiles the public op list into the device-side plan (synchronous copy).
extern "C" int b200_decode_mega_build(const b200_mega_op* ops, int32_t n_ops, void* plan_device) {
    if (!ops || !plan_device || n_ops <= 0) return fail(B200_EINVAL, "b200_decode_mega_build: bad arguments");
    MegaOp* host = new MegaOp[n_ops];
    int rc = B200_OK;
    for (int i = 0; i < n_ops && rc == B200_OK; ++i) {
        MegaOp& m = host[i];
        memset(&m, 0, sizeof(m));
        if (ops[i].type == 0) {
            const b200_gemm_args& a = ops[i].gemm;
            if (!a.w_packed || !a.x || !a.out || !a.scratch || !a.counters)
                rc = fail(B200_EINVAL, "mega op %d: null GEMM pointer", i);
            else if (a.T <= 0 || a.T > kMegaBN)
                rc = fail(B200_ELIMIT, "mega op %d: T=%d exceeds the decode kernel's %d tokens", i, a.T, kMegaBN);
            else if (a.K <= 0 || a.K % kBK != 0 || a.n_tiles * kTileM < a.N)
                rc = fail(B200_EINVAL, "mega op %d: bad GEMM sizes N=%d K=%d n_tiles=%d", i, a.N, a.K, a.n_tiles);
            else if ((a.dual != 0) != (a.epilogue == B200_EPI_SWIGLU))
                rc = fail(B200_EINVAL, "mega op %d: dual packing and SwiGLU epilogue must come together", i);
            else if (a.norm_mode != B200_NORM_NONE && !a.norm_w)
                rc = fail(B200_EINVAL, "mega op %d: norm without weights", i);
            else if (a.fp8)
                rc = fail(B200_EINVAL, "mega op %d: fp8 weights are served by the per-op path", i);
            else if ((a.ldx % 8) != 0 || (reinterpret_cast<uintptr_t>(a.x) & 15) != 0)
                rc = fail(B200_EINVAL, "mega op %d: misaligned activations", i);
            if (rc != B200_OK) break;
            m.type = 0;
            m.dual = a.dual ? 1 : 0;
            GemmParams& p = m.g;
            p.w = static_cast<const uint8_t*>(a.w_packed);
            p.x = static_cast<const uint16_t*>(a.x);
            p.out = a.out;
            p.residual = static_cast<const uint16_t*>(a.residual);
            p.bias = static_cast<const uint16_t*>(a.bias);
            p.norm_w = static_cast<const uint16_t*>(a.norm_w);
            p.norm_b = static_cast<const uint16_t*>(a.norm_b);
            p.scratch = a.scratch;
            p.counters = a.counters;
            p.ll_words = nullptr; p.ll_parts = 0;   // the persistent kernel keeps the ticket fix-up
            p.rstd_in = nullptr;
            p.pf_ptr[0] = p.pf_ptr[1] = nullptr; p.pf_bytes[0] = p.pf_bytes[1] = 0;
            p.T = a.T; p.N = a.N; p.K = a.K;
            p.ldx = a.ldx; p.ldo = a.ldo; p.ldr = a.ldr;
            p.n_tiles = a.n_tiles; p.t_tiles = 1; p.kbpt = a.K / kBK;
            p.Q = (long long)a.n_tiles * p.kbpt;
            p.epilogue = a.epilogue; p.norm_mode = a.norm_mode; p.norm_eps = a.norm_eps;
            p.is_bf16 = (a.dtype == B200_BF16) ? 1 : 0;
            p.out_fp32 = a.out_fp32;
            p.debug = 0;
        } else if (ops[i].type == 1) {
            const b200_attn_decode_args& a = ops[i].attn;
            if (!a.qkv || !a.out || !a.kv_ptrs || !a.pos || !a.scratch || !a.counters)
                rc = fail(B200_EINVAL, "mega op %d: null attention pointer", i);
            else if (a.n_kv_heads <= 0 || a.n_heads % a.n_kv_heads != 0)
                rc = fail(B200_EINVAL, "mega op %d: bad head counts", i);
            if (rc != B200_OK) break;
            const int group = a.n_heads / a.n_kv_heads;
            const bool ok = (a.head_dim == 128 && (group == 1 || group == 2 || group == 4 || group == 8)) ||
                            (a.head_dim == 64 && group == 1);
            if (!ok) { rc = fail(B200_EINVAL, "mega op %d: unsupported head_dim=%d group=%d", i, a.head_dim, group); break; }
            m.type = 1;
            m.attn_d = a.head_dim;
            m.attn_group = group;
            AttnDecodeParams& p = m.a;
            p.qkv = static_cast<const uint16_t*>(a.qkv);
            p.out = static_cast<uint16_t*>(a.out);
            p.kv_ptrs = reinterpret_cast<const unsigned long long*>(a.kv_ptrs);
            p.pos = a.pos;
            p.inv_freq = a.inv_freq;
            p.scratch = a.scratch;
            p.counters = a.counters;
            p.ll = nullptr;   // the persistent kernel keeps the ticket merge
            p.T = a.T; p.n_heads = a.n_heads; p.n_kv = a.n_kv_heads; p.s_max = a.s_max;
            p.splits = a.splits > 0 ? a.splits : 1;
            p.ldqkv = a.ldqkv; p.ldo = a.ldo;
            p.scale = a.scale;
            p.is_bf16 = (a.dtype == B200_BF16) ? 1 : 0;
        } else {
            rc = fail(B200_EINVAL, "mega op %d: unknown type %d", i, ops[i].type);
        }
    }
    if (rc == B200_OK) {
        cudaError_t e = cudaMemcpy(plan_device, host, (size_t)n_ops * sizeof(MegaOp), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) rc = fail(B200_ECUDA, "b200_decode_mega_build: copy failed: %s", cudaGetErrorString(e));
    }
    delete[] host;
    return rc;
}

static unsigned long long* g_mega_trace = nullptr;
// Experiment hook: when set, CTA 0 and the last CTA record (start, end) global-timer stamps per op
// into trace[2][n_ops][2] (device memory owned by the caller).
extern "C" void b200_decode_mega_set_trace(void* trace) { g_mega_trace = static_cast<unsigned long long*>(trace); }

extern "C" int b200_decode_mega_launch(const void* plan_device, int32_t n_ops, int32_t* grid_barrier, int32_t grid,
                                       int32_t cooperative, void* stream_) {
    if (!plan_device || !grid_barrier || n_ops <= 0) return fail(B200_EINVAL, "b200_decode_mega_launch: bad arguments");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    static thread_local int configured_dev = -1;
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    if (configured_dev != dev) {
        B200_CUDA_CHECK(cudaFuncSetAttribute(decode_mega_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)kMegaSmemBytes));
        configured_dev = dev;
    }
    int g = grid > 0 ? grid : sm_count_cached();
    if (g > sm_count_cached()) g = sm_count_cached();   // every CTA must be resident: the ops are separated by grid barriers
    B200_CUDA_CHECK(cudaMemsetAsync(grid_barrier, 0, sizeof(int), stream));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(g);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = kMegaSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    if (cooperative) {
        attr[0].id = cudaLaunchAttributeCooperative;
        attr[0].val.cooperative = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
    }
    const MegaOp* ops = static_cast<const MegaOp*>(plan_device);
    B200_CUDA_CHECK(cudaLaunchKernelEx(&cfg, decode_mega_kernel, ops, (int)n_ops, (int*)grid_barrier, g_mega_trace));
    return B200_OK;
}
