// Swap-AB stream-K tcgen05 GEMM for pipeline-stage linears (see b200_petals.h).
//
//   out[t, n] = epi( sum_k norm(x)[t, k] * W[n, k] )
//
// Layout of one CTA (192 threads, persistent over a contiguous k-block range):
//   warp 0      weight loader   one lane; 16/32 KiB cp.async.bulk per k-block
//   warp 1      MMA issuer      one lane; tcgen05.mma 128 x BN x 16, fp32 in TMEM
//   warps 2..5  B producers     activations -> (norm) -> swizzled smem, then the
//                               same warps drain TMEM and run the epilogue
// The weight ring and the activation ring have independent depths, so the
// HBM stream (weights) is never throttled by the L2-latency-bound B side.
#include <stdlib.h>

#include "gemm_core.cuh"
#include "trace.cuh"

namespace b200 {

constexpr int kLLMaxCtasDev = 512;   // CTAs one half of the word buffer has slots for (b200_gemm_ll_bytes)

// WIDE = 1: the ticket-free fix-up for 5..16 tokens (every CTA of a tile publishes its whole partial and then
// reduces its own share of the tile's elements).  A separate instantiation so that its registers never weigh on
// the batch-1 kernel.
// LLIO = 1: link of a flag-chained decode step (b200_ll_chain): no grid-dependency wait; x and the residual are polled
// as LL words, the output is published as LL words, the stream-K exchange uses per-launch tags in alternating halves
// of the word buffer (nothing to clear).
template <int BN, int DUAL, int FP8, int WIDE = 0, int LLIO = 0>
__global__ void __launch_bounds__(kThreads, GemmCfg<BN, DUAL>::kMinBlocks)
gemm_streamk_kernel(const GemmParams p) {
    constexpr int MX = (FP8 == 2) ? 1 : 0;      // FP8: 0 = 16-bit weights, 1 = e4m3 with one scale per row, 2 = MX block scales
    using Cfg = GemmCfg<BN, DUAL, MX>;
    constexpr int SA = Cfg::kSA, SB = Cfg::kSB, APER = Cfg::kAPer;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + (size_t)SA * Cfg::kStageA;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kSmemData);
    uint64_t* a_full = bars;
    uint64_t* a_empty = a_full + SA;
    uint64_t* b_full = a_empty + SA;
    uint64_t* b_empty = b_full + SB;
    uint64_t* acc_full = b_empty + SB;
    uint64_t* acc_empty = acc_full + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 1);
    volatile int* fin_flag = reinterpret_cast<volatile int*>(tmem_slot + 1);  // [2]
    float* stat_mean = reinterpret_cast<float*>(tmem_slot + 4);               // [32]
    float* stat_rstd = stat_mean + 32;                                        // [32]
    float* stat_red = stat_rstd + 32;                                         // [2][4][32]
    uint8_t* smem_sfb = smem + Cfg::kSmemData + 1024;                         // MX: activation scale factors, 512 B per k-block

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int cta = blockIdx.x;

    if (threadIdx.x == 0) {
        trace_event(1, 0, (p.n_tiles & 0xfff) | ((p.kbpt & 0xfff) << 12));
        for (int i = 0; i < SA; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 1); }
        for (int i = 0; i < SB; ++i) { mbar_init(&b_full[i], 4); mbar_init(&b_empty[i], 1); }
        mbar_init(acc_full, 1);
        mbar_init(acc_empty, 4);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // Let the next kernel in the stream get resident now: its weight loader does not
    // depend on our output and can fill its ring while we finish (it still waits for
    // our completion before touching activations).
    pdl_launch_dependents();

    const long long q_begin = range_begin(p, cta);
    const long long q_end = range_begin(p, cta + 1);

    if (warp == 0) {
        // ------------------------------------------------ weight loader
        // The whole warp walks the loop (warp-uniform control flow keeps addresses and barriers in
        // uniform registers); one elected lane issues.  A lane-0-only loop makes ptxas wrap every
        // bulk copy / MMA in a divergent "waterfall" and roughly halves the per-CTA streaming rate.
        {
            // Weights never depend on the previous kernel: start streaming at once.
            const uint64_t policy = (p.t_tiles == 1) ? l2_policy_evict_first() : l2_policy_evict_last();
            // (u, kb) walk the global k-block index without divisions in the loop.
            int u = (int)(q_begin / p.kbpt);
            int kb = (int)(q_begin - (long long)u * p.kbpt);
            int n_tile = u / p.t_tiles, t_tile = u - n_tile * p.t_tiles;
            long long left = q_end - q_begin;
            int s = 0;
            uint32_t ph = 0;
            const uint8_t* tile_base = p.w + (size_t)n_tile * p.kbpt * Cfg::kBlockA;
            while (left > 0) {
                int n = p.kbpt - kb;                          // stay inside this tile ...
                if ((long long)n > left) n = (int)left;       // ... and inside our range
                if (n > Cfg::kKps) n = Cfg::kKps;
                mbar_wait(&a_empty[s], ph ^ 1);
                const uint32_t bytes = (uint32_t)n * Cfg::kBlockA;
                if (elect_one()) {
                    mbar_arrive_expect_tx(&a_full[s], bytes);
                    bulk_g2s_hint(smem_a + (size_t)s * Cfg::kStageA, tile_base + (size_t)kb * Cfg::kBlockA, bytes,
                                  &a_full[s], policy);
                }
                __syncwarp();
                if (++s == SA) { s = 0; ph ^= 1; }
                left -= n;
                kb += n;
                if (kb == p.kbpt) {
                    kb = 0;
                    if (++t_tile == p.t_tiles) { t_tile = 0; ++n_tile; tile_base += (size_t)p.kbpt * Cfg::kBlockA; }
                }
            }
            // Our own stream is issued; keep the HBM busy through the fix-up / epilogue and the next launch
            // by pulling the head of the following kernels' weights into L2 (16 KiB pieces, CTA-interleaved).
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                constexpr long long kPiece = 16384;
                const long long pieces = (p.pf_bytes[r] + kPiece - 1) / kPiece;
                for (long long c = cta; c < pieces; c += gridDim.x) {
                    const long long off = c * kPiece;
                    const long long n = p.pf_bytes[r] - off < kPiece ? p.pf_bytes[r] - off : kPiece;
                    if (elect_one()) bulk_prefetch_l2(p.pf_ptr[r] + off, (uint32_t)n);
                    __syncwarp();
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------ MMA issuer (whole warp, one elected lane issues)
        {
            const uint32_t idesc = FP8 ? umma_idesc(kTileM, BN, 0, 0, 0, 0)   // e4m3 x e4m3
                                            : umma_idesc(kTileM, BN, p.is_bf16, p.is_bf16, 0, 0);
            const uint32_t sf_base = tmem_base + BN * APER * Cfg::kNAcc;      // MX: scale-factor columns behind the accumulators
            const uint32_t sfb_smem = smem_u32(smem_sfb);
            uint32_t sf_slot = 0;
            int u = (int)(q_begin / p.kbpt);
            int kb0 = (int)(q_begin - (long long)u * p.kbpt);
            long long left = q_end - q_begin;
            int sa = 0, sb = 0;
            uint32_t pha = 0, phb = 0, php = 0;
            const uint32_t a_base = smem_u32(smem_a), b_base = smem_u32(smem_b);
            while (left > 0) {
                int nk = p.kbpt - kb0;
                if ((long long)nk > left) nk = (int)left;      // k-blocks of this part
                mbar_wait(acc_empty, php ^ 1);
                tc_fence_after();
                uint32_t acc = 0;
                int done = 0, bofs = 0;
                while (done < nk) {
                    int na = nk - done;
                    if (na > Cfg::kKps) na = Cfg::kKps;
                    mbar_wait(&a_full[sa], pha);
                    for (int j = 0; j < na; ++j) {
                        if (bofs == 0 && !(p.debug & 2)) mbar_wait(&b_full[sb], phb);
                        tc_fence_after();
                        const uint64_t bdesc = umma_desc_kmajor_sw128(b_base + sb * Cfg::kStageB + bofs * Cfg::kSubB, 1024);
                        if (MX && !(p.debug & 1) && elect_one()) {
                            // scale factors of this k-block: smem -> tensor memory (in order with the MMAs that use them)
                            const uint32_t cols = sf_base + sf_slot * (APER * 4 + 4);
                            const uint32_t sfa_smem = a_base + sa * Cfg::kStageA + j * Cfg::kBlockA + APER * kABytes;
#pragma unroll
                            for (int a = 0; a < APER; ++a)
                                tmem_cp_32x128b(cols + a * 4, umma_desc_kmajor_noswizzle(sfa_smem + a * kMxSfBytes, 128, 128));
                            tmem_cp_32x128b(cols + APER * 4,
                                            umma_desc_kmajor_noswizzle(sfb_smem + (sb * Cfg::kGroup + bofs) * kMxSfBytes, 128, 128));
#pragma unroll
                            for (int a = 0; a < APER; ++a) {
                                const uint64_t adesc = umma_desc_kmajor_sw128(
                                    a_base + sa * Cfg::kStageA + j * Cfg::kBlockA + a * kABytes, 1024);
#pragma unroll
                                for (int k = 0; k < kBK / 16; ++k) {    // four 32-element steps, one scale byte each
                                    const uint32_t d_tmem = tmem_base + (a * Cfg::kNAcc + (k % Cfg::kNAcc)) * BN;
                                    umma_mxf8(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2),
                                              umma_idesc_mxf8(kTileM, BN, k, k), acc, cols + a * 4, cols + APER * 4);
                                }
                            }
                        }
                        if (MX) sf_slot = (sf_slot + 1) % kMxSfSlots;
                        if (!MX && !(p.debug & 1) && elect_one()) {
#pragma unroll
                            for (int a = 0; a < APER; ++a) {
                                const uint64_t adesc = umma_desc_kmajor_sw128(
                                    a_base + sa * Cfg::kStageA + j * Cfg::kBlockA + a * kABytes, 1024);
#pragma unroll
                                for (int k = 0; k < kBK / 16; ++k) {
                                    // +32 B per UMMA_K inside the 128 B swizzle row -> +2 in the address field
                                    const uint32_t d_tmem = tmem_base + (a * Cfg::kNAcc + (k % Cfg::kNAcc)) * BN;
                                    const uint32_t accf = (k >= Cfg::kNAcc) ? 1u : acc;
                                    if (FP8) umma_f8f6f4(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, accf);
                                    else umma_f16(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, accf);
                                }
                            }
                        }
                        acc = 1;
                        ++bofs;
                        __syncwarp();
                        if (bofs == Cfg::kGroup || done + j + 1 == nk) {
                            if (!(p.debug & 2) && elect_one()) umma_commit(&b_empty[sb]);
                            __syncwarp();
                            bofs = 0;
                            if (++sb == SB) { sb = 0; phb ^= 1; }
                        }
                    }
                    if (elect_one()) umma_commit(&a_empty[sa]);
                    __syncwarp();
                    if (++sa == SA) { sa = 0; pha ^= 1; }
                    done += na;
                }
                if (elect_one()) umma_commit(acc_full);
                __syncwarp();
                php ^= 1;
                left -= nk;
                kb0 = 0;   // every later part starts at the beginning of a tile
            }
        }
        __syncwarp();
    } else {
        // ------------------------------------------------ B producers + epilogue
        const int e = threadIdx.x - 64;         // 0..127
        const int lg = warp & 3;                // TMEM lane quarter this warp may read
        const int row = lg * 32 + lane;         // accumulator row (weight row in tile)
        const int ch = e & 7;                   // 16-byte chunk of the 128 B k-row
        const int r0 = e >> 3;                  // first token row this thread stages

        uint32_t xtag = 0, rtag = 0, otag = 0;
        if (LLIO) {
            // no grid dependency: the words this CTA needs are polled where they are used
            const int seq = ll_read_seq(p.ch.seq);
            xtag = ll_tag(seq, p.ch.x_op); rtag = ll_tag(seq, p.ch.res_op); otag = ll_tag(seq, p.ch.out_op);
        } else {
            pdl_wait_prior_grid();  // activations come from the previous kernel
        }
        if (LLIO) {
            // wait until the input words exist - the whole vector when a norm needs its statistics anyway, else the
            // k-range this CTA multiplies; afterwards they are read with plain pipelined loads
            int w0 = 0, w1 = p.K >> 1;
            if (p.norm_mode == B200_NORM_NONE) {
                const int kb_lo = (int)(q_begin % p.kbpt);
                const long long span = q_end - q_begin;
                if (kb_lo + span <= p.kbpt) { w0 = kb_lo * (kBK / 2); w1 = (kb_lo + (int)span) * (kBK / 2); }
            }
            ll_validate_range(p.ch.x, p.ch.ldx, min(BN, p.T), w0, w1, xtag, e, kEpiThreads, 1);
        }
        if (e == 0) trace_event(1, 1, (p.n_tiles & 0xfff) | ((p.kbpt & 0xfff) << 12));
        if (p.gather_world > 1) tp_gather_prologue(p, e);   // x = residual + sum of the TP partials of the previous linear

        if (p.norm_mode == B200_NORM_LAYER) {
            // Per-token statistics (decode-size T only; host guarantees t_tiles == 1, BN <= 32).
            // RMSNorm needs no pre-pass: its 1/rms is applied in the epilogue (see rms_stats_128).
            const int pw = warp - 2;
            for (int t = pw; t < BN; t += 4) {
                float s1 = 0.f, s2 = 0.f;
                if (t < p.T) {
                    const uint16_t* xr = p.x + (size_t)t * p.ldx;
                    for (int c = lane; c < p.K / 8; c += 32) {
                        const uint4 v = LLIO ? ll_load8_valid(p.ch.x + (size_t)t * p.ch.ldx + c * 4) : ld_cg_v4(xr + c * 8);
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
                        const float var = fmaxf(s2 * inv - m * m, 0.f);
                        stat_mean[t] = m;
                        stat_rstd[t] = rsqrtf(var + p.norm_eps);
                    }
                }
            }
            named_bar_sync(1, kEpiThreads);
        }

        if (FP8 == 1) fp8_stats_128(p, min(BN, p.T), e, stat_red, stat_mean, stat_rstd);   // qmul, epilogue scale
        uint32_t part = 0, phb = 0;
        int sb = 0;
        bool rms_ready = (FP8 == 1);   // per-row fp8: the 1/rms is already folded into stat_rstd
        const int tp_seq = (p.tp_world > 1) ? (__ldcv(p.tp_seq) + 1) : 0;
        bool tp_bumped = false;
        constexpr int ll_seq = 1;   // tag of a published word; the consumer clears the word again (tag 0) once it has read it
        // chain: the exchange words carry this launch's own tag and live in the half named by the caller - the
        // previous GEMM may still be reading the other half, and nothing has to be cleared
        const uint32_t wtag = LLIO ? otag : 1u;
        unsigned long long* const ll_words = LLIO && p.ll_words
            ? p.ll_words + (size_t)p.ch.half * ((size_t)kLLMaxCtasDev * 2 * 16 * kTileM) : p.ll_words;
        long long q = q_begin;
        while (q < q_end) {
            const int u = (int)(q / p.kbpt);
            const int kb0 = (int)(q - (long long)u * p.kbpt);
            const long long rem = q_end - q;
            const int kb1 = (rem < (long long)(p.kbpt - kb0)) ? kb0 + (int)rem : p.kbpt;
            const int n_tile = u / p.t_tiles;
            const int t_tile = u - n_tile * p.t_tiles;

            // ---- stage activations for k-blocks [kb0, kb1), kGroup k-blocks per B stage
            {
                constexpr int G = Cfg::kGroup, NI = BN / 16;
                uint4 cur[G][NI], cgw[G], cgb[G];
                uint4 nxt[Cfg::kPrefetch ? G : 1][Cfg::kPrefetch ? NI : 1], ngw[Cfg::kPrefetch ? G : 1], ngb[Cfg::kPrefetch ? G : 1];
                auto load_stage = [&](int kb, auto& v, auto& gw, auto& gb) {
#pragma unroll
                    for (int g = 0; g < G; ++g) {
                        if (kb + g < kb1) {
                            if (FP8) {
                                // 128 elements per k-block; this thread quantises 16 of them (two 16-byte loads)
                                const int kcol8 = (kb + g) * 128 + ch * 16;
#pragma unroll
                                for (int i = 0; i < NI; ++i) {
                                    const int t = t_tile * BN + r0 + 16 * i;
                                    uint4 lo = make_uint4(0, 0, 0, 0), hi = lo;
                                    if (t < p.T) {
                                        lo = ld_cg_v4(p.x + (size_t)t * p.ldx + kcol8);
                                        hi = ld_cg_v4(p.x + (size_t)t * p.ldx + kcol8 + 8);
                                    }
                                    uint4 gl = make_uint4(0, 0, 0, 0), gh = gl;
                                    if (p.norm_mode != B200_NORM_NONE) {
                                        gl = ld_nc_v4(p.norm_w + kcol8);
                                        gh = ld_nc_v4(p.norm_w + kcol8 + 8);
                                    }
                                    const float qm = MX ? 1.0f : ((t < p.T) ? stat_mean[r0 + 16 * i] : 0.f);
                                    const uint32_t xs[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
                                    const uint32_t gs[8] = {gl.x, gl.y, gl.z, gl.w, gh.x, gh.y, gh.z, gh.w};
                                    float f[16];
#pragma unroll
                                    for (int w = 0; w < 8; ++w) {
                                        float a = h16_to_float((uint16_t)(xs[w] & 0xffff), p.is_bf16);
                                        float b = h16_to_float((uint16_t)(xs[w] >> 16), p.is_bf16);
                                        if (p.norm_mode != B200_NORM_NONE) {
                                            a *= h16_to_float((uint16_t)(gs[w] & 0xffff), p.is_bf16);
                                            b *= h16_to_float((uint16_t)(gs[w] >> 16), p.is_bf16);
                                        }
                                        f[2 * w] = a * qm;
                                        f[2 * w + 1] = b * qm;
                                    }
                                    if (MX) {
                                        // one UE8M0 scale per 32 elements = this thread's 16 and its neighbour's (ch ^ 1):
                                        // 2^e with e = ceil(log2(amax / 448)), so that every value fits e4m3
                                        float am = 0.f;
#pragma unroll
                                        for (int w = 0; w < 16; ++w) am = fmaxf(am, fabsf(f[w]));
                                        am = fmaxf(am, __shfl_xor_sync(0xffffffffu, am, 1));
                                        int ex = -127;
                                        if (am > 0.f) {
                                            const float m = frexpf(am * (1.0f / 448.0f), &ex);     // am/448 = m * 2^ex, m in [0.5, 1)
                                            if (m == 0.5f) --ex;
                                            ex = max(-127, min(126, ex));
                                        }
                                        const float inv = __uint_as_float((uint32_t)(127 - ex) << 23);   // 2^-ex
#pragma unroll
                                        for (int w = 0; w < 16; ++w) f[w] *= inv;
                                        const uint32_t sfb = (uint32_t)(ex + 127) << (8 * i);
                                        if (i == 0) gw[g].x = sfb; else gw[g].x |= sfb;
                                    }
                                    v[g][i] = make_uint4(pack_e4m3x4(f[0], f[1], f[2], f[3]), pack_e4m3x4(f[4], f[5], f[6], f[7]),
                                                         pack_e4m3x4(f[8], f[9], f[10], f[11]), pack_e4m3x4(f[12], f[13], f[14], f[15]));
                                }
                                continue;
                            }
                            const int kcol = (kb + g) * kBK + ch * 8;
#pragma unroll
                            for (int i = 0; i < NI; ++i) {
                                const int t = t_tile * BN + r0 + 16 * i;
                                if (t >= p.T) v[g][i] = make_uint4(0, 0, 0, 0);
                                else if (LLIO) v[g][i] = ll_load8_valid(p.ch.x + (size_t)t * p.ch.ldx + (kcol >> 1));
                                else v[g][i] = ld_cg_v4(p.x + (size_t)t * p.ldx + kcol);
                            }
                            if (p.norm_mode != B200_NORM_NONE) {
                                gw[g] = ld_nc_v4(p.norm_w + kcol);
                                gb[g] = p.norm_b ? ld_nc_v4(p.norm_b + kcol) : make_uint4(0, 0, 0, 0);
                            }
                        }
                    }
                };
                if (!(p.debug & 2)) load_stage(kb0, cur, cgw, cgb);
                for (int kb = kb0; kb < kb1 && !(p.debug & 2); kb += G) {
                    if (Cfg::kPrefetch && kb + G < kb1) load_stage(kb + G, nxt, ngw, ngb);
                    mbar_wait(&b_empty[sb], phb ^ 1);
#pragma unroll
                    for (int g = 0; g < G; ++g) {
                        if (kb + g < kb1) {
                            const uint32_t base = smem_u32(smem_b + (size_t)sb * Cfg::kStageB + (size_t)g * Cfg::kSubB);
#pragma unroll
                            for (int i = 0; i < NI; ++i) {
                                const int r = r0 + 16 * i;
                                uint4 val = cur[g][i];
                                if (p.norm_mode != B200_NORM_NONE && !FP8) {
                                    const int t = t_tile * BN + r;
                                    if (t < p.T) {
                                        const float mean = stat_mean[r], rstd = stat_rstd[r];
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
                                                // HF LlamaRMSNorm: weight * (x * rstd).to(dtype)
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
                                }
                                st_shared_v4(base + r * 128 + ((ch ^ (r & 7)) << 4), val);
                                if (MX && !(ch & 1)) {   // the pair's scale byte: atom position (row % 32) * 16 + (row / 32) * 4 + group
                                    const uint32_t sfa = smem_u32(smem_sfb) + (sb * G + g) * kMxSfBytes + (r & 31) * 16 + (r >> 5) * 4 + (ch >> 1);
                                    asm volatile("st.shared.u8 [%0], %1;\n" ::"r"(sfa), "r"((cgw[g].x >> (8 * i)) & 0xffu) : "memory");
                                }
                            }
                        }
                    }
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&b_full[sb]);
                    if (++sb == SB) { sb = 0; phb ^= 1; }
                    if (Cfg::kPrefetch) {
#pragma unroll
                        for (int g = 0; g < G; ++g) {
#pragma unroll
                            for (int i = 0; i < NI; ++i) cur[g][i] = nxt[g][i];
                            cgw[g] = ngw[g];
                            cgb[g] = ngb[g];
                        }
                    } else if (kb + G < kb1) {
                        load_stage(kb + G, cur, cgw, cgb);
                    }
                }
            }

            // ---- tensor parallel: move the launch number on for the NEXT launch.  Done here, while the tensor core
            // still works on the staged k-blocks, instead of at CTA exit (a fence + atomic round trip there delays
            // the end of the kernel by 1.5-3 us).  Every thread of this CTA read the number before the barrier, the
            // CTA's add comes after it, and only the last add of the grid writes - so no CTA can see the new value.
            if (p.tp_world > 1 && !tp_bumped) {
                tp_bumped = true;
                named_bar_sync(1, kEpiThreads);
                if (e == 0 && atomicAdd(p.tp_exit, 1) == (int)gridDim.x - 1) {
                    *p.tp_exit = 0;
                    *p.tp_seq = tp_seq;
                }
            }
            // ---- deferred RMSNorm statistics: computed once, while the tensor core chews on the staged k-blocks
            if (p.norm_mode == B200_NORM_RMS && !rms_ready && p.rstd_in) {
                if (e < min(BN, p.T)) stat_rstd[e] = __ldcg(p.rstd_in + e);
                named_bar_sync(1, kEpiThreads);
                rms_ready = true;
            }
            if (p.norm_mode == B200_NORM_RMS && !rms_ready) {
                rms_stats_128<LLIO>(p, min(BN, p.T), e, stat_red, stat_rstd, xtag);
                rms_ready = true;
            }
            const float* tok_scale = (p.norm_mode == B200_NORM_RMS || FP8 == 1) ? stat_rstd : nullptr;
            // ---- drain the accumulator of this part
            const bool full_tile = (kb0 == 0) && (kb1 == p.kbpt);
            const int slot = (part == 0) ? 0 : 1;
            float* my_partial = p.scratch + ((size_t)cta * 2 + slot) * (size_t)(APER * BN * kTileM);
            mbar_wait(acc_full, part & 1);
            tc_fence_after();
            if (e == 0) trace_event(1, 2, (p.n_tiles & 0xfff) | ((p.kbpt & 0xfff) << 12));
            const int t_live = min(BN, p.T - t_tile * BN);  // token columns that exist in this tile
#pragma unroll 1
            for (int c = 0; c * 16 < t_live; ++c) {
                const int ncol = min(16, t_live - c * 16);
                uint32_t rg[16], ru[16];
                const uint32_t taddr = tmem_base + ((uint32_t)(lg * 32) << 16) + c * 16;
                float vg[16], vu[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) { vg[j] = 0.f; vu[j] = 0.f; }
#pragma unroll
                for (int ac = 0; ac < Cfg::kNAcc; ++ac) {
                    tmem_ld16(taddr + ac * BN, rg);
                    if (DUAL) tmem_ld16(taddr + (Cfg::kNAcc + ac) * BN, ru);
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        vg[j] += __uint_as_float(rg[j]);
                        if (DUAL) vu[j] += __uint_as_float(ru[j]);
                    }
                }
                if (full_tile) {
                    if (p.tp_world > 1) tp_allreduce_tile(p, n_tile, row, ncol, e, tp_seq, vg);
                    if (!(p.tp_world > 1 && p.tp_defer))
                    epilogue_store_chunk<BN, DUAL, LLIO>(p, n_tile, t_tile, row, c, vg, vu, tok_scale, rtag, otag);
                } else if (WIDE && p.ll_words) {
                    // ---- wide ticket-free fix-up, part 1: publish the whole partial (reduced below, after the drain)
                    constexpr int kLLStrideW = APER * 16 * kTileM;        // words per CTA
                    uint2* mine = reinterpret_cast<uint2*>(ll_words) + (size_t)cta * kLLStrideW + row;
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        if (j < ncol) {
                            ll_store(mine + j * kTileM, vg[j], (int)wtag);
                            if (DUAL) ll_store(mine + (16 + j) * kTileM, vu[j], (int)wtag);
                        }
                    }
                } else if (p.ll_words) {
                    // ---- ticket-free fix-up (host guarantees T <= 4, one part per CTA, ll_parts CTAs per tile)
                    constexpr int kLLStride = APER * 4 * kTileM;          // words per CTA
                    const int P = p.ll_parts;
                    const int first = u * P;                              // CTAs [first, first + P) share tile u
                    if (cta != first + P - 1) {
                        uint2* mine = reinterpret_cast<uint2*>(p.ll_words) + (size_t)cta * kLLStride + row;
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            if (j < ncol) {
                                ll_store(mine + j * kTileM, vg[j], ll_seq);
                                if (DUAL) ll_store(mine + (4 + j) * kTileM, vu[j], ll_seq);
                            }
                        }
                    } else {
                        // last slice of the tile: sum p_0 + ... + p_{P-2} + own, in that order (same order as the
                        // ticket path).  Words are polled eight at a time so the L2 round trips overlap.
                        float sg[4] = {0.f, 0.f, 0.f, 0.f}, su[4] = {0.f, 0.f, 0.f, 0.f};
                        const uint2* base = reinterpret_cast<const uint2*>(p.ll_words) + (size_t)first * kLLStride + row;
                        const int lanes_k = APER * ncol;                  // (accumulator, column) pairs of this thread
                        int k = 0, r = 0;                                 // walk (k, r) with r fastest
                        const int total = lanes_k * (P - 1);
                        for (int done = 0; done < total; done += 8) {
                            const uint2* src[8];
                            int kk[8];
                            const int nb = min(8, total - done);
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                const int a = (k >= ncol) ? 1 : 0, j = k - a * ncol;
                                src[i] = base + (size_t)r * kLLStride + (a * 4 + j) * kTileM;
                                kk[i] = k;
                                if (i < nb && ++r == P - 1) { r = 0; ++k; }
                            }
                            uint32_t val[8], tag[8];
                            long long t0 = 0;
                            for (;;) {
                                bool ok = true;
#pragma unroll
                                for (int i = 0; i < 8; ++i)
                                    if (i < nb)
                                        asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];\n"
                                                     : "=r"(val[i]), "=r"(tag[i]) : "l"(src[i]) : "memory");
#pragma unroll
                                for (int i = 0; i < 8; ++i)
                                    if (i < nb) ok = ok && ((int)tag[i] == ll_seq);
                                if (ok) break;
                                if (t0 == 0) t0 = clock64();
                                else if (clock64() - t0 > 8000000000LL) {
                                    printf("b200: stream-K partial exchange timed out (block %d thread %d seq %d)\n",
                                           blockIdx.x, threadIdx.x, ll_seq);
                                    __trap();
                                }
                            }
#pragma unroll
                            for (int i = 0; i < 8; ++i)   // leave the slot empty for the next launch (ordered by the kernel boundary)
                                if (i < nb)
                                    asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %1};\n" ::"l"(src[i]), "r"(0) : "memory");
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                if (i < nb) {
                                    const float v = __uint_as_float(val[i]);
                                    const int a = (kk[i] >= ncol) ? 1 : 0, j = kk[i] - a * ncol;
#pragma unroll
                                    for (int jj = 0; jj < 4; ++jj) {
                                        if (jj == j) { if (a) su[jj] += v; else sg[jj] += v; }
                                    }
                                }
                            }
                        }
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            vg[j] = sg[j] + vg[j];
                            if (DUAL) vu[j] = su[j] + vu[j];
                        }
                        if (p.tp_world > 1) tp_allreduce_tile(p, n_tile, row, ncol, e, tp_seq, vg);
                        if (!(p.tp_world > 1 && p.tp_defer))
                        epilogue_store_chunk<BN, DUAL>(p, n_tile, t_tile, row, c, vg, vu, tok_scale);
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        if (j < ncol) {
                            my_partial[(size_t)(c * 16 + j) * kTileM + row] = vg[j];
                            if (DUAL) my_partial[(size_t)(BN + c * 16 + j) * kTileM + row] = vu[j];
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acc_empty);

            if (WIDE && !full_tile && p.ll_words) {
                // ---- wide fix-up, part 2: this CTA reduces its share of the tile's 128 x T elements (column-major
                // index j * 128 + row): polls the P words of an element together, adds them in CTA order (the order
                // of every other path), clears them (not in a chain: tags are per launch there), applies the epilogue
                // for that single element.  Shares are even-sized and every lane of a warp walks the loop, so that in
                // a chain two neighbouring lanes can publish their two 16-bit results as one LL word.
                constexpr int kLLStrideW = APER * 16 * kTileM;
                const int P = p.ll_parts;
                const int first = u * P;
                const int E = kTileM * t_live;
                const int per = (((E + P - 1) / P) + 1) & ~1;
                const int e_lo = min(E, (cta - first) * per), e_hi = min(E, e_lo + per);
                uint2* base = reinterpret_cast<uint2*>(ll_words) + (size_t)first * kLLStrideW;
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
                                                 : "=r"(vgw[c]), "=r"(tgw[c]) : "l"(w0 + (size_t)c * kLLStrideW) : "memory");
                                    if (DUAL)
                                        asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];\n"
                                                     : "=r"(vuw[c]), "=r"(tuw[c])
                                                     : "l"(w0 + (size_t)c * kLLStrideW + 16 * kTileM) : "memory");
                                }
                            }
#pragma unroll
                            for (int c = 0; c < 9; ++c)
                                if (c < P) ok = ok && tgw[c] == wtag && (!DUAL || tuw[c] == wtag);
                            if (ok) break;
                            if (t0 == 0) t0 = clock64();
                            else if (clock64() - t0 > 8000000000LL) {
                                printf("b200: stream-K wide exchange timed out (block %d thread %d)\n", blockIdx.x, threadIdx.x);
                                __trap();
                            }
                        }
#pragma unroll
                        for (int c = 0; c < 9; ++c) {
                            if (c < P) {
                                sg += __uint_as_float(vgw[c]);
                                if (!LLIO)
                                    asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %1};\n" ::"l"(w0 + (size_t)c * kLLStrideW), "r"(0) : "memory");
                                if (DUAL) {
                                    su += __uint_as_float(vuw[c]);
                                    if (!LLIO)
                                        asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %1};\n" ::"l"(w0 + (size_t)c * kLLStrideW + 16 * kTileM), "r"(0) : "memory");
                                }
                            }
                        }
                    }
                    epilogue_store_scalar<DUAL, LLIO>(p, n_tile * kTileM + r, valid ? j : 0, sg, su,
                                                      (tok_scale && valid) ? tok_scale[j] : 1.0f, valid, rtag, otag);
                }
            }
            if (!full_tile && !p.ll_words) {
                // Publish the partial with ONE acq_rel ticket: the CTA barrier orders every thread's
                // partial stores before thread 0's release, and thread 0's acquire before every
                // thread's reads (cumulativity), so no per-thread membar.gl is needed (~1 us per kernel).
                named_bar_sync(1, kEpiThreads);
                if (e == 0) {
                    const int add = kb1 - kb0;
                    int old;
                    asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], %2;\n"
                                 : "=r"(old) : "l"(&p.counters[u]), "r"(add) : "memory");
                    fin_flag[part & 1] = (old + add == p.kbpt) ? 1 : 0;
                }
                named_bar_sync(1, kEpiThreads);
                if (fin_flag[part & 1]) {
                    // Last contributor of tile u: sum all partials in CTA order (deterministic).
                    const long long t_lo = (long long)u * p.kbpt, t_hi = t_lo + p.kbpt;
                    int c_first = (int)((t_lo * (long long)gridDim.x) / p.Q);
                    while (c_first + 1 < (int)gridDim.x && range_begin(p, c_first + 1) <= t_lo) ++c_first;
                    while (c_first > 0 && range_begin(p, c_first) > t_lo) --c_first;
#pragma unroll 1
                    for (int c = 0; c * 16 < t_live; ++c) {
                        const int ncol = min(16, t_live - c * 16);
                        float vg[16], vu[16];
#pragma unroll
                        for (int j = 0; j < 16; ++j) { vg[j] = 0.f; vu[j] = 0.f; }
#pragma unroll 4
                        for (int cc = c_first; cc < (int)gridDim.x; ++cc) {
                            const long long rb = range_begin(p, cc), re = range_begin(p, cc + 1);
                            if (rb >= t_hi) break;
                            if (re <= rb || re <= t_lo) continue;
                            const int cslot = (rb >= t_lo) ? 0 : 1;
                            const float* src = p.scratch + ((size_t)cc * 2 + cslot) * (size_t)(APER * BN * kTileM);
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                if (j < ncol) {
                                    vg[j] += __ldcg(src + (size_t)(c * 16 + j) * kTileM + row);
                                    if (DUAL) vu[j] += __ldcg(src + (size_t)(BN + c * 16 + j) * kTileM + row);
                                }
                            }
                        }
                        if (p.tp_world > 1) tp_allreduce_tile(p, n_tile, row, ncol, e, tp_seq, vg);
                        if (!(p.tp_world > 1 && p.tp_defer))
                        epilogue_store_chunk<BN, DUAL, LLIO>(p, n_tile, t_tile, row, c, vg, vu, tok_scale, rtag, otag);
                    }
                    if (e == 0) p.counters[u] = 0;  // leave the counters clean for the next launch
                }
            }
            q += kb1 - kb0;
            ++part;
        }
        if (e == 0) trace_event(1, 3, (p.n_tiles & 0xfff) | ((p.kbpt & 0xfff) << 12));
        if (p.tp_world > 1 && !tp_bumped) {   // a CTA without work still counts
            named_bar_sync(1, kEpiThreads);
            if (e == 0 && atomicAdd(p.tp_exit, 1) == (int)gridDim.x - 1) {
                *p.tp_exit = 0;
                *p.tp_seq = tp_seq;
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
}

template <int BN, int DUAL, int FP8 = 0, int WIDE = 0, int LLIO = 0>
static int launch_gemm(const GemmParams& p, int grid, bool pdl, cudaStream_t stream) {
    using Cfg = GemmCfg<BN, DUAL, (FP8 == 2) ? 1 : 0>;
    static_assert(Cfg::kSmemBytes <= 227 * 1024, "shared memory budget exceeded");
    static thread_local int configured_dev = -1;
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    if (configured_dev != dev) {
        B200_CUDA_CHECK(cudaFuncSetAttribute(gemm_streamk_kernel<BN, DUAL, FP8, WIDE, LLIO>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)Cfg::kSmemBytes));
        configured_dev = dev;
    }
    B200_CUDA_CHECK(launch(gemm_streamk_kernel<BN, DUAL, FP8, WIDE, LLIO>, dim3(grid), dim3(kThreads), Cfg::kSmemBytes,
                           stream, pdl, p));
    return B200_OK;
}

static int pick_token_tile(int T) {
    if (T <= 16) return 16;
    if (T <= 32) return 32;
    if (T <= 64) return 64;
    if (T <= 128) return 128;
    return 256;
}

}  // namespace b200

using namespace b200;

B200_DEFINE_TRACE_SETTER(b200_trace_set_gemm)

extern "C" int b200_gemm_debug_occupancy(void) {
    using Cfg = GemmCfg<16, 0>;
    cudaFuncSetAttribute(gemm_streamk_kernel<16, 0, 0, 1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::kSmemBytes);
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, gemm_streamk_kernel<16, 0, 0, 1, 0>);
    int n = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, gemm_streamk_kernel<16, 0, 0, 1, 0>, kThreads, Cfg::kSmemBytes);
    printf("gemm<16,0,0,1,0>: regs %d static %zu local %zu dyn %zu -> %d blocks/SM\n", fa.numRegs, fa.sharedSizeBytes,
           fa.localSizeBytes, (size_t)Cfg::kSmemBytes, n);
    return n;
}

extern "C" int32_t b200_gemm_token_tile(int32_t T, int32_t dual) {
    (void)dual;
    return pick_token_tile(T);
}

// ll_buf layout: [kLLHeaderBytes reserved][kLLMaxCtas x 32 KiB of (value, tag) words, all zero between launches]
constexpr int kLLMaxCtas = kLLMaxCtasDev;
constexpr int kLLHeaderBytes = 256;
// two halves: the launches of a flag chain alternate between them (b200_ll_chain.half)
extern "C" int64_t b200_gemm_ll_bytes(void) { return (int64_t)kLLHeaderBytes + 2 * (int64_t)kLLMaxCtas * 2 * 16 * kTileM * 8; }

extern "C" int64_t b200_gemm_scratch_bytes(int32_t max_tokens) {
    const int bn = pick_token_tile(max_tokens);
    // grid <= SM count; two partial slots per CTA; dual doubles the accumulator
    return (int64_t)sm_count_cached() * 2 /*CTAs per SM*/ * 2 * 2 * bn * kTileM * (int64_t)sizeof(float);
}

static int gemm_impl(const b200_gemm_args* a, const b200_tp_args* tp, void* stream_, const b200_tp_args* gather = nullptr,
                     const void* gather_residual = nullptr, const void* gather_bias = nullptr, int gather_ldr = 0);

extern "C" int b200_gemm(const b200_gemm_args* a, void* stream_) { return gemm_impl(a, nullptr, stream_); }

extern "C" int b200_gemm_allreduce(const b200_gemm_args* a, const b200_tp_args* tp, void* stream_) {
    if (!tp || tp->world < 2 || tp->world > 8 || tp->rank < 0 || tp->rank >= tp->world)
        return fail(B200_EINVAL, "b200_gemm_allreduce: bad tensor-parallel group (world %d)", tp ? tp->world : -1);
    if (!a) return fail(B200_EINVAL, "b200_gemm_allreduce: null args");
    if (a->T > 16 || a->dual || a->norm_mode != B200_NORM_NONE || a->epilogue != B200_EPI_NONE)
        return fail(B200_ELIMIT, "b200_gemm_allreduce: supports row-parallel linears with T <= 16 (no norm / activation)");
    if (a->N > tp->n_max || a->n_tiles > tp->tiles_max)
        return fail(B200_ELIMIT, "b200_gemm_allreduce: N=%d / %d tiles exceed the workspace (%d / %d)", a->N, a->n_tiles,
                    tp->n_max, tp->tiles_max);
    return gemm_impl(a, tp, stream_);
}

// Consumer of a deferred b200_gemm_allreduce (tp->mode == 1): before staging, every CTA rebuilds
// x[t, :] = residual[t, :] + bias + sum over ranks of the partial rows pushed by the producer, writes it to
// a->x (so later kernels can use it as their residual) and then runs the normal GEMM on it.
extern "C" int b200_gemm_gather(const b200_gemm_args* a, const b200_tp_args* tp, const void* residual, int32_t ldr,
                                const void* bias, void* stream_) {
    if (!a || !tp || !residual) return fail(B200_EINVAL, "b200_gemm_gather: null argument");
    if (tp->world < 2 || tp->world > 8) return fail(B200_EINVAL, "b200_gemm_gather: bad world %d", tp->world);
    if (a->T > 16 || a->K > tp->n_max || a->K % 512 != 0)
        return fail(B200_ELIMIT, "b200_gemm_gather: T <= 16 and K a multiple of 512 within the workspace (T=%d K=%d)", a->T, a->K);
    return gemm_impl(a, nullptr, stream_, tp, residual, bias, ldr);
}

static int gemm_impl(const b200_gemm_args* a, const b200_tp_args* tp, void* stream_, const b200_tp_args* gather,
                     const void* gather_residual, const void* gather_bias, int gather_ldr) {
    if (!a || !a->w_packed || !a->x || !a->out || !a->scratch || !a->counters)
        return fail(B200_EINVAL, "b200_gemm: null pointer argument");
    if (a->T <= 0 || a->N <= 0 || a->K <= 0 || (a->K % kBK) != 0)
        return fail(B200_EINVAL, "b200_gemm: bad sizes T=%d N=%d K=%d (K must be a multiple of 64)", a->T, a->N, a->K);
    if (a->n_tiles * kTileM < a->N)
        return fail(B200_EINVAL, "b200_gemm: n_tiles=%d does not cover N=%d", a->n_tiles, a->N);
    if ((a->ldx % 8) != 0 || (reinterpret_cast<uintptr_t>(a->x) & 15) != 0)
        return fail(B200_EINVAL, "b200_gemm: x must be 16-byte aligned with ldx %% 8 == 0");
    if (a->dual && a->epilogue != B200_EPI_SWIGLU)
        return fail(B200_EINVAL, "b200_gemm: dual-packed weights require the SwiGLU epilogue");
    if (!a->dual && a->epilogue == B200_EPI_SWIGLU)
        return fail(B200_EINVAL, "b200_gemm: SwiGLU epilogue requires dual-packed weights");
    const int bn = pick_token_tile(a->T);
    const int t_tiles = (a->T + bn - 1) / bn;
    if (a->fp8) {
        if (t_tiles != 1 || bn > 32) return fail(B200_ELIMIT, "b200_gemm: fp8 weights support T <= 32 per call (got %d)", a->T);
        if (a->K % 128 != 0) return fail(B200_EINVAL, "b200_gemm: fp8 weights need K %% 128 == 0 (got %d)", a->K);
        if (a->fp8 == 1 && !a->w_scale) return fail(B200_EINVAL, "b200_gemm: fp8 weights need w_scale");
        if (a->fp8 != 1 && a->fp8 != 2) return fail(B200_EINVAL, "b200_gemm: fp8 must be 0, 1 (per-row scales) or 2 (MX block scales)");
        if (a->norm_mode == B200_NORM_LAYER) return fail(B200_EINVAL, "b200_gemm: fp8 + LayerNorm prologue is not supported");
    }
    if (a->norm_mode != B200_NORM_NONE) {
        if (t_tiles != 1 || bn > 32)
            return fail(B200_EINVAL, "b200_gemm: fused norm prologue supports T <= 32 (got %d)", a->T);
        if (!a->norm_w) return fail(B200_EINVAL, "b200_gemm: norm_mode set without norm_w");
    }
    GemmParams p;
    p.w = static_cast<const uint8_t*>(a->w_packed);
    p.x = static_cast<const uint16_t*>(a->x);
    p.out = a->out;
    p.residual = static_cast<const uint16_t*>(a->residual);
    p.bias = static_cast<const uint16_t*>(a->bias);
    p.norm_w = static_cast<const uint16_t*>(a->norm_w);
    p.norm_b = static_cast<const uint16_t*>(a->norm_b);
    p.scratch = a->scratch;
    p.counters = a->counters;
    p.T = a->T; p.N = a->N; p.K = a->K;
    p.ldx = a->ldx; p.ldo = a->ldo; p.ldr = a->ldr;
    p.n_tiles = a->n_tiles; p.t_tiles = t_tiles; p.kbpt = a->fp8 ? a->K / 128 : a->K / kBK;
    p.is_fp8 = a->fp8 ? 1 : 0;
    p.w_scale = a->fp8 == 1 ? a->w_scale : nullptr;
    p.Q = (long long)a->n_tiles * t_tiles * p.kbpt;
    p.epilogue = a->epilogue; p.norm_mode = a->norm_mode; p.norm_eps = a->norm_eps;
    p.is_bf16 = (a->dtype == B200_BF16) ? 1 : 0;
    p.out_fp32 = a->out_fp32;
    p.debug = a->debug;
    for (int r = 0; r < 2; ++r) {
        p.pf_ptr[r] = static_cast<const uint8_t*>(a->prefetch[r]);
        p.pf_bytes[r] = a->prefetch[r] ? (a->prefetch_bytes[r] & ~15LL) : 0;
    }
    p.tp_world = 0;
    p.tp_defer = 0;
    if (tp) {
        p.tp_world = tp->world; p.tp_rank = tp->rank; p.tp_n_max = tp->n_max; p.tp_tiles_max = tp->tiles_max;
        for (int r = 0; r < tp->world; ++r) {
            p.tp_stage[r] = static_cast<float*>(tp->stage[r]);
            p.tp_flags[r] = static_cast<int*>(tp->flags[r]);
        }
        p.tp_seq = tp->seq;
        p.tp_exit = tp->exit_counter;
        p.tp_defer = tp->mode == 1 ? 1 : 0;
    }
    p.gather_world = 0;
    if (gather) {
        p.gather_world = gather->world; p.gather_rank = gather->rank;
        p.gather_n_max = gather->n_max; p.gather_tiles_max = gather->tiles_max;
        p.gather_stage = static_cast<const float*>(gather->stage[gather->rank]);
        p.gather_flags = static_cast<const int*>(gather->flags[gather->rank]);
        p.gather_seq = gather->seq;
        p.gather_residual = static_cast<const uint16_t*>(gather_residual);
        p.gather_bias = static_cast<const uint16_t*>(gather_bias);
        p.gather_ldr = gather_ldr;
    }

    // The decode tile keeps two CTAs per SM resident (111 KiB of shared memory each): twice as many
    // loader / MMA-issue threads per SM (measured 280 -> 297 tok/s on Llama-2-7B) and the next
    // kernel's CTAs can start under PDL while this kernel drains.  B200_GEMM_CTAS_PER_SM=1 restores one.
    static const int ctas_per_sm = [] { const char* e = getenv("B200_GEMM_CTAS_PER_SM"); return (e && atoi(e) == 1) ? 1 : 2; }();
    const int max_grid = sm_count_cached() * ((bn == 16 && a->fp8 != 1) ? ctas_per_sm : 1);
    int grid = a->grid > 0 ? a->grid : max_grid;
    if (grid > max_grid) grid = max_grid;
    if (a->grid <= 0 && t_tiles == 1) {
        // Few output tiles (o_proj, down_proj): give every tile the same whole number of CTAs so each
        // CTA writes one partial and each tile has equally many contributors (measured 17.4 -> 14.1 us
        // for 4096x4096 at 128 instead of 148 CTAs).
        int per_tile = grid / a->n_tiles;
        if (a->chain.seq) {   // a flag chain needs the wide exchange: at most 9 CTAs per tile, at least one k-block each
            if (per_tile > 9) per_tile = 9;
            if (per_tile > a->K / kBK) per_tile = a->K / kBK;
        }
        if (per_tile >= 2) grid = per_tile * a->n_tiles;
    }
    if ((long long)grid > p.Q) grid = (int)p.Q;
    // Ticket-free fix-up: only when every tile is shared by the same whole number of CTAs with one part each.
    p.ll_words = nullptr; p.ll_parts = 0;
    p.rstd_in = (a->norm_mode == B200_NORM_RMS && !a->fp8 && t_tiles == 1) ? a->rstd : nullptr;
    // 1..16 tokens: every CTA of a tile reduces a share of its elements (measured per-layer GEMM chain, T=2: 82.9 -> 78.1 us,
    // T=4: 93.7 -> 82.5, T=16: 138 -> 108; T=1: 359.1 -> 362.2 tok/s end of the decode bench).  With tensor parallelism or
    // fp8 weights (T <= 4) the tile's last CTA collects.
    static const int wide_min = [] { const char* e = getenv("B200_LL_WIDE_MIN"); return e ? atoi(e) : 1; }();
    const bool ll_wide = a->T >= wide_min && a->T <= 16 && !tp && !gather && a->fp8 != 1;
    if (a->ll_buf && bn == 16 && t_tiles == 1 && (a->T <= 4 || ll_wide) && grid % a->n_tiles == 0 && grid <= kLLMaxCtas) {
        const int parts = grid / a->n_tiles;
        if (parts >= 2 && p.kbpt >= parts && (!ll_wide || parts <= 9)) {
            p.ll_words = reinterpret_cast<unsigned long long*>(static_cast<uint8_t*>(a->ll_buf) + kLLHeaderBytes);
            p.ll_parts = parts;
        }
    }
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const bool pdl = a->pdl != 0;
    p.ch = LLChain{};
    if (a->chain.seq) {
        const b200_ll_chain& c = a->chain;
        if (bn != 16 || a->fp8 || tp || gather || a->rstd)
            return fail(B200_ELIMIT, "b200_gemm: a flag chain needs T <= 16, 16-bit weights, no tensor parallelism, no rstd");
        if (!c.x_ll || (a->residual && !c.res_ll) || (!a->out_fp32 && !c.out_ll))
            return fail(B200_EINVAL, "b200_gemm: flag chain without x_ll / res_ll / out_ll");
        if ((c.ldx_ll % 4) != 0 || (reinterpret_cast<uintptr_t>(c.x_ll) & 31) != 0)
            return fail(B200_EINVAL, "b200_gemm: chain.x_ll must be 32-byte aligned with ldx_ll %% 4 == 0");
        if (c.x_op <= 0 || c.x_op >= (1 << kLLOpBits) || c.out_op <= 0 || c.out_op >= (1 << kLLOpBits) || (c.half & ~1))
            return fail(B200_EINVAL, "b200_gemm: chain launch ids must be in 1..%d, half 0/1", (1 << kLLOpBits) - 1);
        if (!a->pdl) return fail(B200_EINVAL, "b200_gemm: the launches of a flag chain need pdl = 1");
        p.ch.seq = c.seq;
        p.ch.x = static_cast<const uint2*>(c.x_ll);
        p.ch.res = a->residual ? static_cast<const uint2*>(c.res_ll) : nullptr;
        p.ch.out = a->out_fp32 ? nullptr : static_cast<uint2*>(c.out_ll);
        p.ch.ldx = c.ldx_ll; p.ch.ldr = c.ldr_ll; p.ch.ldo = c.ldo_ll;
        p.ch.x_op = c.x_op; p.ch.res_op = c.res_op; p.ch.out_op = c.out_op; p.ch.half = c.half;
        if (p.ll_words && !ll_wide) { p.ll_words = nullptr; p.ll_parts = 0; }   // only the wide exchange is chain-safe
        if (p.ll_words)
            return a->dual ? launch_gemm<16, 1, 0, 1, 1>(p, grid, pdl, stream) : launch_gemm<16, 0, 0, 1, 1>(p, grid, pdl, stream);
        return a->dual ? launch_gemm<16, 1, 0, 0, 1>(p, grid, pdl, stream) : launch_gemm<16, 0, 0, 0, 1>(p, grid, pdl, stream);
    }
    if (a->fp8 == 2) {   // MX block-scaled e4m3: the decode-size tiles; split tiles meet through the wide exchange
        if (bn == 16 && p.ll_words && ll_wide)
            return a->dual ? launch_gemm<16, 1, 2, 1>(p, grid, pdl, stream) : launch_gemm<16, 0, 2, 1>(p, grid, pdl, stream);
        if (p.ll_words) { p.ll_words = nullptr; p.ll_parts = 0; }
        if (bn == 16) return a->dual ? launch_gemm<16, 1, 2>(p, grid, pdl, stream) : launch_gemm<16, 0, 2>(p, grid, pdl, stream);
        return a->dual ? launch_gemm<32, 1, 2>(p, grid, pdl, stream) : launch_gemm<32, 0, 2>(p, grid, pdl, stream);
    }
    if (a->fp8) {   // only the decode-size tiles are instantiated for e4m3 weights
        if (bn == 16) return a->dual ? launch_gemm<16, 1, 1>(p, grid, pdl, stream) : launch_gemm<16, 0, 1>(p, grid, pdl, stream);
        return a->dual ? launch_gemm<32, 1, 1>(p, grid, pdl, stream) : launch_gemm<32, 0, 1>(p, grid, pdl, stream);
    }
    if (p.ll_words && ll_wide)
        return a->dual ? launch_gemm<16, 1, 0, 1>(p, grid, pdl, stream) : launch_gemm<16, 0, 0, 1>(p, grid, pdl, stream);
#define B200_GEMM_CASE(BN_)                                                        \
    case BN_:                                                                      \
        return a->dual ? launch_gemm<BN_, 1>(p, grid, pdl, stream)                 \
                       : launch_gemm<BN_, 0>(p, grid, pdl, stream);
    switch (bn) {
        B200_GEMM_CASE(16)
        B200_GEMM_CASE(32)
        B200_GEMM_CASE(64)
        B200_GEMM_CASE(128)
        B200_GEMM_CASE(256)
    }
#undef B200_GEMM_CASE
    return fail(B200_EINVAL, "b200_gemm: unsupported token tile %d", bn);
}

// ---------------------------------------------------------------------------------------------
// b200_gemm_static_scratch: the same GEMM for hosts that may not allocate device memory.
//
// Who needs this: programs that give this library pointers but no allocator.  Typical cases are a C
// plugin inside an inference server whose own pool already owns the whole device (a cudaMalloc from
// here would fail or eat into a budget the host has promised to its sessions), and token loops with
// a latency contract, where a device allocation can stall the GPU at an arbitrary point.  Such a
// caller still passes x, the packed weight and out, but cannot provide the stream-K workspace that
// b200_gemm expects in args->scratch / args->counters.
//
// What it does instead: the workspace lives in g_gemm_static_ws below, a fixed array that is part of
// the loaded library image (every GPU that loads the library holds these ~5 MiB until the process
// exits).  Layout: [kStaticCounterBytes of tile counters][kStaticScratchBytes of fp32 partial tiles].
// The array name is only meaningful to device code; host code must translate it with
// cudaGetSymbolAddress, which this function does once per device, with that device current, on the
// first call made on that device.  That first call does the look-up (and may make the driver load the
// module), so make it before any stream capture is begun on that device.
//
// Limits (checked, B200_ELIMIT with a message otherwise): T <= 16 tokens, at most 2048 output tiles,
// at most 160 CTAs.  All calls on one device share the one array and it keeps whatever the last call
// left in it: the counters this call needs are cleared on the stream before the launch, the partial
// tiles need no clearing (they are written before they are read).  Two calls whose kernels overlap on
// the GPU (different streams, or different threads/processes on the same device) race on the array;
// keep calls to this entry point on one stream, or serialise them with events.
namespace b200 {
namespace {
constexpr int kStaticMaxTokens = 16;
constexpr int kStaticMaxCtas = 160;
constexpr int kStaticMaxUnits = 2048;
constexpr size_t kStaticCounterBytes = (size_t)kStaticMaxUnits * sizeof(int);                    // 8 KiB
constexpr size_t kStaticScratchBytes = (size_t)kStaticMaxCtas * 2 /*slots*/ * 2 /*gate+up*/ *
                                       kStaticMaxTokens * kTileM * sizeof(float);                // 5 MiB
static __device__ __align__(256) unsigned char g_gemm_static_ws[kStaticCounterBytes + kStaticScratchBytes];

constexpr int kMaxDevices = 64;
void* g_static_ws_addr[kMaxDevices] = {nullptr};

// Device address of g_gemm_static_ws on the CURRENT device (resolved once per device).
int static_ws_for_current_device(void** out) {
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(B200_ELIMIT, "b200_gemm_static_scratch: device index %d > %d", dev, kMaxDevices - 1);
    void* p = __atomic_load_n(&g_static_ws_addr[dev], __ATOMIC_ACQUIRE);
    if (p == nullptr) {
        void* resolved = nullptr;
        B200_CUDA_CHECK(cudaGetSymbolAddress(&resolved, g_gemm_static_ws));
        if (resolved == nullptr) return fail(B200_ECUDA, "b200_gemm_static_scratch: symbol address is null");
        // Concurrent first calls resolve the same address; last store wins harmlessly.
        __atomic_store_n(&g_static_ws_addr[dev], resolved, __ATOMIC_RELEASE);
        p = resolved;
    }
    *out = p;
    return B200_OK;
}
}  // namespace
}  // namespace b200

extern "C" int b200_gemm_static_scratch(const b200_gemm_args* a, void* stream_) {
    if (!a) return fail(B200_EINVAL, "b200_gemm_static_scratch: null args");
    if (a->T > kStaticMaxTokens)
        return fail(B200_ELIMIT, "b200_gemm_static_scratch: T=%d exceeds the built-in workspace limit of %d tokens "
                                 "(use b200_gemm with a caller-owned workspace)", a->T, kStaticMaxTokens);
    if (a->n_tiles > kStaticMaxUnits)
        return fail(B200_ELIMIT, "b200_gemm_static_scratch: %d output tiles exceed the built-in limit of %d",
                    a->n_tiles, kStaticMaxUnits);
    if (sm_count_cached() > kStaticMaxCtas)
        return fail(B200_ELIMIT, "b200_gemm_static_scratch: %d SMs exceed the %d CTAs the built-in workspace holds",
                    sm_count_cached(), kStaticMaxCtas);
    // Argument errors are reported before anything is put on the stream (the checks b200_gemm would make).
    if (!a->w_packed || !a->x || !a->out) return fail(B200_EINVAL, "b200_gemm_static_scratch: null pointer argument");
    if (a->T <= 0 || a->N <= 0 || a->K <= 0 || (a->K % kBK) != 0)
        return fail(B200_EINVAL, "b200_gemm_static_scratch: bad sizes T=%d N=%d K=%d (K must be a multiple of 64)", a->T, a->N, a->K);
    if (a->n_tiles * kTileM < a->N) return fail(B200_EINVAL, "b200_gemm_static_scratch: n_tiles=%d does not cover N=%d", a->n_tiles, a->N);
    if ((a->ldx % 8) != 0 || (reinterpret_cast<uintptr_t>(a->x) & 15) != 0)
        return fail(B200_EINVAL, "b200_gemm_static_scratch: x must be 16-byte aligned with ldx %% 8 == 0");
    if (a->chain.seq) return fail(B200_EINVAL, "b200_gemm_static_scratch: flag chains need the caller's own buffers (b200_gemm)");
    void* ws = nullptr;
    const int rc = static_ws_for_current_device(&ws);
    if (rc != B200_OK) return rc;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int n_units = a->n_tiles > 0 ? a->n_tiles : 1;   // T <= 16 -> one token tile per output tile
    // The counters must read zero when the kernel starts; whatever an earlier (possibly failed) call left is wiped.
    B200_CUDA_CHECK(cudaMemsetAsync(ws, 0, (size_t)n_units * sizeof(int), stream));
    b200_gemm_args local = *a;
    local.grid = sm_count_cached();   // one CTA per SM: the built-in workspace has slots for kStaticMaxCtas CTAs
    local.pdl = 0;                    // the node before the kernel is the memset above: nothing to overlap with
    local.ll_buf = nullptr;           // the partial tiles meet in the built-in array only
    local.counters = static_cast<int32_t*>(ws);
    local.scratch = reinterpret_cast<float*>(static_cast<unsigned char*>(ws) + kStaticCounterBytes);
    return b200_gemm(&local, stream_);
}
