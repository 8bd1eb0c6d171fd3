// Shared device-side body of the split-KV decode attention (see attention.cu for the contract).
#pragma once
#include "common.h"
#include "ptx.cuh"
#include "trace.cuh"
#include "llchain.cuh"

namespace b200 {

struct AttnDecodeParams {
    const uint16_t* __restrict__ qkv;
    uint16_t* __restrict__ out;
    const unsigned long long* __restrict__ kv_ptrs;  // [T][2]
    const int* __restrict__ pos;                     // [T]
    const float* __restrict__ inv_freq;              // [D/2] or null
    float* __restrict__ scratch;                     // [T][n_heads][splits][D+2]
    int* __restrict__ counters;                      // [T][n_kv]
    unsigned long long* __restrict__ ll;             // optional [T][n_heads][splits][D+2] (value, tag) words, or null
    int T, n_heads, n_kv, s_max, splits, ldqkv, ldo;
    float scale;
    int is_bf16;
    LLChain ch;   // flag-chained decode step (ch.seq != null): qkv polled from ch.x, the output also published to ch.out
};

constexpr int kAttnThreads = 128;
#ifndef B200_ATTN_MIN_ROWS
#define B200_ATTN_MIN_ROWS 32   /* one pass of a CTA covers 32 cache rows; measured 64 -> 32: +0.6 % decode tokens/s */
#endif
constexpr int kMinRowsPerSplit = B200_ATTN_MIN_ROWS;

// (value, tag) words of the ticket-free split merge: tag 1 = published, 0 = empty (cleared by the consumer).
__device__ __forceinline__ void ll_word_store(uint2* dst, float v) {
    asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %2};\n" ::"l"(dst), "r"(__float_as_uint(v)), "r"(1u) : "memory");
}
__device__ __forceinline__ void ll_word_load(const uint2* src, uint32_t& v, uint32_t& tag) {
    asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];\n" : "=r"(v), "=r"(tag) : "l"(src) : "memory");
}
__device__ __forceinline__ void ll_word_clear(uint2* dst) {
    asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %1};\n" ::"l"(dst), "r"(0u) : "memory");
}

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8], int is_bf16) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        f[2 * i] = h16_to_float((uint16_t)(w[i] & 0xffff), is_bf16);
        f[2 * i + 1] = h16_to_float((uint16_t)(w[i] >> 16), is_bf16);
    }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8], int is_bf16) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
        w[i] = (uint32_t)float_to_h16(f[2 * i], is_bf16) | ((uint32_t)float_to_h16(f[2 * i + 1], is_bf16) << 16);
    return make_uint4(w[0], w[1], w[2], w[3]);
}

// Rotates one head vector held as (lo = dims [c, c+8), hi = dims [c+D/2, c+D/2+8)).
// HF convention: out = x*cos + rotate_half(x)*sin with cos/sin rounded to the model dtype.
__device__ __forceinline__ void rope_sincos8(int c, int pos, const float* inv_freq, int is_bf16, float (&sn)[8],
                                             float (&cs)[8]) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float s_, c_;
        sincosf((float)pos * inv_freq[c + i], &s_, &c_);
        cs[i] = h16_to_float(float_to_h16(c_, is_bf16), is_bf16);
        sn[i] = h16_to_float(float_to_h16(s_, is_bf16), is_bf16);
    }
}
__device__ __forceinline__ void rope_apply8(float (&lo)[8], float (&hi)[8], const float (&sn)[8], const float (&cs)[8]) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float a = lo[i], b = hi[i];
        lo[i] = a * cs[i] - b * sn[i];
        hi[i] = b * cs[i] + a * sn[i];
    }
}
__device__ __forceinline__ void rope8(float (&lo)[8], float (&hi)[8], int c, int pos, const float* inv_freq,
                                      int is_bf16) {
    float sn[8], cs[8];
    rope_sincos8(c, pos, inv_freq, is_bf16, sn, cs);
    rope_apply8(lo, hi, sn, cs);
}

// One CTA = (token t, kv head g, split s).  LPR = D/8 lanes cooperate on one KV row.
template <int D, int GROUP>
struct AttnDecodeSmem {
    float q[GROUP][D];
    float m[kAttnThreads / 32][GROUP], l[kAttnThreads / 32][GROUP];
    float o[kAttnThreads / 32][GROUP][D];
    int last;
};

// Cache rows of a split staged through shared memory by bulk async copies (cp.async.bulk, SASS UBLKCP): one pass of
// the CTA = kAttnRowsPerPass<D> rows = 8 KiB of K and 8 KiB of V, kAttnStages passes in flight.  The rows of earlier tokens do
// are requested as soon as the previous kernels are known to be complete (their L2 prefetch is issued above the
// programmatic-launch wait) and land while q is rotated; later passes are copied while the current one is reduced.
template <int D>
constexpr int kAttnRowsPerPass = (kAttnThreads / 32) * (32 / (D / 8)) * 4;   // warps x rows per warp step x unroll
#ifndef B200_ATTN_BULK_STAGES
#define B200_ATTN_BULK_STAGES 3   /* 48 KiB of tiles per CTA: four CTAs per SM stay resident (>= 512 at batch 1) */
#endif
constexpr int kAttnStages = B200_ATTN_BULK_STAGES;
template <int D>
struct AttnBulkTiles {
    alignas(128) uint16_t k[kAttnStages][kAttnRowsPerPass<D> * D];
    alignas(128) uint16_t v[kAttnStages][kAttnRowsPerPass<D> * D];
    alignas(16) uint16_t newk[D];      // the new token's own row (RoPE'd k, v): its owner split has it in registers
    alignas(16) uint16_t newv[D];
    alignas(8) uint64_t full[kAttnStages];
};

// One work item = (token t, kv head g, split s), executed by kAttnThreads threads that share
// `sm` and synchronise through `sync()` (a __syncthreads() or a named barrier).
// `before_qkv()` runs right before the first access to p.qkv (the only input the previous kernel produces):
// the stand-alone kernel waits for the previous grid there, so everything above it - positions, cache
// pointers, the RoPE sines and cosines, an L2 prefetch of this split's cache rows - overlaps that grid's tail.
struct AttnNoHook { __device__ __forceinline__ void operator()() const {} };
template <int D, int GROUP, typename Sync, typename Hook = AttnNoHook>
__device__ __forceinline__ void attn_decode_body(const AttnDecodeParams& p, int t, int g, int s, int tid,
                                                 AttnDecodeSmem<D, GROUP>& sm, Sync sync, Hook before_qkv = Hook(),
                                                 AttnBulkTiles<D>* tiles = nullptr) {
    constexpr int LPR = D / 8;          // lanes per row: 16 (D=128) or 8 (D=64)
    constexpr int RPW = 32 / LPR;       // rows per warp per step: 2 or 4
    constexpr int NW = kAttnThreads / 32;
    constexpr int UNROLL = 4;
    constexpr int ROWS = NW * RPW * UNROLL;     // rows per pass of the CTA
    static_assert(ROWS == kAttnRowsPerPass<D>, "tile size");
    auto& s_q = sm.q;
    auto& s_m = sm.m;
    auto& s_l = sm.l;
    auto& s_o = sm.o;
    int& s_last = sm.last;
    const int splits = p.splits;
    const int warp = tid >> 5, lane = tid & 31;

    const int pos = p.pos[t];
    const int len = pos + 1;  // keys visible to the new token, itself included
    int eff = (len + kMinRowsPerSplit - 1) / kMinRowsPerSplit;
    if (eff > splits) eff = splits;
    if (s >= eff) return;
    const int chunk = (len + eff - 1) / eff;
    const int r_begin = s * chunk;
    const int r_end = min(len, r_begin + chunk);

    uint16_t* kbase = reinterpret_cast<uint16_t*>(p.kv_ptrs[2 * t]) + (size_t)g * p.s_max * D;
    uint16_t* vbase = reinterpret_cast<uint16_t*>(p.kv_ptrs[2 * t + 1]) + (size_t)g * p.s_max * D;
    const uint16_t* row = p.qkv + (size_t)t * p.ldqkv;
    const int q_off = g * GROUP * D;
    const int k_off = p.n_heads * D + g * D;
    const int v_off = (p.n_heads + p.n_kv) * D + g * D;

    // ---- q (RoPE'd, pre-scaled) -> smem; new k/v -> cache (by the split that owns `pos`)
    // D/2/8 threads per head vector, each owning dims [c, c+8) and [c+D/2, c+D/2+8).
    constexpr int TPH = D / 16;
    static_assert((GROUP + 1) * TPH <= kAttnThreads, "one RoPE item per thread");
    float rs[8], rc[8];
    if (p.inv_freq && tid < (GROUP + 1) * TPH) rope_sincos8((tid % TPH) * 8, pos, p.inv_freq, p.is_bf16, rs, rc);
    const bool bulk = tiles != nullptr;
    const int n_old = max(0, min(r_end, pos) - r_begin);      // rows written by earlier steps
    const int n_pass = bulk ? (n_old + ROWS - 1) / ROWS : 0;
    auto issue = [&](int pass, int st) {                       // one thread: request pass `pass` into stage `st`
        const int rows = min(ROWS, n_old - pass * ROWS);
        const uint32_t bytes = (uint32_t)(rows * D * 2);
        mbar_arrive_expect_tx(&tiles->full[st], 2 * bytes);
        bulk_g2s(tiles->k[st], kbase + (size_t)(r_begin + pass * ROWS) * D, bytes, &tiles->full[st]);
        bulk_g2s(tiles->v[st], vbase + (size_t)(r_begin + pass * ROWS) * D, bytes, &tiles->full[st]);
    };
    if (tid == 0) {   // cache rows of earlier tokens do not depend on the previous kernel: start fetching them
        if (bulk) {
#pragma unroll
            for (int st = 0; st < kAttnStages; ++st) mbar_init(&tiles->full[st], 1);
            fence_mbar_init();
        }
        if (n_old > 0) {   // (a hint only: the copies themselves wait for the previous kernels, see below)
            bulk_prefetch_l2(kbase + (size_t)r_begin * D, (uint32_t)(n_old * D * 2));
            bulk_prefetch_l2(vbase + (size_t)r_begin * D, (uint32_t)(n_old * D * 2));
        }
    }
    before_qkv();
    const bool chain = p.ch.seq != nullptr;
    uint32_t xtag = 0, otag = 0;
    const uint2* row_ll = nullptr;
    if (chain) {
        const int seq = ll_read_seq(p.ch.seq);
        xtag = ll_tag(seq, p.ch.x_op);
        otag = ll_tag(seq, p.ch.out_op);
        row_ll = p.ch.x + (size_t)t * p.ch.ldx;
    }
    if (chain) {   // validate this item's q heads, k and v once (see ll_validate_range); the loads below are then plain
        for (int hv = 0; hv < GROUP + 2; ++hv) {
            const int off = (hv < GROUP) ? q_off + hv * D : (hv == GROUP ? k_off : v_off);
            ll_validate_range(row_ll, 0, 1, off >> 1, (off + D) >> 1, xtag, tid, kAttnThreads, -1);
        }
        sync();
    }
    // The rows are READ only now: the newest of them was written by this layer's launch of the previous step, and a
    // chain of programmatic launches (or a flag chain) may let this kernel's prologue run before that one has finished.
    if (bulk && tid == 0) {
#pragma unroll
        for (int st = 0; st < kAttnStages; ++st)
            if (st < n_pass) issue(st, st);
    }
    auto load8 = [&](int off) -> uint4 {       // 8 elements of this token's qkv row starting at element `off`
        return chain ? ll_load8_valid(row_ll + (off >> 1)) : ld_cg_v4(row + off);
    };
    for (int item = tid; item < (GROUP + 1) * TPH; item += kAttnThreads) {
        const int h = item / TPH, c = (item % TPH) * 8;
        const bool is_k = (h == GROUP);
        if (is_k && !(pos >= r_begin && pos < r_end)) continue;
        const int src = is_k ? k_off : q_off + h * D;
        float lo[8], hi[8];
        unpack8(load8(src + c), lo, p.is_bf16);
        unpack8(load8(src + c + D / 2), hi, p.is_bf16);
        if (p.inv_freq) rope_apply8(lo, hi, rs, rc);
        if (is_k) {
            *reinterpret_cast<uint4*>(kbase + (size_t)pos * D + c) = pack8(lo, p.is_bf16);
            *reinterpret_cast<uint4*>(kbase + (size_t)pos * D + c + D / 2) = pack8(hi, p.is_bf16);
            // v is copied verbatim alongside
            const uint4 v_lo = load8(v_off + c), v_hi = load8(v_off + c + D / 2);
            *reinterpret_cast<uint4*>(vbase + (size_t)pos * D + c) = v_lo;
            *reinterpret_cast<uint4*>(vbase + (size_t)pos * D + c + D / 2) = v_hi;
            if (bulk) {
                *reinterpret_cast<uint4*>(tiles->newk + c) = pack8(lo, p.is_bf16);
                *reinterpret_cast<uint4*>(tiles->newk + c + D / 2) = pack8(hi, p.is_bf16);
                *reinterpret_cast<uint4*>(tiles->newv + c) = v_lo;
                *reinterpret_cast<uint4*>(tiles->newv + c + D / 2) = v_hi;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                // round through the storage type like the unfused path (q is stored after RoPE)
                s_q[h][c + i] = h16_to_float(float_to_h16(lo[i], p.is_bf16), p.is_bf16) * p.scale;
                s_q[h][c + D / 2 + i] = h16_to_float(float_to_h16(hi[i], p.is_bf16), p.is_bf16) * p.scale;
            }
        }
    }
    sync();

    // ---- online-softmax pass over this split's rows
    const int sub = lane / LPR;         // which of the warp's RPW rows
    const int dl = (lane % LPR) * 8;    // this lane's 8 dims
    float q[GROUP][8], acc[GROUP][8], m[GROUP], l[GROUP];
#pragma unroll
    for (int h = 0; h < GROUP; ++h) {
        m[h] = -INFINITY;
        l[h] = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) { q[h][i] = s_q[h][dl + i]; acc[h][i] = 0.f; }
    }
    // one row (this lane's 8 dims of it) into the running (m, l, acc) of every q head of the group
    auto accumulate = [&](const uint4& kr, const uint4& vr, bool valid) {
        float kf[8], vf[8];
        unpack8(kr, kf, p.is_bf16);
        unpack8(vr, vf, p.is_bf16);
#pragma unroll
        for (int h = 0; h < GROUP; ++h) {
            float sc = 0.f;
#pragma unroll
            for (int i = 0; i < 8; ++i) sc += q[h][i] * kf[i];
#pragma unroll
            for (int o = LPR / 2; o > 0; o >>= 1) sc += __shfl_xor_sync(0xffffffffu, sc, o);
            if (valid) {
                const float mn = fmaxf(m[h], sc);
                const float corr = __expf(m[h] - mn);
                const float pr = __expf(sc - mn);
                l[h] = l[h] * corr + pr;
#pragma unroll
                for (int i = 0; i < 8; ++i) acc[h][i] = acc[h][i] * corr + pr * vf[i];
                m[h] = mn;
            }
        }
    };
    if (bulk) {
        const int old_end = r_begin + n_old;
        for (int pi = 0; pi < n_pass; ++pi) {
            const int st = pi % kAttnStages;
            mbar_wait(&tiles->full[st], (uint32_t)((pi / kAttnStages) & 1));
            const int base = r_begin + pi * ROWS;
            uint4 kraw[UNROLL], vraw[UNROLL];
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) {
                const int r = base + warp * RPW * UNROLL + j * RPW + sub;
                if (r < old_end) {
                    kraw[j] = *reinterpret_cast<const uint4*>(tiles->k[st] + (size_t)(r - base) * D + dl);
                    vraw[j] = *reinterpret_cast<const uint4*>(tiles->v[st] + (size_t)(r - base) * D + dl);
                } else {
                    kraw[j] = make_uint4(0, 0, 0, 0);
                    vraw[j] = make_uint4(0, 0, 0, 0);
                }
            }
#pragma unroll
            for (int j = 0; j < UNROLL; ++j)
                accumulate(kraw[j], vraw[j], base + warp * RPW * UNROLL + j * RPW + sub < old_end);
            if (pi + kAttnStages < n_pass) {
                sync();                                   // every thread is done with stage st
                if (tid == 0) issue(pi + kAttnStages, st);
            }
        }
        if (pos >= r_begin && pos < r_end && warp == 0) {  // the new token's own row, from shared memory
            const uint4 kr = *reinterpret_cast<const uint4*>(tiles->newk + dl);
            const uint4 vr = *reinterpret_cast<const uint4*>(tiles->newv + dl);
            accumulate(kr, vr, sub == 0);
        }
    } else {
        for (int r0 = r_begin + warp * RPW * UNROLL; r0 < r_end; r0 += NW * RPW * UNROLL) {
            uint4 kraw[UNROLL], vraw[UNROLL];
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) {
                const int r = r0 + j * RPW + sub;
                if (r < r_end) {
                    kraw[j] = *reinterpret_cast<const uint4*>(kbase + (size_t)r * D + dl);
                    vraw[j] = *reinterpret_cast<const uint4*>(vbase + (size_t)r * D + dl);
                } else {
                    kraw[j] = make_uint4(0, 0, 0, 0);
                    vraw[j] = make_uint4(0, 0, 0, 0);
                }
            }
#pragma unroll
            for (int j = 0; j < UNROLL; ++j) accumulate(kraw[j], vraw[j], r0 + j * RPW + sub < r_end);
        }
    }
    // ---- merge the RPW row-groups of a warp, then the warps, then the splits
#pragma unroll
    for (int h = 0; h < GROUP; ++h) {
#pragma unroll
        for (int o = LPR; o < 32; o <<= 1) {
            const float m2 = __shfl_xor_sync(0xffffffffu, m[h], o);
            const float l2 = __shfl_xor_sync(0xffffffffu, l[h], o);
            const float mn = fmaxf(m[h], m2);
            const float c1 = (m[h] == -INFINITY) ? 0.f : __expf(m[h] - mn);
            const float c2 = (m2 == -INFINITY) ? 0.f : __expf(m2 - mn);
            l[h] = l[h] * c1 + l2 * c2;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float a2 = __shfl_xor_sync(0xffffffffu, acc[h][i], o);
                acc[h][i] = acc[h][i] * c1 + a2 * c2;
            }
            m[h] = mn;
        }
        if (lane < LPR) {
            if (lane == 0) { s_m[warp][h] = m[h]; s_l[warp][h] = l[h]; }
#pragma unroll
            for (int i = 0; i < 8; ++i) s_o[warp][h][dl + i] = acc[h][i];
        }
    }
    sync();
    // thread -> (head, dim) pairs
    float* my_part = p.scratch + ((size_t)(t * p.n_heads + g * GROUP) * splits + s) * (D + 2);
    for (int idx = tid; idx < GROUP * D; idx += kAttnThreads) {
        const int h = idx / D, d = idx % D;
        float mm = -INFINITY;
#pragma unroll
        for (int w = 0; w < NW; ++w) mm = fmaxf(mm, s_m[w][h]);
        float ll = 0.f, oo = 0.f;
#pragma unroll
        for (int w = 0; w < NW; ++w) {
            const float c = (s_m[w][h] == -INFINITY) ? 0.f : __expf(s_m[w][h] - mm);
            ll += s_l[w][h] * c;
            oo += s_o[w][h][d] * c;
        }
        if (eff == 1) {
            const uint16_t hv = float_to_h16(oo / ll, p.is_bf16);
            p.out[(size_t)t * p.ldo + (g * GROUP + h) * D + d] = hv;
            if (chain) ll_store_pairwise(p.ch.out + (size_t)t * p.ch.ldo, (g * GROUP + h) * D + d, hv, true, otag);
        } else if (p.ll) {
            // ---- ticket-free merge: splits 0..eff-2 publish (value, tag) words, split eff-1 polls for them, merges
            // in split order (same arithmetic as the ticket path) and clears the words again.
            uint2* wbase = reinterpret_cast<uint2*>(p.ll) + ((size_t)(t * p.n_heads + g * GROUP + h) * splits) * (D + 2);
            if (s != eff - 1) {
                uint2* dst = wbase + (size_t)s * (D + 2);
                ll_word_store(dst + d, oo);
                if (d == 0) { ll_word_store(dst + D, mm); ll_word_store(dst + D + 1, ll); }
            } else {
                float pm[16], pl[16], po[16];   // splits <= 16
                for (int j0 = 0; j0 < eff - 1; j0 += 4) {
                    const int nb = min(4, eff - 1 - j0);
                    uint32_t v[12], tg[12];
                    long long t0 = 0;
                    for (;;) {
                        bool ok = true;
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            if (i < nb) {
                                const uint2* src = wbase + (size_t)(j0 + i) * (D + 2);
                                ll_word_load(src + d, v[3 * i], tg[3 * i]);
                                ll_word_load(src + D, v[3 * i + 1], tg[3 * i + 1]);
                                ll_word_load(src + D + 1, v[3 * i + 2], tg[3 * i + 2]);
                            }
                        }
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            if (i < nb) ok = ok && tg[3 * i] == 1u && tg[3 * i + 1] == 1u && tg[3 * i + 2] == 1u;
                        if (ok) break;
                        if (t0 == 0) t0 = clock64();
                        else if (clock64() - t0 > 8000000000LL) {
                            printf("b200: split-KV merge timed out (block %d thread %d)\n", blockIdx.x, threadIdx.x);
                            __trap();
                        }
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        if (i < nb) {
                            po[j0 + i] = __uint_as_float(v[3 * i]);
                            pm[j0 + i] = __uint_as_float(v[3 * i + 1]);
                            pl[j0 + i] = __uint_as_float(v[3 * i + 2]);
                            ll_word_clear(wbase + (size_t)(j0 + i) * (D + 2) + d);
                        }
                    }
                }
                float gm = mm;
                for (int j = 0; j < eff - 1; ++j) gm = fmaxf(gm, pm[j]);
                float gl = 0.f, go = 0.f;
                for (int j = 0; j < eff - 1; ++j) {
                    const float cj = __expf(pm[j] - gm);
                    gl += pl[j] * cj;
                    go += po[j] * cj;
                }
                const float cs = __expf(mm - gm);
                gl += ll * cs;
                go += oo * cs;
                const uint16_t hv = float_to_h16(go / gl, p.is_bf16);
                p.out[(size_t)t * p.ldo + (g * GROUP + h) * D + d] = hv;
                if (chain) ll_store_pairwise(p.ch.out + (size_t)t * p.ch.ldo, (g * GROUP + h) * D + d, hv, true, otag);
            }
        } else {
            float* dst = my_part + (size_t)h * splits * (D + 2);
            dst[d] = oo;
            if (d == 0) { dst[D] = mm; dst[D + 1] = ll; }
        }
    }
    if (eff == 1) return;
    if (p.ll) {
        if (s != eff - 1) return;
        sync();   // every thread of the merging split has read the shared (m, l) words: clear them
        for (int idx = tid; idx < GROUP * (eff - 1); idx += kAttnThreads) {
            const int h = idx / (eff - 1), j = idx % (eff - 1);
            uint2* w = reinterpret_cast<uint2*>(p.ll) + ((size_t)(t * p.n_heads + g * GROUP + h) * splits + j) * (D + 2);
            ll_word_clear(w + D);
            ll_word_clear(w + D + 1);
        }
        return;
    }
    sync();   // orders every thread's partial stores before thread 0's release below
    if (tid == 0) {
        int old;
        asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], %2;\n"
                     : "=r"(old) : "l"(&p.counters[t * p.n_kv + g]), "r"(1) : "memory");
        s_last = (old + 1 == eff) ? 1 : 0;
    }
    sync();
    if (!s_last) return;
    for (int idx = tid; idx < GROUP * D; idx += kAttnThreads) {
        const int h = idx / D, d = idx % D;
        const float* src = p.scratch + ((size_t)(t * p.n_heads + g * GROUP + h) * splits) * (D + 2);
        float mm = -INFINITY;
        for (int j = 0; j < eff; ++j) mm = fmaxf(mm, __ldcg(src + (size_t)j * (D + 2) + D));
        float ll = 0.f, oo = 0.f;
        for (int j = 0; j < eff; ++j) {
            const float c = __expf(__ldcg(src + (size_t)j * (D + 2) + D) - mm);
            ll += __ldcg(src + (size_t)j * (D + 2) + D + 1) * c;
            oo += __ldcg(src + (size_t)j * (D + 2) + d) * c;
        }
        p.out[(size_t)t * p.ldo + (g * GROUP + h) * D + d] = float_to_h16(oo / ll, p.is_bf16);
    }
    if (tid == 0) p.counters[t * p.n_kv + g] = 0;
}


template <int D, int GROUP, int BULK = 0>
__global__ void __launch_bounds__(kAttnThreads) attn_decode_kernel(const AttnDecodeParams p) {
    __shared__ AttnDecodeSmem<D, GROUP> sm;
    extern __shared__ __align__(128) unsigned char attn_dyn_smem[];      // BULK: the K/V tiles
    AttnBulkTiles<D>* tiles = BULK ? reinterpret_cast<AttnBulkTiles<D>*>(attn_dyn_smem) : nullptr;
    int bid = blockIdx.x;
    const int s = bid % p.splits; bid /= p.splits;
    const int g = bid % p.n_kv;
    const int t = bid / p.n_kv;
    pdl_launch_dependents();
    if (threadIdx.x == 0) trace_event(2, 0);
    const bool chain = p.ch.seq != nullptr;    // a link of a flag chain polls its inputs instead of waiting for the grid
    attn_decode_body<D, GROUP>(p, t, g, s, (int)threadIdx.x, sm, [] { __syncthreads(); },
                               [chain] { if (!chain) pdl_wait_prior_grid(); if (threadIdx.x == 0) trace_event(2, 1); },
                               tiles);
    if (threadIdx.x == 0) trace_event(2, 3);
}

}  // namespace b200
