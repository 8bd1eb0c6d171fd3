// Device-side helpers of the flag-chained decode step (b200_ll_chain in b200_petals.h): activation vectors travel
// between the kernels of a step as 8-byte words {two 16-bit values, tag}; a reader polls until the tag matches.
#pragma once
#include "ptx.cuh"

namespace b200 {

constexpr int kLLOpBits = 12;

struct LLChain {                 // kernel-parameter form of b200_ll_chain
    const int* seq;
    const uint2* x;
    const uint2* res;
    uint2* out;
    int ldx, ldr, ldo;
    int x_op, res_op, out_op, half;
};

__device__ __forceinline__ uint32_t ll_tag(int seq, int op) { return ((uint32_t)seq << kLLOpBits) | (uint32_t)op; }

__device__ __forceinline__ int ll_read_seq(const int* seq) {
    int v;
    asm volatile("ld.volatile.global.s32 %0, [%1];\n" : "=r"(v) : "l"(seq) : "memory");
    return v;
}

__device__ __forceinline__ void ll_timeout(long long& t0, const char* what, uint32_t tag) {
    if (t0 == 0) t0 = clock64();
    else if (clock64() - t0 > 6000000000LL) {
        printf("b200: flag chain timed out in %s (block %d thread %d, tag %u)\n", what, blockIdx.x, threadIdx.x, tag);
        __trap();
    }
}

// 8 consecutive 16-bit elements = 4 words (32-byte aligned).
__device__ __forceinline__ uint4 ll_load8(const uint2* w, uint32_t tag) {
    uint32_t a0, t0_, a1, t1_, a2, t2_, a3, t3_;
    long long t0 = 0;
    for (;;) {
        asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];\n"
                     : "=r"(a0), "=r"(t0_), "=r"(a1), "=r"(t1_) : "l"(w) : "memory");
        asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];\n"
                     : "=r"(a2), "=r"(t2_), "=r"(a3), "=r"(t3_) : "l"(w + 2) : "memory");
        if (t0_ == tag && t1_ == tag && t2_ == tag && t3_ == tag) break;
        __nanosleep(64);    // back off: thousands of threads polling flat out starve the producers and the weight stream
        ll_timeout(t0, "ll_load8", tag);
    }
    return make_uint4(a0, a1, a2, a3);
}

// Validation of a range of words before they are read WITHOUT checks: every word of [w_begin, w_end) (both multiples
// of 2) of T rows is polled by exactly one of the `nthreads` callers - four 16-byte loads in flight per round, a back-off
// between rounds - until it carries `tag`; then the callers meet at named barrier `bar` (bar < 0: the caller
// synchronises).  A producer writes each word once per step, so a word that carries the tag is final and the loads that
// follow can be plain, pipelined loads (ll_load8_valid) instead of one blocking poll per load - which would put an L2
// round trip on the critical path of every k-block - and the L2 is not hammered by thousands of spinning threads.
__device__ __forceinline__ void ll_validate_range(const uint2* x, int ld, int T, int w_begin, int w_end, uint32_t tag,
                                                  int tid, int nthreads, int bar) {
    const int pairs = (w_end - w_begin) >> 1;          // 16-byte cells per row
    const int total = pairs * T;
    for (int i0 = tid; i0 < total; i0 += 4 * nthreads) {
        long long t0 = 0;
        for (;;) {
            uint32_t tg[8];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int i = i0 + k * nthreads;
                tg[2 * k] = tag; tg[2 * k + 1] = tag;
                if (i < total) {
                    const int t = i / pairs, c = i - t * pairs;
                    uint32_t a, b;
                    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];\n"
                                 : "=r"(a), "=r"(tg[2 * k]), "=r"(b), "=r"(tg[2 * k + 1])
                                 : "l"(x + (size_t)t * ld + w_begin + 2 * c) : "memory");
                }
            }
            bool ok = true;
#pragma unroll
            for (int k = 0; k < 8; ++k) ok = ok && tg[k] == tag;
            if (ok) break;
            __nanosleep(100);
            ll_timeout(t0, "ll_validate_range", tag);
        }
    }
    if (bar >= 0) named_bar_sync(bar, nthreads);
}

// 8 consecutive elements of a VALIDATED range: two plain 16-byte loads, no poll.
__device__ __forceinline__ uint4 ll_load8_valid(const uint2* w) {
    uint32_t a0, t0_, a1, t1_, a2, t2_, a3, t3_;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];\n" : "=r"(a0), "=r"(t0_), "=r"(a1), "=r"(t1_) : "l"(w));
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];\n" : "=r"(a2), "=r"(t2_), "=r"(a3), "=r"(t3_) : "l"(w + 2));
    return make_uint4(a0, a1, a2, a3);
}

// One word = elements 2i, 2i+1.
__device__ __forceinline__ uint32_t ll_load2(const uint2* w, uint32_t tag) {
    uint32_t a, t_;
    long long t0 = 0;
    for (;;) {
        asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];\n" : "=r"(a), "=r"(t_) : "l"(w) : "memory");
        if (t_ == tag) break;
        __nanosleep(64);
        ll_timeout(t0, "ll_load2", tag);
    }
    return a;
}

__device__ __forceinline__ void ll_store2(uint2* w, uint32_t payload, uint32_t tag) {
    asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %2};\n" ::"l"(w), "r"(payload), "r"(tag) : "memory");
}

// Element n (16-bit) of an LL row.
__device__ __forceinline__ uint16_t ll_load1(const uint2* row, int n, uint32_t tag) {
    const uint32_t w = ll_load2(row + (n >> 1), tag);
    return (uint16_t)((n & 1) ? (w >> 16) : (w & 0xffff));
}

// The lanes of a warp hold consecutive elements (lane i -> element base + i, base even): even lanes store the
// word of their pair.  Every lane of the warp must call (full-mask shuffle); `valid` = this lane's element exists.
__device__ __forceinline__ void ll_store_pairwise(uint2* row, int n, uint16_t v, bool valid, uint32_t tag) {
    const uint32_t mine = valid ? (uint32_t)v : 0u;
    const uint32_t other = __shfl_down_sync(0xffffffffu, mine, 1);
    if (valid && !(n & 1)) ll_store2(row + (n >> 1), mine | (other << 16), tag);
}

}  // namespace b200
