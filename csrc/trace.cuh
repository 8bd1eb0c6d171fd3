// Device-side event trace (opt-in, for scripts/trace_decode.py): selected kernels stamp %globaltimer at entry, after
// their programmatic-dependent-launch wait, when their first accumulator is ready and at exit, for the first and the
// last CTA of the grid.
// Every translation unit that stamps owns a copy of the pointer; b200_trace_set() sets all of them.
#pragma once
#include <cuda_runtime.h>

namespace b200 {

static __device__ unsigned long long* t_trace_buf = nullptr;   // [0] = events written, then (time, tag) pairs
static __device__ int t_trace_cap = 0;

// Compiled in only with -DB200_TRACE (python csrc/build.py --trace -> ops/libb200petals_trace.so): even a disabled
// stamp is a dependent global load at the head of every CTA, which cost the decode step 2.5 % when it was always on.
__device__ __forceinline__ void trace_event(int kernel_id, int phase, int aux = 0) {
#ifndef B200_TRACE
    (void)kernel_id; (void)phase; (void)aux;
    return;
#endif
    unsigned long long* b = t_trace_buf;
    if (b == nullptr) return;
    const bool edge = blockIdx.x == 0 || blockIdx.x == gridDim.x - 1;
    if (!edge && (kernel_id == 3 || (phase != 3 && phase != 0))) return;   // per-launch kernels: entries and exits of every CTA
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t));
    const unsigned long long slot = atomicAdd(b, 1ULL);
    if (slot < (unsigned long long)t_trace_cap) {
        b[1 + 2 * slot] = t;
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;\n" : "=r"(smid));
        b[2 + 2 * slot] = ((unsigned long long)kernel_id << 56) | ((unsigned long long)(aux & 0xffffff) << 32) |
                          ((unsigned long long)phase << 16) | ((unsigned long long)(smid & 0xfff) << 2) | (blockIdx.x == 0 ? 0ULL : (blockIdx.x == gridDim.x - 1 ? 1ULL : 2ULL));
    }
}

#define B200_DEFINE_TRACE_SETTER(name)                                                          \
    extern "C" int name(void* buf, int cap) {                                                   \
        unsigned long long* p = static_cast<unsigned long long*>(buf);                          \
        if (cudaMemcpyToSymbol(::b200::t_trace_buf, &p, sizeof(p)) != cudaSuccess) return -2;   \
        if (cudaMemcpyToSymbol(::b200::t_trace_cap, &cap, sizeof(cap)) != cudaSuccess) return -2; \
        return 0;                                                                               \
    }

}  // namespace b200
