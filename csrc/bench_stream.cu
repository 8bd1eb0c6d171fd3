// Roofline probe: how fast can CTAs pull HBM through cp.async.bulk rings?
// (No math: one lane issues copies into an S-deep ring and recycles a stage as
// soon as its bytes have landed.)  Used by scripts/exp_stream.py to size the
// GEMM weight ring; not part of the serving path.
#include "common.h"
#include "ptx.cuh"

namespace b200 {

__global__ void __launch_bounds__(64, 1)
stream_probe_kernel(const uint8_t* __restrict__ src, long long total_bytes, int stage_bytes, int stages,
                    int hint, unsigned long long* sink) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)stages * stage_bytes);
    if (threadIdx.x == 0) {
        for (int i = 0; i < stages; ++i) mbar_init(&full[i], 1);
        fence_mbar_init();
    }
    __syncthreads();
    const long long nblk = total_bytes / stage_bytes;
    const long long b0 = (long long)blockIdx.x * nblk / gridDim.x;
    const long long b1 = (long long)(blockIdx.x + 1) * nblk / gridDim.x;
    if (threadIdx.x == 0) {
        const uint64_t policy = l2_policy_evict_first();
        long long issued = b0, done = b0;
        // prime
        for (; issued < b1 && issued - b0 < stages; ++issued) {
            const int s = (int)((issued - b0) % stages);
            mbar_arrive_expect_tx(&full[s], stage_bytes);
            if (hint) bulk_g2s_hint(smem + (size_t)s * stage_bytes, src + issued * stage_bytes, stage_bytes, &full[s], policy);
            else bulk_g2s(smem + (size_t)s * stage_bytes, src + issued * stage_bytes, stage_bytes, &full[s]);
        }
        for (; done < b1; ++done) {
            const long long i = done - b0;
            const int s = (int)(i % stages);
            mbar_wait(&full[s], (uint32_t)((i / stages) & 1));
            if (issued < b1) {
                mbar_arrive_expect_tx(&full[s], stage_bytes);
                if (hint) bulk_g2s_hint(smem + (size_t)s * stage_bytes, src + issued * stage_bytes, stage_bytes, &full[s], policy);
                else bulk_g2s(smem + (size_t)s * stage_bytes, src + issued * stage_bytes, stage_bytes, &full[s]);
                ++issued;
            }
        }
        if (sink && smem[0] == 0x7f && smem[1] == 0x7e) atomicAdd(sink, 1ULL);
    }
}

}  // namespace b200

extern "C" int b200_stream_probe(const void* src, int64_t total_bytes, int32_t stage_bytes, int32_t stages,
                                 int32_t grid, int32_t hint, void* sink, void* stream) {
    using namespace b200;
    const size_t smem = (size_t)stages * stage_bytes + 1024 + 8 * stages + 64;
    if (smem > 227 * 1024) return fail(B200_EINVAL, "stream probe: %zu B of shared memory requested", smem);
    B200_CUDA_CHECK(cudaFuncSetAttribute(stream_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    stream_probe_kernel<<<grid, 64, smem, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const uint8_t*>(src), total_bytes, stage_bytes, stages, hint,
        static_cast<unsigned long long*>(sink));
    B200_CUDA_CHECK(cudaGetLastError());
    return B200_OK;
}
