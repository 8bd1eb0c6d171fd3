// Tensor-parallel all-reduce of a row-parallel linear's output through the NVSwitch (NVLS), with the linear's
// epilogue in the same kernel:   out = sum over ranks(partial_r) + bias + residual   on every rank.
//
// Every rank's GEMM leaves its partial [T, N] (16-bit) in a symmetric buffer that is also mapped as ONE multicast
// object.  Rank r owns a contiguous slice of the T*N elements: it reads the slice through the multicast address with
// multimem.ld_reduce - the switch adds the R ranks' values in flight, in fp32 - adds bias and residual, rounds once and
// writes the result back through the multicast address of the output buffer with multimem.st, which the switch
// replicates into every rank's memory.  Each element crosses this GPU's links once in and once out instead of R-1
// times each way (reduce-scatter by load, all-gather by store), and no rank polls its peers one after the other.
// Two cross-rank barriers (release/acquire flags in peer memory, numbered by a launch counter) bracket the exchange.
//
// Used for the prefill chunks of a tensor-parallel span (T > 16; decode-size steps use the exchange fused into the
// GEMM, gemm_core.cuh).  The reference's dead TP path calls an NCCL all-reduce here and adds bias / residual in
// separate torch ops (petals/server/backend.py:134-136 via tensor_parallel).
#include "common.h"
#include "ptx.cuh"

namespace b200 {

struct NvlsParams {
    const uint16_t* mc_partial;   // multicast address of the partial buffer, rows ldp apart
    uint16_t* mc_out;             // multicast address of the output buffer, rows ldo apart
    const uint16_t* residual;     // local, [T, ldr] or null
    const uint16_t* bias;         // [N] or null
    int T, N, ldp, ldo, ldr, is_bf16;
    int world, rank;
    int* flags_local;             // [2][8] arrival numbers, this rank's copy
    int* flags_peer[8];           // the same array on every rank (peer-mapped)
    int* seq;                     // local: launches completed
    int* cta_done;                // local: CTAs past their stores in this launch
};

__device__ __forceinline__ void st_release_sys(int* p, int v) {
    asm volatile("st.release.sys.global.s32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void wait_all(const NvlsParams& p, int phase, int want) {
    if (threadIdx.x < p.world) {
        long long t0 = clock64();
        while (ld_acquire_sys(p.flags_local + phase * 8 + threadIdx.x) - want < 0) {
            if (clock64() - t0 > 20000000000LL) {
                printf("b200: NVLS all-reduce barrier %d timed out (rank %d waits for rank %d, launch %d)\n", phase, p.rank,
                       (int)threadIdx.x, want);
                __trap();
            }
        }
    }
    __syncthreads();
}

template <int BF16>
__global__ void __launch_bounds__(256) tp_allreduce_nvls_kernel(const NvlsParams p) {
    pdl_wait_prior_grid();          // this rank's partial comes from the GEMM before us
    pdl_launch_dependents();
    const int want = *p.seq + 1;
    // ---- barrier 1: every rank's partial is complete
    if (blockIdx.x == 0 && threadIdx.x < p.world) {
        __threadfence_system();
        st_release_sys(p.flags_peer[threadIdx.x] + p.rank, want);
    }
    wait_all(p, 0, want);
    // ---- my slice: 16-byte cells [c0, c1) of the T x N/8 cells
    const int cells_per_row = p.N >> 3;
    const long long cells = (long long)p.T * cells_per_row;
    const long long per = (cells + p.world - 1) / p.world;
    const long long c0 = per * p.rank, c1 = (c0 + per < cells) ? c0 + per : cells;
    for (long long c = c0 + (long long)blockIdx.x * blockDim.x + threadIdx.x; c < c1; c += (long long)gridDim.x * blockDim.x) {
        const int t = (int)(c / cells_per_row), n = (int)(c - (long long)t * cells_per_row) << 3;
        uint32_t v[4];
        const uint16_t* src = p.mc_partial + (size_t)t * p.ldp + n;
        if (BF16)
            asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];\n"
                         : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "l"(src) : "memory");
        else
            asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0, %1, %2, %3}, [%4];\n"
                         : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "l"(src) : "memory");
        uint4 r4 = make_uint4(0, 0, 0, 0), b4 = make_uint4(0, 0, 0, 0);
        if (p.residual) r4 = ld_cg_v4(p.residual + (size_t)t * p.ldr + n);
        if (p.bias) b4 = ld_nc_v4(p.bias + n);
        const uint32_t rw[4] = {r4.x, r4.y, r4.z, r4.w}, bw[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float lo = h16_to_float((uint16_t)(v[k] & 0xffff), BF16), hi = h16_to_float((uint16_t)(v[k] >> 16), BF16);
            if (p.bias) {   // round after the bias like the one-GPU epilogue: the linear's output is a 16-bit tensor
                lo = h16_to_float(float_to_h16(lo + h16_to_float((uint16_t)(bw[k] & 0xffff), BF16), BF16), BF16);
                hi = h16_to_float(float_to_h16(hi + h16_to_float((uint16_t)(bw[k] >> 16), BF16), BF16), BF16);
            }
            if (p.residual) {
                lo += h16_to_float((uint16_t)(rw[k] & 0xffff), BF16);
                hi += h16_to_float((uint16_t)(rw[k] >> 16), BF16);
            }
            v[k] = (uint32_t)float_to_h16(lo, BF16) | ((uint32_t)float_to_h16(hi, BF16) << 16);
        }
        uint16_t* dst = p.mc_out + (size_t)t * p.ldo + n;
        if (BF16)
            asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1, %2, %3, %4};\n" ::"l"(dst), "r"(v[0]), "r"(v[1]),
                         "r"(v[2]), "r"(v[3]) : "memory");
        else
            asm volatile("multimem.st.relaxed.sys.global.v4.f16x2 [%0], {%1, %2, %3, %4};\n" ::"l"(dst), "r"(v[0]), "r"(v[1]),
                         "r"(v[2]), "r"(v[3]) : "memory");
    }
    // ---- barrier 2: every rank has written its slice into everybody's output
    __threadfence_system();
    __syncthreads();
    __shared__ int s_last;
    if (threadIdx.x == 0) s_last = (atomicAdd(p.cta_done, 1) == (int)gridDim.x - 1) ? 1 : 0;
    __syncthreads();
    if (s_last && threadIdx.x < p.world) {
        __threadfence_system();
        st_release_sys(p.flags_peer[threadIdx.x] + 8 + p.rank, want);
    }
    wait_all(p, 1, want);
    if (s_last && threadIdx.x == 0) {     // every CTA of this launch has read *seq and counted itself by now
        *p.cta_done = 0;
        *p.seq = want;
    }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_tp_allreduce_nvls(const b200_nvls_args* a, void* stream_) {
    if (!a || !a->mc_partial || !a->mc_out || !a->flags_local || !a->seq || !a->cta_done)
        return fail(B200_EINVAL, "b200_tp_allreduce_nvls: null pointer argument");
    if (a->world < 2 || a->world > 8 || a->rank < 0 || a->rank >= a->world)
        return fail(B200_EINVAL, "b200_tp_allreduce_nvls: bad group (world %d rank %d)", a->world, a->rank);
    if (a->T <= 0 || a->N <= 0 || (a->N % 8) || (a->ldp % 8) || (a->ldo % 8) || (a->residual && (a->ldr % 8)))
        return fail(B200_EINVAL, "b200_tp_allreduce_nvls: N and the row strides must be multiples of 8 (T=%d N=%d)", a->T, a->N);
    NvlsParams p;
    p.mc_partial = static_cast<const uint16_t*>(a->mc_partial);
    p.mc_out = static_cast<uint16_t*>(a->mc_out);
    p.residual = static_cast<const uint16_t*>(a->residual);
    p.bias = static_cast<const uint16_t*>(a->bias);
    p.T = a->T; p.N = a->N; p.ldp = a->ldp; p.ldo = a->ldo; p.ldr = a->ldr;
    p.is_bf16 = a->dtype == B200_BF16 ? 1 : 0;
    p.world = a->world; p.rank = a->rank;
    p.flags_local = a->flags_local;
    for (int r = 0; r < a->world; ++r) {
        if (!a->flags_peer[r]) return fail(B200_EINVAL, "b200_tp_allreduce_nvls: flags_peer[%d] is null", r);
        p.flags_peer[r] = static_cast<int*>(a->flags_peer[r]);
    }
    p.seq = a->seq; p.cta_done = a->cta_done;
    const long long cells = (long long)a->T * (a->N / 8);
    const long long per = (cells + a->world - 1) / a->world;
    int grid = (int)((per + 255) / 256);
    if (grid > 64) grid = 64;         // every CTA must be resident: they meet at the second barrier
    if (grid < 1) grid = 1;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (p.is_bf16) B200_CUDA_CHECK(launch(tp_allreduce_nvls_kernel<1>, dim3(grid), dim3(256), 0, stream, a->pdl != 0, p));
    else B200_CUDA_CHECK(launch(tp_allreduce_nvls_kernel<0>, dim3(grid), dim3(256), 0, stream, a->pdl != 0, p));
    return B200_OK;
}
