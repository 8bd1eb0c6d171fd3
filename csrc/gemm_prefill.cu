// Prefill-size GEMM (T >= 64): out[T,N] = epi(x[T,K] . W[N,K]^T) on tcgen05 with every operand
// moved by the TMA engine and the epilogue overlapped with the next tile's main loop.
//
//   warp 0      producer: per k-block one cp.async.bulk of the pre-tiled weight block (16/32 KiB)
//               and one cp.async.bulk.tensor.2d of the [BN tokens x 64] activation box (128B swizzle,
//               rows past T are zero-filled by the TMA unit) on the same mbarrier
//   warp 1      MMA issuer: 128 x BN x 16 tcgen05.mma into one of TWO TMEM accumulator buffers
//   warps 2..5  epilogue: drain the other buffer (bias / GELU / SwiGLU / residual) while the tensor
//               core already works on the next tile
// Persistent grid, tiles assigned round-robin; tiles that share a weight block run on neighbouring
// CTAs at the same time, so the weight block is read from HBM once and from L2 afterwards.
// The decode-size kernel (gemm.cu) stages activations with the worker warps instead (fused norm,
// fp8 quantisation) and splits K across CTAs stream-K style.
//
// Short prompts (few token tiles): a Llama-2-7B GEMM then has only 32-96 output tiles for 148 SMs and every CTA
// would stream a whole tile's weights alone (~1-3 MB at the ~90 GB/s one CTA reaches).  Such launches split K:
// every tile is worked on by `ksplit` units (k-ranges), each unit leaves its fp32 partial tile in the scratch
// buffer and takes a ticket; the unit that takes the tile's last ticket adds the partials in k order (bit-reproducible)
// and runs the epilogue.  Nobody waits for anybody.
#include <cuda.h>
#include <stdlib.h>
#include <string.h>

#include "gemm_core.cuh"
#include "tma_host.h"

namespace b200 {

template <int BN, int DUAL>
struct PrefillCfg {
    static constexpr int kAPer = DUAL ? 2 : 1;
    static constexpr int kBytesA = kABytes * kAPer;          // weight bytes per k-block
    static constexpr int kBytesB = BN * 128;                 // activation bytes per k-block
    static constexpr int kStage = kBytesA + kBytesB;
    static constexpr int kStages = (200 * 1024) / kStage;    // 4 for (256, single) and (128, dual)
    static constexpr int kAccCols = BN * kAPer;              // TMEM columns of one accumulator buffer
    static constexpr int kTmemCols = 2 * kAccCols <= 256 ? 256 : 512;
    static constexpr size_t kSmemBytes = (size_t)kStages * kStage + 1024 + 1024 + 4 * 2048 /*epilogue transposes*/;
    static_assert(2 * kAccCols <= 512, "two accumulator buffers must fit TMEM");
};

__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* tmap, int c0, int c1, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];\n" ::
            "r"(smem_u32(smem_dst)),
        "l"(tmap), "r"(c0), "r"(c1), "r"(smem_u32(bar))
        : "memory");
}

template <int BN, int DUAL>
__global__ void __launch_bounds__(kThreads, 1)
gemm_prefill_kernel(const GemmParams p, const __grid_constant__ CUtensorMap tmap_x) {
    using Cfg = PrefillCfg<BN, DUAL>;
    constexpr int S = Cfg::kStages, APER = Cfg::kAPer;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)S * Cfg::kStage);
    uint64_t* full = bars;              // [S]
    uint64_t* empty = bars + S;         // [S]
    uint64_t* acc_full = empty + S;     // [2]
    uint64_t* acc_empty = acc_full + 2; // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
    float* epi_smem = reinterpret_cast<float*>(smem + (size_t)S * Cfg::kStage + 1024);   // [4 warps][16][32]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int i = 0; i < S; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 4); }
        fence_mbar_init();
        asm volatile("prefetch.tensormap [%0];\n" ::"l"(&tmap_x) : "memory");
    }
    if (warp == 1) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();

    const int P = p.ll_parts > 1 ? p.ll_parts : 1;            // k-splits per tile (host: prefill_ksplit)
    const int units = p.n_tiles * p.t_tiles * P;
    volatile int* s_last = reinterpret_cast<volatile int*>(tmem_slot + 1);

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer
        pdl_wait_prior_grid();   // activations come from the previous kernel
        int s = 0;
        uint32_t ph = 0;
        for (int u = blockIdx.x; u < units; u += gridDim.x) {
            const int tu = u / P, part = u - tu * P;
            const int n_tile = tu / p.t_tiles, t_tile = tu - n_tile * p.t_tiles;
            const uint8_t* wt = p.w + (size_t)n_tile * p.kbpt * Cfg::kBytesA;
            const int kb0 = part * p.kbpt / P, kb1 = (part + 1) * p.kbpt / P;
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(&empty[s], ph ^ 1);
                if (elect_one()) {
                    uint8_t* st = smem + (size_t)s * Cfg::kStage;
                    mbar_arrive_expect_tx(&full[s], Cfg::kStage);
                    bulk_g2s(st, wt + (size_t)kb * Cfg::kBytesA, Cfg::kBytesA, &full[s]);
                    tma_load_2d(st + Cfg::kBytesA, &tmap_x, kb * kBK, t_tile * BN, &full[s]);
                }
                __syncwarp();
                if (++s == S) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        const uint32_t idesc = umma_idesc(kTileM, BN, p.is_bf16, p.is_bf16, 0, 0);
        const uint32_t base = smem_u32(smem);
        int s = 0;
        uint32_t ph = 0, it = 0;
        for (int u = blockIdx.x; u < units; u += gridDim.x, ++it) {
            const int buf = it & 1;
            mbar_wait(&acc_empty[buf], ((it >> 1) & 1) ^ 1);
            tc_fence_after();
            const uint32_t d0 = tmem_base + buf * Cfg::kAccCols;
            const int part = u % P;
            const int kb0 = part * p.kbpt / P, kb1 = (part + 1) * p.kbpt / P;
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(&full[s], ph);
                tc_fence_after();
                if (elect_one()) {
                    const uint32_t a_addr = base + s * Cfg::kStage, b_addr = a_addr + Cfg::kBytesA;
                    const uint64_t bdesc = umma_desc_kmajor_sw128(b_addr, 1024);
#pragma unroll
                    for (int a = 0; a < APER; ++a) {
                        const uint64_t adesc = umma_desc_kmajor_sw128(a_addr + a * kABytes, 1024);
#pragma unroll
                        for (int k = 0; k < kBK / 16; ++k)
                            umma_f16(d0 + a * BN, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc,
                                     (kb > kb0 || k > 0) ? 1u : 0u);
                    }
                    umma_commit(&empty[s]);
                }
                __syncwarp();
                if (++s == S) { s = 0; ph ^= 1; }
            }
            if (elect_one()) umma_commit(&acc_full[buf]);
            __syncwarp();
        }
    } else {
        // ------------------------------------------------------------ epilogue warps
        const int lg = warp & 3;
        const int row = lg * 32 + lane;
        const bool vec_ok = !p.out_fp32 && (p.N % 8 == 0) && (p.ldo % 8 == 0) && (!p.residual || p.ldr % 8 == 0);
        float* my_smem = epi_smem + (warp - 2) * 512;
        uint32_t it = 0;
        const int e = threadIdx.x - 64;
        for (int u = blockIdx.x; u < units; u += gridDim.x, ++it) {
            const int tu = u / P;
            const int n_tile = tu / p.t_tiles, t_tile = tu - n_tile * p.t_tiles;
            const int buf = it & 1;
            mbar_wait(&acc_full[buf], (it >> 1) & 1);
            tc_fence_after();
            const int t_live = min(BN, p.T - t_tile * BN);
            const uint32_t taddr0 = tmem_base + buf * Cfg::kAccCols + ((uint32_t)(lg * 32) << 16);
            if (P > 1) {
                // ---- split K: leave the fp32 partial of this unit in the scratch buffer, take a ticket
                constexpr size_t kTileFloats = (size_t)APER * BN * kTileM;
                float* mine = p.scratch + (size_t)u * kTileFloats;
#pragma unroll 1
                for (int c = 0; c * 16 < t_live; ++c) {
                    uint32_t rg[16], ru[16];
                    tmem_ld16(taddr0 + c * 16, rg);
                    if (DUAL) tmem_ld16(taddr0 + BN + c * 16, ru);
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        mine[(size_t)(c * 16 + j) * kTileM + row] = __uint_as_float(rg[j]);
                        if (DUAL) mine[(size_t)(BN + c * 16 + j) * kTileM + row] = __uint_as_float(ru[j]);
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&acc_empty[buf]);     // the tensor core may have the buffer back
                named_bar_sync(1, kEpiThreads);
                if (e == 0) {
                    int old;
                    asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], %2;\n" : "=r"(old) : "l"(&p.counters[tu]), "r"(1) : "memory");
                    *s_last = (old + 1 == P) ? 1 : 0;
                }
                named_bar_sync(1, kEpiThreads);
                if (*s_last) {
                    // last unit of the tile: add the partials in k order and finish the tile
                    const float* base = p.scratch + (size_t)tu * P * kTileFloats;
#pragma unroll 1
                    for (int c = 0; c * 16 < t_live; ++c) {
                        float vg[16], vu[16];
#pragma unroll
                        for (int j = 0; j < 16; ++j) { vg[j] = 0.f; vu[j] = 0.f; }
                        for (int q = 0; q < P; ++q) {
                            const float* src = base + (size_t)q * kTileFloats;
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                vg[j] += __ldcg(src + (size_t)(c * 16 + j) * kTileM + row);
                                if (DUAL) vu[j] += __ldcg(src + (size_t)(BN + c * 16 + j) * kTileM + row);
                            }
                        }
                        if (vec_ok) epilogue_store_chunk_vec<BN, DUAL>(p, n_tile, t_tile, lane, lg, c, vg, vu, my_smem);
                        else epilogue_store_chunk<BN, DUAL>(p, n_tile, t_tile, row, c, vg, vu, nullptr);
                    }
                    if (e == 0) p.counters[tu] = 0;      // clean for the next launch
                }
                named_bar_sync(1, kEpiThreads);           // s_last is rewritten by the next unit
                continue;
            }
#pragma unroll 1
            for (int c = 0; c * 16 < t_live; ++c) {
                uint32_t rg[16], ru[16];
                tmem_ld16(taddr0 + c * 16, rg);
                if (DUAL) tmem_ld16(taddr0 + BN + c * 16, ru);
                tmem_ld_wait();
                float vg[16], vu[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    vg[j] = __uint_as_float(rg[j]);
                    vu[j] = DUAL ? __uint_as_float(ru[j]) : 0.f;
                }
                if (vec_ok) epilogue_store_chunk_vec<BN, DUAL>(p, n_tile, t_tile, lane, lg, c, vg, vu, my_smem);
                else epilogue_store_chunk<BN, DUAL>(p, n_tile, t_tile, row, c, vg, vu, nullptr);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[buf]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
}

template <int BN, int DUAL>
static int launch_prefill(const GemmParams& p, const CUtensorMap& tmap, int grid, bool pdl, cudaStream_t stream) {
    using Cfg = PrefillCfg<BN, DUAL>;
    static_assert(Cfg::kSmemBytes <= 227 * 1024, "shared memory budget exceeded");
    static thread_local int configured_dev = -1;
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    if (configured_dev != dev) {
        B200_CUDA_CHECK(cudaFuncSetAttribute(gemm_prefill_kernel<BN, DUAL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)Cfg::kSmemBytes));
        configured_dev = dev;
    }
    B200_CUDA_CHECK(launch(gemm_prefill_kernel<BN, DUAL>, dim3(grid), dim3(kThreads), Cfg::kSmemBytes, stream, pdl, p, tmap));
    return B200_OK;
}

}  // namespace b200

using namespace b200;

// Same argument struct as b200_gemm; valid for T >= 64, no norm prologue, 16-bit weights.
extern "C" int b200_gemm_prefill(const b200_gemm_args* a, void* stream_) {
    if (!a || !a->w_packed || !a->x || !a->out) return fail(B200_EINVAL, "b200_gemm_prefill: null pointer argument");
    if (a->T <= 0 || a->N <= 0 || a->K <= 0 || (a->K % kBK) != 0)
        return fail(B200_EINVAL, "b200_gemm_prefill: bad sizes T=%d N=%d K=%d", a->T, a->N, a->K);
    if (a->norm_mode != B200_NORM_NONE || a->fp8)
        return fail(B200_EINVAL, "b200_gemm_prefill: norm prologue / fp8 weights belong to b200_gemm");
    if (a->n_tiles * kTileM < a->N) return fail(B200_EINVAL, "b200_gemm_prefill: n_tiles does not cover N");
    if ((a->ldx % 8) != 0 || (reinterpret_cast<uintptr_t>(a->x) & 15) != 0)
        return fail(B200_EINVAL, "b200_gemm_prefill: x must be 16-byte aligned with ldx %% 8 == 0");
    if (a->dual && a->epilogue != B200_EPI_SWIGLU) return fail(B200_EINVAL, "b200_gemm_prefill: dual weights need SwiGLU");
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return fail(B200_ECUDA, "b200_gemm_prefill: cuTensorMapEncodeTiled is not available from the driver");
    // 128 x 256 tiles (128 x 128 x 2 matrices for SwiGLU); inputs of at most 128 tokens use the 128-wide
    // tile so half of every MMA is not spent on zero-filled rows
    const int bn = (a->dual || a->T <= 128) ? 128 : 256;
    GemmParams p;
    memset(&p, 0, sizeof(p));
    p.w = static_cast<const uint8_t*>(a->w_packed);
    p.x = static_cast<const uint16_t*>(a->x);
    p.out = a->out;
    p.residual = static_cast<const uint16_t*>(a->residual);
    p.bias = static_cast<const uint16_t*>(a->bias);
    p.T = a->T; p.N = a->N; p.K = a->K;
    p.ldx = a->ldx; p.ldo = a->ldo; p.ldr = a->ldr;
    p.n_tiles = a->n_tiles; p.t_tiles = (a->T + bn - 1) / bn; p.kbpt = a->K / kBK;
    p.epilogue = a->epilogue;
    p.is_bf16 = (a->dtype == B200_BF16) ? 1 : 0;
    p.out_fp32 = a->out_fp32;

    CUtensorMap tmap;
    const cuuint64_t dims[2] = {(cuuint64_t)a->K, (cuuint64_t)a->T};
    const cuuint64_t strides[1] = {(cuuint64_t)a->ldx * 2};
    const cuuint32_t box[2] = {(cuuint32_t)kBK, (cuuint32_t)bn};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(&tmap, p.is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                           const_cast<void*>(a->x), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(B200_ECUDA, "b200_gemm_prefill: cuTensorMapEncodeTiled failed (%d)", (int)r);

    int grid = a->grid > 0 ? a->grid : sm_count_cached();
    if (grid > sm_count_cached()) grid = sm_count_cached();
    // Few tiles for the SMs (short prompts): split K over up to 4 units per tile when that fills the grid better
    // (see the kernel comment); needs the caller's scratch buffer and zeroed counters like b200_gemm.
    const int tiles = p.n_tiles * p.t_tiles;
    int ksplit = 1;
    static const int split_on = [] { const char* e = getenv("B200_PREFILL_KSPLIT"); return e ? atoi(e) : 1; }();
    if (split_on && a->scratch && a->counters && tiles < grid) {
        // One wave only: a short-prompt GEMM is a weight stream, ~90 GB/s per CTA, so what a split buys is more SMs
        // streaming at once.  From ~75 streaming CTAs the HBM is the limit and a second wave of units only adds
        // their fixed costs (measured: 3-way split of the 96-tile qkv GEMM made a 128-token prefill slower).
        for (int q = 2; q <= 4; ++q) {
            if (p.kbpt / q < 8) break;
            const int un = tiles * q;
            const size_t need = (size_t)un * (a->dual ? 2 : 1) * bn * kTileM * sizeof(float);
            if (tiles * (q - 1) < 75 && un <= grid && need <= (size_t)b200_gemm_scratch_bytes(a->T)) ksplit = q;
        }
    }
    p.ll_parts = ksplit;
    p.scratch = a->scratch;
    p.counters = a->counters;
    const int units = tiles * ksplit;
    if (grid > units) grid = units;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (a->dual) return launch_prefill<128, 1>(p, tmap, grid, a->pdl != 0, stream);
    if (bn == 128) return launch_prefill<128, 0>(p, tmap, grid, a->pdl != 0, stream);
    return launch_prefill<256, 0>(p, tmap, grid, a->pdl != 0, stream);
}
