// Causal prefill attention on the 5th-gen tensor cores (tcgen05 + TMEM).
//
// One CTA = 128 queries of one q head.  For every block of 128 cached keys:
//     S = Q K^T          tcgen05.mma, A = Q tile, B = K tile (both K-major, 128-byte swizzle), S in TMEM
//     online softmax     128 threads, thread = query row = TMEM lane; P (16-bit) -> shared memory
//     O_blk = P V        tcgen05.mma, A = P (K-major over keys), B = V tile as an MN-major operand
//                        (V is stored [key, d]: no transposed copy of the cache is needed)
//     O = O * corr + O_blk in registers
// The T x S score matrix never exists in memory (the reference materialises [nh, T, S] in fp32:
// petals/llama/block.py:134-141) and the causal mask is applied (the reference's live path has none).
// v1 is not software-pipelined: loads, MMAs and softmax of a block run back to back.
#include "attn_core.cuh"

namespace b200 {

struct AttnPrefillTcParams {
    const uint16_t* __restrict__ q;     // rotated q inside the qkv buffer
    uint16_t* __restrict__ out;
    const uint16_t* __restrict__ kcache;
    const uint16_t* __restrict__ vcache;
    int T, n_heads, n_kv, s_max, ldq, ldo, pos0, is_bf16;
    float scale_log2e;
};

constexpr int kPfThreads = 192;   // warp 0: spare, warp 1: MMA issuer, warps 2-5: load / softmax / accumulate
constexpr int kPfBQ = 128, kPfBK = 128;

template <int D>
__global__ void __launch_bounds__(kPfThreads, 1) attn_prefill_tc_kernel(const AttnPrefillTcParams p) {
    constexpr int HALVES = D / 64;                 // 64-element (128-byte) column groups of a [rows x D] tile
    constexpr int TILE_BYTES = 128 * 128 * HALVES; // Q, K, V tiles
    constexpr int P_BYTES = 128 * 128 * 2;         // P: 128 queries x 128 keys
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* s_q = smem;
    uint8_t* s_k = s_q + TILE_BYTES;
    uint8_t* s_v = s_k + TILE_BYTES;
    uint8_t* s_p = s_v + TILE_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_p + P_BYTES);
    uint64_t* kv_ready = bars;      // 128 arrivals: K/V (and at first Q) tiles are in shared memory
    uint64_t* s_full = bars + 1;    // S is in TMEM
    uint64_t* p_ready = bars + 2;   // 128 arrivals: P is in shared memory
    uint64_t* o_full = bars + 3;    // O_blk is in TMEM
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int h = blockIdx.y, q0 = blockIdx.x * kPfBQ;
    const int g = h / (p.n_heads / p.n_kv);

    if (threadIdx.x == 0) {
        mbar_init(kv_ready, 128);
        mbar_init(s_full, 1);
        mbar_init(p_ready, 128);
        mbar_init(o_full, 1);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<256>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tmem_s = tmem_base, tmem_o = tmem_base + 128;

    pdl_wait_prior_grid();
    pdl_launch_dependents();

    const int last_q = min(p.T, q0 + kPfBQ) - 1;
    const int kv_end = p.pos0 + last_q + 1;
    const int n_blocks = (kv_end + kPfBK - 1) / kPfBK;

    if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        const uint32_t idesc_s = umma_idesc(128, kPfBK, p.is_bf16, p.is_bf16, 0, 0);
        const uint32_t idesc_o = umma_idesc(128, D, p.is_bf16, p.is_bf16, 0, 1);   // B = V is MN-major
        const uint32_t q_addr = smem_u32(s_q), k_addr = smem_u32(s_k), v_addr = smem_u32(s_v), p_addr = smem_u32(s_p);
        uint32_t ph = 0;
        for (int j = 0; j < n_blocks; ++j, ph ^= 1) {
            mbar_wait(kv_ready, ph);
            tc_fence_after();
            if (elect_one()) {
#pragma unroll
                for (int ks = 0; ks < D / 16; ++ks) {
                    const uint32_t off = (ks / 4) * (128 * 128) + (ks % 4) * 32;
                    umma_f16(tmem_s, umma_desc_kmajor_sw128(q_addr + off, 1024), umma_desc_kmajor_sw128(k_addr + off, 1024),
                             idesc_s, ks > 0 ? 1u : 0u);
                }
                umma_commit(s_full);
            }
            __syncwarp();
            mbar_wait(p_ready, ph);
            tc_fence_after();
            if (elect_one()) {
#pragma unroll
                for (int ks = 0; ks < kPfBK / 16; ++ks) {
                    const uint32_t a_off = (ks / 4) * (128 * 128) + (ks % 4) * 32;     // P: K-major over keys
                    const uint32_t b_off = ks * 2048;                                   // V: 16 key rows = 2 swizzle atoms
                    umma_f16(tmem_o, umma_desc_kmajor_sw128(p_addr + a_off, 1024),
                             umma_desc_mnmajor_sw128(v_addr + b_off, 128 * 128, 1024), idesc_o, ks > 0 ? 1u : 0u);
                }
                umma_commit(o_full);
            }
            __syncwarp();
        }
    } else if (warp >= 2) {
        // ------------------------------------------------------------ load / softmax / accumulate
        const int e = threadIdx.x - 64;             // 0..127
        const int lg = warp & 3;
        const int row = lg * 32 + lane;             // query row == TMEM lane
        const int t_row = q0 + row;
        const int q_pos = p.pos0 + t_row;
        const uint16_t* kb = p.kcache + (size_t)g * p.s_max * D;
        const uint16_t* vb = p.vcache + (size_t)g * p.s_max * D;

        // tile loader: [128 rows x D] 16-bit, 16-byte chunks, 128-byte swizzle, column halves 16 KiB apart
        auto load_tile = [&](uint8_t* dst, const uint16_t* src, int ld, int row0, int row_limit) {
            constexpr int CPR = D / 8;   // 16-byte chunks per row
            const uint32_t base = smem_u32(dst);
            for (int i = e; i < 128 * CPR; i += 128) {
                const int r = i / CPR, c = i % CPR;
                uint4 v = make_uint4(0, 0, 0, 0);
                if (row0 + r < row_limit) v = ld_cg_v4(src + (size_t)(row0 + r) * ld + c * 8);
                st_shared_v4(base + (c / 8) * (128 * 128) + r * 128 + (((c & 7) ^ (r & 7)) << 4), v);
            }
        };
        load_tile(s_q, p.q + (size_t)h * D, p.ldq, q0, p.T);

        float acc[D];
#pragma unroll
        for (int i = 0; i < D; ++i) acc[i] = 0.f;
        float m_run = -INFINITY, l_run = 0.f;
        uint32_t ph = 0;
        const uint32_t t_lane = (uint32_t)(lg * 32) << 16;
        for (int j = 0; j < n_blocks; ++j, ph ^= 1) {
            const int k0 = j * kPfBK;
            load_tile(s_k, kb, D, k0, kv_end);
            load_tile(s_v, vb, D, k0, kv_end);
            fence_proxy_async_smem();
            mbar_arrive(kv_ready);

            mbar_wait(s_full, ph);
            tc_fence_after();
            // ---- pass 1: row maximum of this block (masked)
            float m_blk = -INFINITY;
#pragma unroll 1
            for (int c = 0; c < kPfBK / 16; ++c) {
                uint32_t r[16];
                tmem_ld16(tmem_s + t_lane + c * 16, r);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int key = k0 + c * 16 + i;
                    const float s = (key <= q_pos && key < kv_end) ? __uint_as_float(r[i]) : -INFINITY;
                    m_blk = fmaxf(m_blk, s);
                }
            }
            const float m_new = fmaxf(m_run, m_blk);
            const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
            const float corr = (m_run == -INFINITY) ? 0.f : exp2f((m_run - m_use) * p.scale_log2e);
            // ---- pass 2: probabilities -> shared memory (16-bit, K-major over keys), row sum
            float l_blk = 0.f;
            const uint32_t p_base = smem_u32(s_p);
#pragma unroll 1
            for (int c = 0; c < kPfBK / 16; ++c) {
                uint32_t r[16];
                tmem_ld16(tmem_s + t_lane + c * 16, r);
                tmem_ld_wait();
                uint32_t packed[8];
#pragma unroll
                for (int i = 0; i < 16; i += 2) {
                    float pr[2];
#pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        const int key = k0 + c * 16 + i + u;
                        const bool ok = (key <= q_pos && key < kv_end);
                        pr[u] = ok ? exp2f((__uint_as_float(r[i + u]) - m_use) * p.scale_log2e) : 0.f;
                    }
                    const uint16_t a = float_to_h16(pr[0], p.is_bf16), b = float_to_h16(pr[1], p.is_bf16);
                    // the sum uses the rounded probabilities, i.e. exactly what the P.V product sees
                    l_blk += h16_to_float(a, p.is_bf16) + h16_to_float(b, p.is_bf16);
                    packed[i / 2] = (uint32_t)a | ((uint32_t)b << 16);
                }
                // 16 keys = two 16-byte chunks: chunk index (c*2, c*2+1) of the 128-key row
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    const int chunk = c * 2 + hh;
                    st_shared_v4(p_base + (chunk / 8) * (128 * 128) + row * 128 + (((chunk & 7) ^ (row & 7)) << 4),
                                 make_uint4(packed[hh * 4], packed[hh * 4 + 1], packed[hh * 4 + 2], packed[hh * 4 + 3]));
                }
            }
            tc_fence_before();
            fence_proxy_async_smem();
            mbar_arrive(p_ready);

            l_run = l_run * corr + l_blk;
            m_run = m_new;
            mbar_wait(o_full, ph);
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < D / 16; ++c) {
                uint32_t r[16];
                tmem_ld16(tmem_o + t_lane + c * 16, r);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 16; ++i) acc[c * 16 + i] = acc[c * 16 + i] * corr + __uint_as_float(r[i]);
            }
            tc_fence_before();
        }
        if (t_row < p.T) {
            const float inv = 1.0f / l_run;
            uint16_t* dst = p.out + (size_t)t_row * p.ldo + (size_t)h * D;
#pragma unroll
            for (int c = 0; c < D / 8; ++c) {
                float f[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) f[i] = acc[c * 8 + i] * inv;
                *reinterpret_cast<uint4*>(dst + c * 8) = pack8(f, p.is_bf16);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<256>(tmem_base);
}

}  // namespace b200

using namespace b200;

extern "C" int b200_attn_prefill_tc(const void* q, void* out, const void* kcache, const void* vcache, int32_t T,
                                    int32_t n_heads, int32_t n_kv_heads, int32_t head_dim, int32_t s_max, int32_t ldq,
                                    int32_t ldo, int32_t pos0, float scale, int32_t dtype, int32_t pdl, void* stream_) {
    if (!q || !out || !kcache || !vcache) return fail(B200_EINVAL, "b200_attn_prefill_tc: null pointer argument");
    if (head_dim != 128 && head_dim != 64) return fail(B200_EINVAL, "b200_attn_prefill_tc: unsupported head_dim=%d", head_dim);
    if (n_kv_heads <= 0 || n_heads % n_kv_heads) return fail(B200_EINVAL, "b200_attn_prefill_tc: bad head counts");
    if ((ldq % 8) || (ldo % 8)) return fail(B200_EINVAL, "b200_attn_prefill_tc: row strides must be multiples of 8 elements");
    AttnPrefillTcParams p;
    p.q = static_cast<const uint16_t*>(q);
    p.out = static_cast<uint16_t*>(out);
    p.kcache = static_cast<const uint16_t*>(kcache);
    p.vcache = static_cast<const uint16_t*>(vcache);
    p.T = T; p.n_heads = n_heads; p.n_kv = n_kv_heads; p.s_max = s_max; p.ldq = ldq; p.ldo = ldo; p.pos0 = pos0;
    p.is_bf16 = (dtype == B200_BF16) ? 1 : 0;
    p.scale_log2e = scale * 1.4426950408889634f;
    const dim3 grid((unsigned)((T + kPfBQ - 1) / kPfBQ), (unsigned)n_heads);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const size_t smem = (size_t)3 * 128 * 128 * (head_dim / 64) + 128 * 128 * 2 + 1024 + 256;
    static thread_local int configured[2] = {-1, -1};
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    if (head_dim == 128) {
        if (configured[0] != dev) {
            B200_CUDA_CHECK(cudaFuncSetAttribute(attn_prefill_tc_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            configured[0] = dev;
        }
        B200_CUDA_CHECK(launch(attn_prefill_tc_kernel<128>, grid, dim3(kPfThreads), smem, stream, pdl != 0, p));
    } else {
        if (configured[1] != dev) {
            B200_CUDA_CHECK(cudaFuncSetAttribute(attn_prefill_tc_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            configured[1] = dev;
        }
        B200_CUDA_CHECK(launch(attn_prefill_tc_kernel<64>, grid, dim3(kPfThreads), smem, stream, pdl != 0, p));
    }
    return B200_OK;
}
