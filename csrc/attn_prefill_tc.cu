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
//
// Two kernels live here:
//   attn_prefill_tc_kernel   (v1)  one 128-query tile per CTA, loads / MMAs / softmax back to back
//   attn_prefill_tc2_kernel  (v2)  the production path: warp-specialised and software-pipelined
//       warp 8      TMA producer: Q tiles once, then K(0) V(0) K(1) V(1) ... through a 5-stage ring
//                   (tensor maps bounded at the valid cache length: rows past it arrive as zeros)
//       warp 9      MMA issuer, ping-pong over the two query tiles: PV_A(j), S_A(j+1), PV_B(j), S_B(j+1) ...
//                   while one group's softmax runs, the tensor core works for the other group
//       warps 0-3   softmax group A (query tile 0), warps 4-7 softmax group B (query tile 1), registers moved
//                   to them from the two worker warps' group with setmaxnreg: one thread per
//                   query row reads its 128 scores from TMEM in one go and writes the 16-bit P back over the
//                   first half of the same TMEM columns (P is the tensor-memory A operand of the P V product,
//                   it never touches shared memory); O accumulates in TMEM and is rescaled only when the
//                   running maximum grew by more than 2^8 (exactness is restored by the final 1/l)
//   K/V tiles are loaded once per 256 queries; CTAs are ordered heaviest-first (causal work is triangular).
#include "attn_core.cuh"
#include "tma_host.h"

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


// ---------------------------------------------------------------------------------------------- v2
constexpr int kPf2Threads = 384;   // warps 0-3: softmax group A, 4-7: group B, 8: TMA, 9: MMA, 10-11: idle
constexpr int kPf2Ring = 5;

struct AttnPrefillTc2Params {
    uint16_t* __restrict__ out;
    int T, n_heads, n_kv, ldo, pos0, tiles_per_cta;
    float scale_log2e;
};

__device__ __forceinline__ void tma_load_2d_attn(uint32_t smem_dst, const CUtensorMap* tmap, int c0, int c1, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];\n" ::
            "r"(smem_dst),
        "l"(tmap), "r"(c0), "r"(c1), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void tma_load_3d_attn(uint32_t smem_dst, const CUtensorMap* tmap, int c0, int c1, int c2,
                                                 uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];\n" ::
            "r"(smem_dst),
        "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
        : "memory");
}

struct RingPos {
    int stage;
    uint32_t par;
    __device__ __forceinline__ void advance(int by) {
        stage += by;
        if (stage >= kPf2Ring) { stage -= kPf2Ring; par ^= 1; }
    }
};

template <bool BF16>
__device__ __forceinline__ uint32_t pack2(float a, float b) {
    if (BF16) {
        const __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<const uint32_t*>(&v);
    }
    const __half2 v = __floats2half2_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&v);
}

template <int D, bool BF16>
__global__ void __launch_bounds__(kPf2Threads, 1)
attn_prefill_tc2_kernel(const AttnPrefillTc2Params p, const __grid_constant__ CUtensorMap tmap_q,
                        const __grid_constant__ CUtensorMap tmap_k, const __grid_constant__ CUtensorMap tmap_v) {
    constexpr int HALVES = D / 64;
    constexpr int TILE = 128 * 128 * HALVES;       // bytes of a [128 x D] 16-bit tile
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* s_q = smem;                           // [2][TILE]
    uint8_t* s_ring = s_q + 2 * TILE;              // [kPf2Ring][TILE]
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_ring + kPf2Ring * TILE);
    uint64_t* q_full = bars;                       // [1]
    uint64_t* ring_full = bars + 1;                // [kPf2Ring]
    uint64_t* ring_empty = ring_full + kPf2Ring;   // [kPf2Ring]
    uint64_t* s_full = ring_empty + kPf2Ring;      // [2] scores of tile w are in TMEM
    uint64_t* p_ready = s_full + 2;                // [2] 128 arrivals: P_w is in TMEM, O_w is rescaled
    uint64_t* o_full = p_ready + 2;                // [2] O_w += P_w V has completed
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int h = blockIdx.x;
    const int qb = (int)gridDim.y - 1 - (int)blockIdx.y;      // heaviest query blocks are scheduled first
    const int q0 = qb * 128 * p.tiles_per_cta;
    const int g = h / (p.n_heads / p.n_kv);

    if (threadIdx.x == 0) {
        mbar_init(q_full, 1);
        for (int i = 0; i < kPf2Ring; ++i) { mbar_init(&ring_full[i], 1); mbar_init(&ring_empty[i], 1); }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&s_full[i], 1); mbar_init(&p_ready[i], 128); mbar_init(&o_full[i], 1);
        }
        fence_mbar_init();
        asm volatile("prefetch.tensormap [%0];\n" ::"l"(&tmap_q) : "memory");
        asm volatile("prefetch.tensormap [%0];\n" ::"l"(&tmap_k) : "memory");
        asm volatile("prefetch.tensormap [%0];\n" ::"l"(&tmap_v) : "memory");
    }
    if (warp == 9) tmem_alloc<512>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    pdl_wait_prior_grid();
    pdl_launch_dependents();

    // key blocks per query tile (causal): tile w sees keys [0, pos0 + last query of the tile]
    int n_blk[2];
#pragma unroll
    for (int w = 0; w < 2; ++w) {
        const int first = q0 + w * 128;
        n_blk[w] = (w < p.tiles_per_cta && first < p.T) ? (p.pos0 + min(p.T, first + 128) + 127) / 128 : 0;
    }
    const int n_max = max(n_blk[0], n_blk[1]);

    if (warp >= 8) {
      // the whole worker warpgroup gives registers up with ONE instruction (setmaxnreg is warpgroup-aligned)
      asm volatile("setmaxnreg.dec.sync.aligned.u32 72;\n");
      if (warp == 8) {
        // ------------------------------------------------------------ TMA producer
        if (elect_one()) {
            const int nq = n_blk[1] > 0 ? 2 : 1;
            mbar_arrive_expect_tx(q_full, (uint32_t)(nq * TILE));
            for (int w = 0; w < nq; ++w)
                for (int hh = 0; hh < HALVES; ++hh)
                    tma_load_2d_attn(smem_u32(s_q + w * TILE + hh * (128 * 128)), &tmap_q, h * D + hh * 64, q0 + w * 128, q_full);
            RingPos r{0, 0};
            for (int j = 0; j < n_max; ++j) {
#pragma unroll
                for (int kv = 0; kv < 2; ++kv) {
                    mbar_wait(&ring_empty[r.stage], r.par ^ 1);
                    mbar_arrive_expect_tx(&ring_full[r.stage], (uint32_t)TILE);
                    for (int hh = 0; hh < HALVES; ++hh)
                        tma_load_3d_attn(smem_u32(s_ring + r.stage * TILE + hh * (128 * 128)), kv == 0 ? &tmap_k : &tmap_v,
                                         hh * 64, j * 128, g, &ring_full[r.stage]);
                    r.advance(1);
                }
            }
        }
        __syncwarp();
      } else if (warp == 9) {
        // ------------------------------------------------------------ MMA issuer
        const uint32_t idesc_s = umma_idesc(128, 128, BF16, BF16, 0, 0);
        const uint32_t idesc_o = umma_idesc(128, D, BF16, BF16, 0, 1);   // B = V is MN-major
        const uint32_t q_addr = smem_u32(s_q), ring_addr = smem_u32(s_ring);
        RingPos rk{0, 0}, rv{1, 0};   // K(j) is ring item 2j, V(j) is item 2j + 1
        // S_w(j+1) overwrites the columns P_w(j) lives in: it is issued after P_w(j) V(j), and the tensor
        // core executes in issue order.
        auto issue_s = [&](int w, const RingPos& r) {
            mbar_wait(&ring_full[r.stage], r.par);
            tc_fence_after();
            if (elect_one()) {
#pragma unroll
                for (int ks = 0; ks < D / 16; ++ks) {
                    const uint32_t off = (ks / 4) * (128 * 128) + (ks % 4) * 32;
                    umma_f16(tmem_base + w * 128, umma_desc_kmajor_sw128(q_addr + w * TILE + off, 1024),
                             umma_desc_kmajor_sw128(ring_addr + r.stage * TILE + off, 1024), idesc_s, ks > 0 ? 1u : 0u);
                }
                umma_commit(&s_full[w]);
            }
            __syncwarp();
        };
        auto issue_pv = [&](int w, int j, const RingPos& r) {
            mbar_wait(&ring_full[r.stage], r.par);
            mbar_wait(&p_ready[w], (uint32_t)j & 1);
            tc_fence_after();
            if (elect_one()) {
#pragma unroll
                for (int ks = 0; ks < 8; ++ks) {
                    // A = P_w from tensor memory: 16 keys = 8 columns per k-step; B = V: 16 key rows = 2 KiB
                    umma_f16_ts(tmem_base + 256 + w * D, tmem_base + w * 128 + ks * 8,
                                umma_desc_mnmajor_sw128(ring_addr + r.stage * TILE + ks * 2048, 128 * 128, 1024), idesc_o,
                                (j > 0 || ks > 0) ? 1u : 0u);
                }
                umma_commit(&o_full[w]);
            }
            __syncwarp();
        };
        auto release = [&](const RingPos& r) {
            if (elect_one()) umma_commit(&ring_empty[r.stage]);
            __syncwarp();
        };
        mbar_wait(q_full, 0);
        if (n_blk[0] > 0) issue_s(0, rk);
        if (n_blk[1] > 0) issue_s(1, rk);
        release(rk);
        for (int j = 0; j < n_max; ++j) {
            const bool more = j + 1 < n_max;
            if (more) rk.advance(2);
            if (j < n_blk[0]) issue_pv(0, j, rv);
            if (j + 1 < n_blk[0]) issue_s(0, rk);
            if (j < n_blk[1]) issue_pv(1, j, rv);
            if (j + 1 < n_blk[1]) issue_s(1, rk);
            release(rv);
            if (more) release(rk);
            rv.advance(2);
        }
      }
    } else {
        // ------------------------------------------------------------ softmax groups
        asm volatile("setmaxnreg.inc.sync.aligned.u32 216;\n");
        const int w = warp >> 2;                    // 0: query tile A, 1: query tile B
        const int lg = warp & 3;                    // TMEM lane quarter this warp may touch
        const int row = lg * 32 + lane;
        const int t_row = q0 + w * 128 + row;
        const int q_pos = p.pos0 + t_row;
        const int nb = n_blk[w];
        const uint32_t t_lane = (uint32_t)(lg * 32) << 16;
        const uint32_t tmem_s = tmem_base + w * 128 + t_lane;
        const uint32_t tmem_o = tmem_base + 256 + w * D + t_lane;
        float m_used = -INFINITY, l_run = 0.f;
        for (int j = 0; j < nb; ++j) {
            const int k0 = j * 128;
            mbar_wait(&s_full[w], (uint32_t)j & 1);
            tc_fence_after();
            float s[128];
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                uint32_t r[16];
                tmem_ld16(tmem_s + c * 16, r);
#pragma unroll
                for (int i = 0; i < 16; ++i) s[c * 16 + i] = __uint_as_float(r[i]);
            }
            tmem_ld_wait();
            if (k0 + 127 > p.pos0 + q0 + w * 128) { // the block crosses the diagonal of this tile
                const int lim = q_pos - k0;
#pragma unroll
                for (int i = 0; i < 128; ++i) s[i] = (i <= lim) ? s[i] : -INFINITY;
            }
            // eight independent chains: two warps per scheduler cannot hide a 127-deep dependent FMNMX chain
            float mx[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) mx[i] = s[i];
#pragma unroll
            for (int i = 8; i < 128; ++i) mx[i & 7] = fmaxf(mx[i & 7], s[i]);
            const float m_blk = fmaxf(fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])), fmaxf(fmaxf(mx[4], mx[5]), fmaxf(mx[6], mx[7])));
            const float m_new = fmaxf(m_used, m_blk);
            const bool need = (m_new - m_used) * p.scale_log2e > 8.0f;
            const bool any = __any_sync(0xffffffffu, need);
            if (j > 0) {                            // P_w and O_w are free once the previous P V has completed
                mbar_wait(&o_full[w], (uint32_t)(j - 1) & 1);
                tc_fence_after();
            }
            if (any) {
                const float corr = need ? ex2_approx((m_used - m_new) * p.scale_log2e) : 1.0f;
                if (need) { l_run *= corr; m_used = m_new; }
                if (j > 0) {
#pragma unroll
                    for (int c = 0; c < D / 16; ++c) {
                        uint32_t r[16];
                        tmem_ld16(tmem_o + c * 16, r);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 16; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * corr);
                        tmem_st16(tmem_o + c * 16, r);
                    }
                    tmem_st_wait();
                }
            }
            const float neg_m = -m_used * p.scale_log2e;
            float ls[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int c = 0; c < 4; ++c) {           // 32 keys = 16 packed columns per tcgen05.st
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const float a = ex2_approx(fmaf(s[c * 32 + 2 * i], p.scale_log2e, neg_m));
                    const float b = ex2_approx(fmaf(s[c * 32 + 2 * i + 1], p.scale_log2e, neg_m));
                    ls[i & 3] += a + b;
                    pk[i] = pack2<BF16>(a, b);
                }
                tmem_st16(tmem_s + c * 16, pk);
            }
            l_run += (ls[0] + ls[1]) + (ls[2] + ls[3]);
            tmem_st_wait();
            tc_fence_before();
            mbar_arrive(&p_ready[w]);
        }
        if (nb > 0) {
            mbar_wait(&o_full[w], (uint32_t)(nb - 1) & 1);
            tc_fence_after();
            const float inv = 1.0f / l_run;
            uint16_t* dst = p.out + (size_t)t_row * p.ldo + (size_t)h * D;
#pragma unroll
            for (int c = 0; c < D / 16; ++c) {
                uint32_t r[16];
                tmem_ld16(tmem_o + c * 16, r);
                tmem_ld_wait();
                if (t_row < p.T) {
                    uint32_t w8[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        w8[i] = pack2<BF16>(__uint_as_float(r[2 * i]) * inv, __uint_as_float(r[2 * i + 1]) * inv);
                    *reinterpret_cast<uint4*>(dst + c * 16) = make_uint4(w8[0], w8[1], w8[2], w8[3]);
                    *reinterpret_cast<uint4*>(dst + c * 16 + 8) = make_uint4(w8[4], w8[5], w8[6], w8[7]);
                }
            }
            tc_fence_before();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 9) tmem_dealloc<512>(tmem_base);
}

template <int D, bool BF16>
static int launch_tc2(const AttnPrefillTc2Params& p, const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv,
                      dim3 grid, bool pdl, cudaStream_t stream) {
    constexpr size_t smem = (size_t)(2 + kPf2Ring) * 128 * 128 * (D / 64) + 1024 + 256;
    static_assert(smem <= 227 * 1024, "shared memory budget exceeded");
    static thread_local int configured_dev = -1;
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    if (configured_dev != dev) {
        B200_CUDA_CHECK(cudaFuncSetAttribute(attn_prefill_tc2_kernel<D, BF16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured_dev = dev;
    }
    B200_CUDA_CHECK(launch(attn_prefill_tc2_kernel<D, BF16>, grid, dim3(kPf2Threads), smem, stream, pdl, p, mq, mk, mv));
    return B200_OK;
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
    static const int version = [] { const char* e = getenv("B200_ATTN_PREFILL_V"); return e ? atoi(e) : 2; }();
    if (version != 1 && (reinterpret_cast<uintptr_t>(q) & 15) == 0 && pos0 + T <= s_max) {
        EncodeTiledFn enc = encode_tiled_fn();
        if (!enc) return fail(B200_ECUDA, "b200_attn_prefill_tc: cuTensorMapEncodeTiled is not available from the driver");
        const CUtensorMapDataType dt = (dtype == B200_BF16) ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
        CUtensorMap mq, mk, mv;
        const cuuint32_t estr[3] = {1, 1, 1};
        {
            const cuuint64_t dims[2] = {(cuuint64_t)n_heads * head_dim, (cuuint64_t)T};
            const cuuint64_t strides[1] = {(cuuint64_t)ldq * 2};
            const cuuint32_t box[2] = {64, 128};
            const CUresult r = enc(&mq, dt, 2, const_cast<void*>(q), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) return fail(B200_ECUDA, "b200_attn_prefill_tc: tensor map for q failed (%d)", (int)r);
        }
        for (int i = 0; i < 2; ++i) {
            // bounded at the valid length: rows the cache has not been written to arrive as zeros
            const cuuint64_t dims[3] = {(cuuint64_t)head_dim, (cuuint64_t)(pos0 + T), (cuuint64_t)n_kv_heads};
            const cuuint64_t strides[2] = {(cuuint64_t)head_dim * 2, (cuuint64_t)s_max * head_dim * 2};
            const cuuint32_t box[3] = {64, 128, 1};
            const CUresult r = enc(i == 0 ? &mk : &mv, dt, 3, const_cast<void*>(i == 0 ? kcache : vcache), dims, strides, box, estr,
                                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) return fail(B200_ECUDA, "b200_attn_prefill_tc: tensor map for k/v failed (%d)", (int)r);
        }
        static const int forced_tiles = [] { const char* e = getenv("B200_ATTN_PREFILL_TILES"); return e ? atoi(e) : 0; }();
        AttnPrefillTc2Params p2;
        p2.out = static_cast<uint16_t*>(out);
        p2.T = T; p2.n_heads = n_heads; p2.n_kv = n_kv_heads; p2.ldo = ldo; p2.pos0 = pos0;
        p2.scale_log2e = scale * 1.4426950408889634f;
        // two query tiles per CTA share every K/V load; short prompts keep one tile per CTA to fill the SMs
        const int pairs = (T + 255) / 256;
        p2.tiles_per_cta = forced_tiles ? forced_tiles : ((pairs * n_heads * 4 >= sm_count_cached() * 3) ? 2 : 1);
        const dim3 grid2((unsigned)n_heads, (unsigned)((T + 128 * p2.tiles_per_cta - 1) / (128 * p2.tiles_per_cta)));
        cudaStream_t st = static_cast<cudaStream_t>(stream_);
        if (head_dim == 128)
            return dtype == B200_BF16 ? launch_tc2<128, true>(p2, mq, mk, mv, grid2, pdl != 0, st)
                                      : launch_tc2<128, false>(p2, mq, mk, mv, grid2, pdl != 0, st);
        return dtype == B200_BF16 ? launch_tc2<64, true>(p2, mq, mk, mv, grid2, pdl != 0, st)
                                  : launch_tc2<64, false>(p2, mq, mk, mv, grid2, pdl != 0, st);
    }
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
