// Session-KV attention kernels (see b200_petals.h):
//   b200_attn_decode   RoPE + in-place KV append + split-KV (flash-decoding) attention,
//                      one launch per layer for T independent sessions' new tokens.
//   b200_rope_append   prefill/chunk: RoPE on q (in place) and k, k/v -> session cache.
//   b200_attn_prefill  causal attention of a chunk against the session cache.
// Decode attention is KV-bandwidth bound (one query per session), so it runs on the
// CUDA cores with 16-byte loads; the tensor-core path is the prefill kernel's job.
#include "attn_core.cuh"

namespace b200 {

// ------------------------------------------------------------------ rope + append
struct RopeAppendParams {
    uint16_t* __restrict__ qkv;        // q rotated in place
    uint16_t* __restrict__ kcache;     // [n_kv][s_max][D]
    uint16_t* __restrict__ vcache;
    const float* __restrict__ inv_freq;
    int T, n_heads, n_kv, D, s_max, ldqkv, pos0, is_bf16;
};

__global__ void __launch_bounds__(128) rope_append_kernel(const RopeAppendParams p) {
    pdl_wait_prior_grid();
    pdl_launch_dependents();
    const int D = p.D, tph = D / 16;
    const int heads = p.n_heads + p.n_kv;
    const long long total = (long long)p.T * heads * tph;
    for (long long item = (long long)blockIdx.x * blockDim.x + threadIdx.x; item < total;
         item += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(item % tph) * 8;
        const int h = (int)((item / tph) % heads);
        const int t = (int)(item / ((long long)tph * heads));
        const int pos = p.pos0 + t;
        uint16_t* src = p.qkv + (size_t)t * p.ldqkv + (size_t)h * D;
        float lo[8], hi[8];
        unpack8(*reinterpret_cast<const uint4*>(src + c), lo, p.is_bf16);
        unpack8(*reinterpret_cast<const uint4*>(src + c + D / 2), hi, p.is_bf16);
        if (p.inv_freq) rope8(lo, hi, c, pos, p.inv_freq, p.is_bf16);
        if (h < p.n_heads) {
            *reinterpret_cast<uint4*>(src + c) = pack8(lo, p.is_bf16);
            *reinterpret_cast<uint4*>(src + c + D / 2) = pack8(hi, p.is_bf16);
        } else {
            const int g = h - p.n_heads;
            uint16_t* kd = p.kcache + ((size_t)g * p.s_max + pos) * D;
            uint16_t* vd = p.vcache + ((size_t)g * p.s_max + pos) * D;
            const uint16_t* vs = p.qkv + (size_t)t * p.ldqkv + (size_t)(heads + g) * D;
            *reinterpret_cast<uint4*>(kd + c) = pack8(lo, p.is_bf16);
            *reinterpret_cast<uint4*>(kd + c + D / 2) = pack8(hi, p.is_bf16);
            *reinterpret_cast<uint4*>(vd + c) = *reinterpret_cast<const uint4*>(vs + c);
            *reinterpret_cast<uint4*>(vd + c + D / 2) = *reinterpret_cast<const uint4*>(vs + c + D / 2);
        }
    }
}

// ------------------------------------------------------------------ prefill (chunk vs cache)
// CTA = (32 queries, one q head).  K/V tiles of 32 cached rows go through shared memory;
// lane = key inside the tile for QK^T, lane = output dim for PV.  Scores never leave the SM.
struct AttnPrefillParams {
    const uint16_t* __restrict__ q;    // rotated q inside the qkv buffer, row stride ldq
    uint16_t* __restrict__ out;        // [T, n_heads*D]
    const uint16_t* __restrict__ kcache;
    const uint16_t* __restrict__ vcache;
    int T, n_heads, n_kv, s_max, ldq, ldo, pos0, is_bf16;
    float scale;
};

template <int D>
__global__ void __launch_bounds__(128) attn_prefill_kernel(const AttnPrefillParams p) {
    constexpr int BQ = 32, BK = 32, NW = 4, QPW = BQ / NW;  // 8 queries per warp
    constexpr int KS = D + 2;                               // padded row (halves): conflict-free column reads
    __shared__ __align__(16) uint16_t s_k[BK * KS];
    __shared__ __align__(16) uint16_t s_v[BK * D];
    __shared__ float s_q[BQ][D];
    __shared__ float s_p[NW][QPW][BK];

    pdl_wait_prior_grid();
    pdl_launch_dependents();
    const int h = blockIdx.y, qb = blockIdx.x;
    const int g = h / (p.n_heads / p.n_kv);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int q0 = qb * BQ;
    const uint16_t* kb = p.kcache + (size_t)g * p.s_max * D;
    const uint16_t* vb = p.vcache + (size_t)g * p.s_max * D;

    for (int i = tid; i < BQ * D; i += 128) {
        const int r = i / D, d = i % D;
        const int t = q0 + r;
        s_q[r][d] = (t < p.T) ? h16_to_float(p.q[(size_t)t * p.ldq + (size_t)h * D + d], p.is_bf16) * p.scale : 0.f;
    }
    float m[QPW], l[QPW], acc[QPW][D / 32];
#pragma unroll
    for (int i = 0; i < QPW; ++i) {
        m[i] = -INFINITY; l[i] = 0.f;
#pragma unroll
        for (int j = 0; j < D / 32; ++j) acc[i][j] = 0.f;
    }
    const int last_q = min(p.T, q0 + BQ) - 1;
    const int kv_end = p.pos0 + last_q + 1;  // causal bound for the block
    for (int k0 = 0; k0 < kv_end; k0 += BK) {
        __syncthreads();
        for (int i = tid; i < BK * D / 8; i += 128) {
            const int r = i / (D / 8), c = (i % (D / 8)) * 8;
            uint4 kv4 = make_uint4(0, 0, 0, 0), vv4 = make_uint4(0, 0, 0, 0);
            if (k0 + r < kv_end) {
                kv4 = *reinterpret_cast<const uint4*>(kb + (size_t)(k0 + r) * D + c);
                vv4 = *reinterpret_cast<const uint4*>(vb + (size_t)(k0 + r) * D + c);
            }
            uint32_t* kd = reinterpret_cast<uint32_t*>(s_k + r * KS + c);  // KS*2 bytes = 4-byte aligned rows
            kd[0] = kv4.x; kd[1] = kv4.y; kd[2] = kv4.z; kd[3] = kv4.w;
            *reinterpret_cast<uint4*>(s_v + r * D + c) = vv4;
        }
        __syncthreads();
        const int key = k0 + lane;
#pragma unroll
        for (int qi = 0; qi < QPW; ++qi) {
            const int r = warp * QPW + qi;
            const int t = q0 + r;
            float sc = 0.f;
            const uint32_t* krow = reinterpret_cast<const uint32_t*>(s_k + lane * KS);
#pragma unroll 8
            for (int d2 = 0; d2 < D / 2; ++d2) {
                const uint32_t w = krow[d2];
                sc += s_q[r][2 * d2] * h16_to_float((uint16_t)(w & 0xffff), p.is_bf16) +
                      s_q[r][2 * d2 + 1] * h16_to_float((uint16_t)(w >> 16), p.is_bf16);
            }
            const bool ok = (t < p.T) && (key <= p.pos0 + t);
            sc = ok ? sc : -INFINITY;
            const float bm = warp_max(sc);
            const float mn = fmaxf(m[qi], bm);
            const float pr = (mn == -INFINITY) ? 0.f : __expf(sc - mn);
            const float corr = (mn == -INFINITY) ? 1.f : __expf(m[qi] - mn);
            l[qi] = l[qi] * corr + warp_sum(pr);
#pragma unroll
            for (int j = 0; j < D / 32; ++j) acc[qi][j] *= corr;
            m[qi] = mn;
            s_p[warp][qi][lane] = pr;
        }
        __syncwarp();
#pragma unroll 4
        for (int kk = 0; kk < BK; ++kk) {
            float vv[D / 32];
#pragma unroll
            for (int j = 0; j < D / 32; ++j) vv[j] = h16_to_float(s_v[kk * D + j * 32 + lane], p.is_bf16);
#pragma unroll
            for (int qi = 0; qi < QPW; ++qi) {
                const float pr = s_p[warp][qi][kk];
#pragma unroll
                for (int j = 0; j < D / 32; ++j) acc[qi][j] += pr * vv[j];
            }
        }
        __syncwarp();
    }
#pragma unroll
    for (int qi = 0; qi < QPW; ++qi) {
        const int t = q0 + warp * QPW + qi;
        if (t < p.T) {
#pragma unroll
            for (int j = 0; j < D / 32; ++j)
                p.out[(size_t)t * p.ldo + (size_t)h * D + j * 32 + lane] = float_to_h16(acc[qi][j] / l[qi], p.is_bf16);
        }
    }
}

}  // namespace b200

using namespace b200;

B200_DEFINE_TRACE_SETTER(b200_trace_set_attn)

extern "C" int64_t b200_attn_decode_scratch_bytes(int32_t T, int32_t n_heads, int32_t head_dim, int32_t splits) {
    return (int64_t)T * n_heads * splits * (head_dim + 2) * (int64_t)sizeof(float);
}

extern "C" int b200_attn_decode(const b200_attn_decode_args* a, void* stream_) {
    if (!a || !a->qkv || !a->out || !a->kv_ptrs || !a->pos || !a->scratch || !a->counters)
        return fail(B200_EINVAL, "b200_attn_decode: null pointer argument");
    if (a->n_kv_heads <= 0 || a->n_heads % a->n_kv_heads != 0)
        return fail(B200_EINVAL, "b200_attn_decode: n_heads=%d not a multiple of n_kv_heads=%d", a->n_heads, a->n_kv_heads);
    const int group = a->n_heads / a->n_kv_heads;
    AttnDecodeParams p;
    p.qkv = static_cast<const uint16_t*>(a->qkv);
    p.out = static_cast<uint16_t*>(a->out);
    p.kv_ptrs = reinterpret_cast<const unsigned long long*>(a->kv_ptrs);
    p.pos = a->pos;
    p.inv_freq = a->inv_freq;
    p.scratch = a->scratch;
    p.counters = a->counters;
    p.ll = static_cast<unsigned long long*>(a->ll_scratch);
    p.T = a->T; p.n_heads = a->n_heads; p.n_kv = a->n_kv_heads; p.s_max = a->s_max;
    p.splits = a->splits > 0 ? a->splits : 1;
    p.ldqkv = a->ldqkv; p.ldo = a->ldo;
    p.scale = a->scale;
    p.is_bf16 = (a->dtype == B200_BF16) ? 1 : 0;
    p.ch = LLChain{};
    if (a->chain.seq) {
        const b200_ll_chain& c = a->chain;
        if (!c.x_ll || !c.out_ll || !a->ll_scratch || !a->pdl)
            return fail(B200_EINVAL, "b200_attn_decode: a flag chain needs x_ll, out_ll, ll_scratch and pdl = 1");
        if ((c.ldx_ll % 4) != 0 || (reinterpret_cast<uintptr_t>(c.x_ll) & 31) != 0 || (a->head_dim % 16) != 0)
            return fail(B200_EINVAL, "b200_attn_decode: chain.x_ll must be 32-byte aligned with ldx_ll %% 4 == 0");
        if (c.x_op <= 0 || c.x_op >= (1 << kLLOpBits) || c.out_op <= 0 || c.out_op >= (1 << kLLOpBits))
            return fail(B200_EINVAL, "b200_attn_decode: chain launch ids must be in 1..%d", (1 << kLLOpBits) - 1);
        p.ch.seq = c.seq;
        p.ch.x = static_cast<const uint2*>(c.x_ll);
        p.ch.out = static_cast<uint2*>(c.out_ll);
        p.ch.ldx = c.ldx_ll; p.ch.ldo = c.ldo_ll;
        p.ch.x_op = c.x_op; p.ch.out_op = c.out_op;
    }
    const dim3 grid((unsigned)(a->T * a->n_kv_heads * p.splits));
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const bool pdl = a->pdl != 0;
    // Cache rows reach the SM either by plain 16-byte loads (+ an L2 prefetch above the PDL wait) or through shared
    // memory by bulk async copies.  Measured per layer (32 kv heads, d = 128, 1xB200, scripts/exp_attn.py): with 512
    // CTAs (16 splits) the loads win, 11.4 vs 12.3 us at 1500 cached tokens - 3.5 CTAs per SM hide their latency and
    // the copies' fixed cost shows; with 128 CTAs (4 splits, or a GQA model with 8 kv heads) the copies win, 16.8 vs
    // 20.8 us - one CTA per SM has nothing but its own prefetch depth.  So: bulk copies when the grid leaves at most
    // two CTAs per SM; B200_ATTN_BULK=0/1 forces one or the other.
    static const int bulk_env = [] { const char* e = getenv("B200_ATTN_BULK"); return e ? (e[0] == '0' ? 0 : 1) : -1; }();
    const bool bulk = bulk_env >= 0 ? bulk_env == 1 : (int)grid.x <= 2 * sm_count_cached();
#define B200_ATTN_CASE(D_, G_)                                                                     \
    if (a->head_dim == D_ && group == G_) {                                                        \
        if (bulk) {                                                                                \
            static int configured_dev = -1;                                                        \
            int dev = 0;                                                                           \
            B200_CUDA_CHECK(cudaGetDevice(&dev));                                                  \
            if (configured_dev != dev) {                                                           \
                B200_CUDA_CHECK(cudaFuncSetAttribute(attn_decode_kernel<D_, G_, 1>,                \
                                                     cudaFuncAttributeMaxDynamicSharedMemorySize,  \
                                                     (int)sizeof(AttnBulkTiles<D_>)));             \
                configured_dev = dev;                                                              \
            }                                                                                      \
            B200_CUDA_CHECK(launch(attn_decode_kernel<D_, G_, 1>, grid, dim3(kAttnThreads),        \
                                   sizeof(AttnBulkTiles<D_>), stream, pdl, p));                    \
        } else {                                                                                   \
            B200_CUDA_CHECK(launch(attn_decode_kernel<D_, G_, 0>, grid, dim3(kAttnThreads), 0, stream, pdl, p)); \
        }                                                                                          \
        return B200_OK;                                                                            \
    }
    B200_ATTN_CASE(128, 1) B200_ATTN_CASE(128, 2) B200_ATTN_CASE(128, 4) B200_ATTN_CASE(128, 8)
    B200_ATTN_CASE(64, 1) B200_ATTN_CASE(64, 2) B200_ATTN_CASE(64, 4)
#undef B200_ATTN_CASE
    return fail(B200_EINVAL, "b200_attn_decode: unsupported head_dim=%d / group=%d", a->head_dim, group);
}

extern "C" int b200_rope_append(void* qkv, void* kcache, void* vcache, const float* inv_freq, int32_t T,
                                int32_t n_heads, int32_t n_kv_heads, int32_t head_dim, int32_t s_max,
                                int32_t ldqkv, int32_t pos0, int32_t dtype, int32_t pdl, void* stream_) {
    if (!qkv || !kcache || !vcache) return fail(B200_EINVAL, "b200_rope_append: null pointer argument");
    if (head_dim % 16 != 0) return fail(B200_EINVAL, "b200_rope_append: head_dim %d not a multiple of 16", head_dim);
    if (pos0 + T > s_max) return fail(B200_ELIMIT, "b200_rope_append: %d+%d tokens exceed the cache capacity %d", pos0, T, s_max);
    RopeAppendParams p;
    p.qkv = static_cast<uint16_t*>(qkv);
    p.kcache = static_cast<uint16_t*>(kcache);
    p.vcache = static_cast<uint16_t*>(vcache);
    p.inv_freq = inv_freq;
    p.T = T; p.n_heads = n_heads; p.n_kv = n_kv_heads; p.D = head_dim; p.s_max = s_max; p.ldqkv = ldqkv;
    p.pos0 = pos0; p.is_bf16 = (dtype == B200_BF16) ? 1 : 0;
    const long long total = (long long)T * (n_heads + n_kv_heads) * (head_dim / 16);
    int blocks = (int)((total + 127) / 128);
    if (blocks > 148 * 8) blocks = 148 * 8;
    B200_CUDA_CHECK(launch(rope_append_kernel, dim3(blocks), dim3(128), 0, static_cast<cudaStream_t>(stream_), pdl != 0, p));
    return B200_OK;
}

extern "C" int b200_attn_prefill(const void* q, void* out, const void* kcache, const void* vcache, int32_t T,
                                 int32_t n_heads, int32_t n_kv_heads, int32_t head_dim, int32_t s_max, int32_t ldq,
                                 int32_t ldo, int32_t pos0, float scale, int32_t dtype, int32_t pdl, void* stream_) {
    if (!q || !out || !kcache || !vcache) return fail(B200_EINVAL, "b200_attn_prefill: null pointer argument");
    AttnPrefillParams p;
    p.q = static_cast<const uint16_t*>(q);
    p.out = static_cast<uint16_t*>(out);
    p.kcache = static_cast<const uint16_t*>(kcache);
    p.vcache = static_cast<const uint16_t*>(vcache);
    p.T = T; p.n_heads = n_heads; p.n_kv = n_kv_heads; p.s_max = s_max; p.ldq = ldq; p.ldo = ldo; p.pos0 = pos0;
    p.is_bf16 = (dtype == B200_BF16) ? 1 : 0;
    p.scale = scale;
    const dim3 grid((unsigned)((T + 31) / 32), (unsigned)n_heads);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (head_dim == 128) {
        B200_CUDA_CHECK(launch(attn_prefill_kernel<128>, grid, dim3(128), 0, stream, pdl != 0, p));
    } else if (head_dim == 64) {
        B200_CUDA_CHECK(launch(attn_prefill_kernel<64>, grid, dim3(128), 0, stream, pdl != 0, p));
    } else {
        return fail(B200_EINVAL, "b200_attn_prefill: unsupported head_dim=%d", head_dim);
    }
    return B200_OK;
}
