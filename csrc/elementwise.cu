// Small bandwidth-bound kernels around the GEMMs: embedding gather, standalone
// RMSNorm/LayerNorm (prefill-size inputs; decode fuses the norm into the GEMM
// prologue), and arg-max over the vocabulary.
#include "common.h"
#include "ptx.cuh"
#include "llchain.cuh"

namespace b200 {

// ------------------------------------------------------------------ embedding
// out[t, :] = wte[ids[t], :] (+ wpe[pos0 + t or pos[t], :])      16-byte vectors
__global__ void __launch_bounds__(128) embed_kernel(const long long* __restrict__ ids, const int* __restrict__ pos,
                                                   int pos0, const uint16_t* __restrict__ wte,
                                                   const uint16_t* __restrict__ wpe, uint16_t* __restrict__ out,
                                                   int T, int H, int vocab, int max_pos, int is_bf16) {
    pdl_wait_prior_grid();
    pdl_launch_dependents();
    const int t = blockIdx.x;
    if (t >= T) return;
    long long id = ids[t];
    if (id < 0) id = 0;
    if (id >= vocab) id = vocab - 1;
    int ps = pos ? pos[t] : pos0 + t;
    if (ps >= max_pos) ps = max_pos - 1;
    const uint4* src = reinterpret_cast<const uint4*>(wte + (size_t)id * H);
    const uint4* psrc = wpe ? reinterpret_cast<const uint4*>(wpe + (size_t)ps * H) : nullptr;
    uint4* dst = reinterpret_cast<uint4*>(out + (size_t)t * H);
    for (int c = threadIdx.x; c < H / 8; c += blockDim.x) {
        uint4 v = src[c];
        if (psrc) {
            const uint4 w = psrc[c];
            uint32_t a[4] = {v.x, v.y, v.z, v.w};
            const uint32_t b[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float lo = h16_to_float((uint16_t)(a[i] & 0xffff), is_bf16) + h16_to_float((uint16_t)(b[i] & 0xffff), is_bf16);
                const float hi = h16_to_float((uint16_t)(a[i] >> 16), is_bf16) + h16_to_float((uint16_t)(b[i] >> 16), is_bf16);
                a[i] = (uint32_t)float_to_h16(lo, is_bf16) | ((uint32_t)float_to_h16(hi, is_bf16) << 16);
            }
            v = make_uint4(a[0], a[1], a[2], a[3]);
        }
        dst[c] = v;
    }
}

// ------------------------------------------------------------------ head of a flag chain
// x[T, H] (complete in stream order) -> LL words tagged (*seq << 12 | op): 8 elements = 4 words per thread step.
__global__ void __launch_bounds__(256) ll_publish_kernel(const uint16_t* __restrict__ x, int ldx, uint2* __restrict__ out,
                                                        int ld, int T, int H, const int* __restrict__ seq, int op) {
    pdl_wait_prior_grid();
    pdl_launch_dependents();
    const uint32_t tag = ll_tag(ll_read_seq(seq), op);
    const int per_row = H / 8;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < T * per_row; i += gridDim.x * blockDim.x) {
        const int t = i / per_row, c = i - t * per_row;
        const uint4 v = ld_cg_v4(x + (size_t)t * ldx + c * 8);
        uint2* w = out + (size_t)t * ld + c * 4;
        asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1, %2, %3, %2};\n" ::"l"(w), "r"(v.x), "r"(tag), "r"(v.y) : "memory");
        asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1, %2, %3, %2};\n" ::"l"(w + 2), "r"(v.z), "r"(tag), "r"(v.w) : "memory");
    }
}

// ------------------------------------------------------------------ norm
// One CTA per row; the row stays in registers between the statistics and the scale pass.
template <int VPT>  // 16-byte vectors per thread (H <= 256 * 8 * VPT)
__global__ void __launch_bounds__(256) norm_kernel(const uint16_t* __restrict__ x, const uint16_t* __restrict__ w,
                                                  const uint16_t* __restrict__ b, uint16_t* __restrict__ out, int T,
                                                  int H, int ldx, int ldo, float eps, int layer_norm, int is_bf16) {
    __shared__ float s_red[2][8];
    pdl_wait_prior_grid();
    pdl_launch_dependents();
    const int t = blockIdx.x;
    const int nvec = H / 8;
    uint4 v[VPT];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
        const int c = threadIdx.x + i * 256;
        v[i] = (c < nvec) ? *reinterpret_cast<const uint4*>(x + (size_t)t * ldx + c * 8) : make_uint4(0, 0, 0, 0);
        const uint32_t a[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float lo = h16_to_float((uint16_t)(a[j] & 0xffff), is_bf16);
            const float hi = h16_to_float((uint16_t)(a[j] >> 16), is_bf16);
            s1 += lo + hi;
            s2 += lo * lo + hi * hi;
        }
    }
    s1 = warp_sum(s1);
    s2 = warp_sum(s2);
    if ((threadIdx.x & 31) == 0) { s_red[0][threadIdx.x >> 5] = s1; s_red[1][threadIdx.x >> 5] = s2; }
    __syncthreads();
    s1 = 0.f; s2 = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { s1 += s_red[0][i]; s2 += s_red[1][i]; }
    const float inv = 1.0f / (float)H;
    float mean = 0.f, rstd;
    if (layer_norm) {
        mean = s1 * inv;
        rstd = rsqrtf(fmaxf(s2 * inv - mean * mean, 0.f) + eps);
    } else {
        rstd = rsqrtf(s2 * inv + eps);
    }
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
        const int c = threadIdx.x + i * 256;
        if (c >= nvec) continue;
        const uint4 g4 = *reinterpret_cast<const uint4*>(w + c * 8);
        const uint4 b4 = b ? *reinterpret_cast<const uint4*>(b + c * 8) : make_uint4(0, 0, 0, 0);
        uint32_t a[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
        const uint32_t g[4] = {g4.x, g4.y, g4.z, g4.w};
        const uint32_t bb[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float lo = h16_to_float((uint16_t)(a[j] & 0xffff), is_bf16);
            float hi = h16_to_float((uint16_t)(a[j] >> 16), is_bf16);
            const float g0 = h16_to_float((uint16_t)(g[j] & 0xffff), is_bf16);
            const float g1 = h16_to_float((uint16_t)(g[j] >> 16), is_bf16);
            if (layer_norm) {
                lo = (lo - mean) * rstd * g0 + h16_to_float((uint16_t)(bb[j] & 0xffff), is_bf16);
                hi = (hi - mean) * rstd * g1 + h16_to_float((uint16_t)(bb[j] >> 16), is_bf16);
            } else {
                lo = g0 * h16_to_float(float_to_h16(lo * rstd, is_bf16), is_bf16);
                hi = g1 * h16_to_float(float_to_h16(hi * rstd, is_bf16), is_bf16);
            }
            a[j] = (uint32_t)float_to_h16(lo, is_bf16) | ((uint32_t)float_to_h16(hi, is_bf16) << 16);
        }
        *reinterpret_cast<uint4*>(out + (size_t)t * ldo + c * 8) = make_uint4(a[0], a[1], a[2], a[3]);
    }
}

// ------------------------------------------------------------------ argmax
// logits fp32 [T, V] -> ids[t] = first index of the row maximum (torch.argmax tie rule).
// Stage 1: grid (chunks, T) partial (max, idx); the last CTA of a row reduces the partials.
// Vocabulary shards: indices are reported with idx_offset added; init_pair[t] = {value bits, index} is a candidate from
// another shard (same tie rule: the lower index wins), out_pair[t] receives the winner.
__global__ void __launch_bounds__(256) argmax_kernel(const float* __restrict__ logits, int ld, int V,
                                                    float* __restrict__ part_val, int* __restrict__ part_idx,
                                                    int* __restrict__ counters, long long* __restrict__ out_ids,
                                                    int idx_offset, const uint32_t* __restrict__ init_pair,
                                                    uint32_t* __restrict__ out_pair) {
    __shared__ float s_v[8];
    __shared__ int s_i[8];
    __shared__ int s_last;
    pdl_wait_prior_grid();
    pdl_launch_dependents();
    const int t = blockIdx.y, chunks = gridDim.x;
    const float* row = logits + (size_t)t * ld;
    float best = -INFINITY;
    int bidx = 0x7fffffff;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < V; i += chunks * blockDim.x) {
        const float v = row[i];
        if (v > best || (v == best && i < bidx)) { best = v; bidx = i; }
    }
    auto reduce_pair = [](float& v, int& i) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float v2 = __shfl_xor_sync(0xffffffffu, v, o);
            const int i2 = __shfl_xor_sync(0xffffffffu, i, o);
            if (v2 > v || (v2 == v && i2 < i)) { v = v2; i = i2; }
        }
    };
    reduce_pair(best, bidx);
    if ((threadIdx.x & 31) == 0) { s_v[threadIdx.x >> 5] = best; s_i[threadIdx.x >> 5] = bidx; }
    __syncthreads();
    if (threadIdx.x < 32) {
        best = threadIdx.x < 8 ? s_v[threadIdx.x] : -INFINITY;
        bidx = threadIdx.x < 8 ? s_i[threadIdx.x] : 0x7fffffff;
        reduce_pair(best, bidx);
        if (threadIdx.x == 0) {
            part_val[t * chunks + blockIdx.x] = best;
            part_idx[t * chunks + blockIdx.x] = bidx;
            __threadfence();
            const int old = atomicAdd(&counters[t], 1);
            s_last = (old + 1 == chunks);
        }
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (threadIdx.x < 32) {
        best = -INFINITY;
        bidx = 0x7fffffff;
        for (int c = threadIdx.x; c < chunks; c += 32) {
            const float v = __ldcg(part_val + t * chunks + c);
            const int i = __ldcg(part_idx + t * chunks + c);
            if (v > best || (v == best && i < bidx)) { best = v; bidx = i; }
        }
        reduce_pair(best, bidx);
        if (threadIdx.x == 0) {
            if (bidx != 0x7fffffff) bidx += idx_offset;
            if (init_pair) {
                const float v2 = __uint_as_float(__ldcg(init_pair + 2 * t));
                const int i2 = (int)__ldcg(init_pair + 2 * t + 1);
                if (v2 > best || (v2 == best && i2 < bidx)) { best = v2; bidx = i2; }
            }
            if (out_ids) out_ids[t] = (bidx == 0x7fffffff) ? 0 : bidx;
            if (out_pair) { out_pair[2 * t] = __float_as_uint(best); out_pair[2 * t + 1] = (uint32_t)bidx; }
            counters[t] = 0;
        }
    }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_embed(const void* ids, const void* pos, int32_t pos0, const void* wte, const void* wpe, void* out,
                          int32_t T, int32_t H, int32_t vocab, int32_t max_pos, int32_t dtype, int32_t pdl,
                          void* stream_) {
    if (!ids || !wte || !out) return fail(B200_EINVAL, "b200_embed: null pointer argument");
    if (H % 8 != 0) return fail(B200_EINVAL, "b200_embed: hidden size %d not a multiple of 8", H);
    B200_CUDA_CHECK(launch(embed_kernel, dim3(T), dim3(128), 0, static_cast<cudaStream_t>(stream_), pdl != 0,
                           static_cast<const long long*>(ids), static_cast<const int*>(pos), pos0,
                           static_cast<const uint16_t*>(wte), static_cast<const uint16_t*>(wpe),
                           static_cast<uint16_t*>(out), T, H, vocab, max_pos, dtype == B200_BF16 ? 1 : 0));
    return B200_OK;
}

extern "C" int b200_ll_publish(const void* x, int32_t ldx, void* out_ll, int32_t ld_ll, int32_t T, int32_t H,
                               const int32_t* seq, int32_t op, int32_t pdl, void* stream_) {
    if (!x || !out_ll || !seq) return fail(B200_EINVAL, "b200_ll_publish: null pointer argument");
    if (H % 8 != 0 || ld_ll % 4 != 0 || T <= 0 || op <= 0 || op >= (1 << kLLOpBits))
        return fail(B200_EINVAL, "b200_ll_publish: bad sizes T=%d H=%d ld=%d op=%d", T, H, ld_ll, op);
    const int blocks = (T * (H / 8) + 255) / 256;
    B200_CUDA_CHECK(launch(ll_publish_kernel, dim3(blocks > 64 ? 64 : blocks), dim3(256), 0, static_cast<cudaStream_t>(stream_),
                           pdl != 0, static_cast<const uint16_t*>(x), ldx, static_cast<uint2*>(out_ll), ld_ll, T, H, seq, op));
    return B200_OK;
}

namespace b200 {
// Per-token 1/rms with exactly the reduction tree of rms_stats_128 (gemm_core.cuh): 128 threads per token, thread e
// sums chunks e, e+128, ..., warp sums, then the four warp sums in warp order.  One block per token.
__global__ void __launch_bounds__(128) rms_rstd_kernel(const uint16_t* __restrict__ x, int ldx, int K, float eps, int is_bf16,
                                                       float* __restrict__ rstd) {
    __shared__ float red[4];
    pdl_wait_prior_grid();
    pdl_launch_dependents();
    const int e = threadIdx.x, t = blockIdx.x;
    const uint16_t* xr = x + (size_t)t * ldx;
    float s2 = 0.f;
    for (int c = e; c < K / 8; c += 128) {
        const uint4 v = ld_cg_v4(xr + c * 8);
        const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float a = h16_to_float((uint16_t)(w4[i] & 0xffff), is_bf16);
            const float b = h16_to_float((uint16_t)(w4[i] >> 16), is_bf16);
            s2 += a * a + b * b;
        }
    }
    s2 = warp_sum(s2);
    if ((e & 31) == 0) red[e >> 5] = s2;
    __syncthreads();
    if (e == 0) rstd[t] = rsqrtf((red[0] + red[1] + red[2] + red[3]) / (float)K + eps);
}
}  // namespace b200

// rstd[t] = 1 / sqrt(mean(x[t]^2) + eps): feeds b200_gemm_args.rstd so that a many-token decode step computes the
// RMSNorm statistics once instead of once per CTA of the following GEMM.
extern "C" int b200_rms_rstd(const void* x, float* rstd, int32_t T, int32_t K, int32_t ldx, float eps, int32_t dtype,
                             int32_t pdl, void* stream_) {
    if (!x || !rstd) return fail(B200_EINVAL, "b200_rms_rstd: null pointer argument");
    if (T <= 0 || K <= 0 || K % 8 != 0) return fail(B200_EINVAL, "b200_rms_rstd: bad sizes T=%d K=%d", T, K);
    B200_CUDA_CHECK(launch(b200::rms_rstd_kernel, dim3(T), dim3(128), 0, static_cast<cudaStream_t>(stream_), pdl != 0,
                           static_cast<const uint16_t*>(x), ldx, K, eps, dtype == B200_BF16 ? 1 : 0, rstd));
    return B200_OK;
}

extern "C" int b200_norm(const void* x, const void* w, const void* b, void* out, int32_t T, int32_t H, int32_t ldx,
                         int32_t ldo, float eps, int32_t layer_norm, int32_t dtype, int32_t pdl, void* stream_) {
    if (!x || !w || !out) return fail(B200_EINVAL, "b200_norm: null pointer argument");
    if (H % 8 != 0 || H > 256 * 8 * 8) return fail(B200_EINVAL, "b200_norm: unsupported hidden size %d", H);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const uint16_t* xx = static_cast<const uint16_t*>(x);
    const uint16_t* ww = static_cast<const uint16_t*>(w);
    const uint16_t* bb = static_cast<const uint16_t*>(b);
    uint16_t* oo = static_cast<uint16_t*>(out);
    const int bf = dtype == B200_BF16 ? 1 : 0;
    const int vpt = (H / 8 + 255) / 256;
    if (vpt <= 1) {
        B200_CUDA_CHECK(launch(norm_kernel<1>, dim3(T), dim3(256), 0, stream, pdl != 0, xx, ww, bb, oo, T, H, ldx, ldo, eps, layer_norm, bf));
    } else if (vpt <= 2) {
        B200_CUDA_CHECK(launch(norm_kernel<2>, dim3(T), dim3(256), 0, stream, pdl != 0, xx, ww, bb, oo, T, H, ldx, ldo, eps, layer_norm, bf));
    } else if (vpt <= 4) {
        B200_CUDA_CHECK(launch(norm_kernel<4>, dim3(T), dim3(256), 0, stream, pdl != 0, xx, ww, bb, oo, T, H, ldx, ldo, eps, layer_norm, bf));
    } else {
        B200_CUDA_CHECK(launch(norm_kernel<8>, dim3(T), dim3(256), 0, stream, pdl != 0, xx, ww, bb, oo, T, H, ldx, ldo, eps, layer_norm, bf));
    }
    return B200_OK;
}

extern "C" int b200_argmax_partial(const float* logits, int32_t ld, int32_t T, int32_t V, float* part_val,
                                   int32_t* part_idx, int32_t* counters, void* out_ids, int32_t chunks,
                                   int32_t idx_offset, const void* init_pair, void* out_pair, int32_t pdl,
                                   void* stream_) {
    if (!logits || !part_val || !part_idx || !counters || (!out_ids && !out_pair))
        return fail(B200_EINVAL, "b200_argmax: null pointer argument");
    if (chunks <= 0 || chunks > 1024) return fail(B200_EINVAL, "b200_argmax: chunks=%d out of range", chunks);
    B200_CUDA_CHECK(launch(argmax_kernel, dim3(chunks, T), dim3(256), 0, static_cast<cudaStream_t>(stream_), pdl != 0,
                           logits, ld, V, part_val, part_idx, counters, static_cast<long long*>(out_ids), idx_offset,
                           static_cast<const uint32_t*>(init_pair), static_cast<uint32_t*>(out_pair)));
    return B200_OK;
}

extern "C" int b200_argmax(const float* logits, int32_t ld, int32_t T, int32_t V, float* part_val, int32_t* part_idx,
                           int32_t* counters, void* out_ids, int32_t chunks, int32_t pdl, void* stream_) {
    if (!out_ids) return fail(B200_EINVAL, "b200_argmax: null pointer argument");
    return b200_argmax_partial(logits, ld, T, V, part_val, part_idx, counters, out_ids, chunks, 0, nullptr, nullptr, pdl,
                               stream_);
}
