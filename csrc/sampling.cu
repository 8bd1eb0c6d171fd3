// On-device sampling with the reference's semantics (src/rpc_handler.py:327-403): count-based
// repetition penalty over the last <= 50 tokens (+ the triple-repeat rule), softmax(logits / T),
// top-k, top-p over the descending order (keep entries whose inclusive cumulative mass <= top_p,
// always the largest), renormalise, draw with a caller-supplied uniform number.
//
// One CTA of 1024 threads per row; no sort: the k-th largest probability and the nucleus boundary
// are found by a 4 x 8-bit radix descent over the float bit patterns (probabilities are >= 0, so
// their bit patterns order like integers) with shared-memory histograms — counts for top-k,
// fixed-point (2^-40) masses for top-p so the result does not depend on atomic ordering.
// Ties at the k-th value are all kept (torch.topk keeps an arbitrary subset of them).
#include "common.h"
#include "ptx.cuh"

namespace b200 {

constexpr int kSampThreads = 1024;
constexpr int kMaxRecent = 50;

struct SampleRow {
    float temperature, top_p, penalty;
    int top_k, n_recent;
    int recent[kMaxRecent];
};

// Device-resident generator state of one session (b200_sample_stateful): the uniform number of a draw is
// Philox4x32-10(key = seed, counter = draws made so far), so a session's random stream depends on nothing but its
// seed - not on the batch it is sampled in, the GPU it runs on or how many pipeline stages there are.
struct SampleState {
    unsigned long long seed;
    unsigned long long offset;
};

__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
    const unsigned long long p0 = (unsigned long long)0xD2511F53u * c[0];
    const unsigned long long p1 = (unsigned long long)0xCD9E8D57u * c[2];
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
    c[0] = n0; c[1] = (uint32_t)p1; c[2] = n2; c[3] = (uint32_t)p0;
}
// First word of Philox4x32-10 as a float in [0, 1) with 24 random bits.
__device__ __forceinline__ float philox_uniform(unsigned long long seed, unsigned long long offset) {
    uint32_t c[4] = {(uint32_t)offset, (uint32_t)(offset >> 32), 0u, 0u};
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        philox_round(c, k0, k1);
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return (float)(c[0] >> 8) * (1.0f / 16777216.0f);
}

__device__ __forceinline__ float block_reduce_max(float v, float* red) {
    v = warp_max(v);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = (threadIdx.x < 32) ? red[threadIdx.x] : -INFINITY;
    if (threadIdx.x < 32) { r = warp_max(r); if (threadIdx.x == 0) red[0] = r; }
    __syncthreads();
    r = red[0];
    __syncthreads();
    return r;
}
// Deterministic block sum (fixed tree order).
__device__ __forceinline__ double block_reduce_sum(double v, double* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double r = (threadIdx.x < 32) ? red[threadIdx.x] : 0.0;
    if (threadIdx.x < 32) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
        if (threadIdx.x == 0) red[0] = r;
    }
    __syncthreads();
    r = red[0];
    __syncthreads();
    return r;
}

// Stateless: rows[t] + uniforms[t].  Stateful (states != null): row / state of CTA t = rows / states[row_index ?
// row_index[t] : t]; the uniform comes from the session's Philox stream, and after the draw the CTA appends the token to
// the row's recent-token window and advances the stream - nothing of a session's sampling state lives on the host.
__global__ void __launch_bounds__(kSampThreads) sample_kernel(const float* __restrict__ logits, int ld, int V,
                                                             SampleRow* __restrict__ rows_all,
                                                             const float* __restrict__ uniforms, float* __restrict__ probs,
                                                             long long* __restrict__ out_ids,
                                                             SampleState* __restrict__ states,
                                                             const int* __restrict__ row_index) {
    __shared__ float s_redf[32];
    __shared__ double s_redd[32];
    __shared__ unsigned long long s_hist[256];
    __shared__ int s_pen_tok[kMaxRecent + 1];
    __shared__ float s_pen_fac[kMaxRecent + 1];
    __shared__ int s_npen;
    __shared__ unsigned int s_prefix;
    __shared__ unsigned long long s_above;
    __shared__ double s_seg[kSampThreads];
    __shared__ int s_pick;

    pdl_wait_prior_grid();
    pdl_launch_dependents();
    const int t = blockIdx.x, tid = threadIdx.x;
    const int ri = (states && row_index) ? row_index[t] : t;
    SampleRow& R = rows_all[ri];
    const float u01 = states ? philox_uniform(states[ri].seed, states[ri].offset) : uniforms[t];
    const float* x = logits + (size_t)t * ld;
    float* p = probs + (size_t)t * V;

    // ---- 1. repetition penalty factors (first occurrences only)
    if (tid == 0) s_npen = 0;
    __syncthreads();
    const int nr = min(R.n_recent, kMaxRecent);
    if (tid < nr && R.penalty != 1.0f) {
        const int tok = R.recent[tid];
        bool first = tok >= 0 && tok < V;
        int count = 0;
        for (int j = 0; j < nr; ++j) {
            if (R.recent[j] == tok) { if (j < tid) first = false; ++count; }
        }
        if (first) {
            float f = powf(R.penalty, (float)count);
            if (nr >= 3 && R.recent[nr - 1] == tok && R.recent[nr - 2] == tok && R.recent[nr - 3] == tok)
                f *= R.penalty * R.penalty * R.penalty;
            const int slot = atomicAdd(&s_npen, 1);
            s_pen_tok[slot] = tok;
            s_pen_fac[slot] = f;
        }
    }
    __syncthreads();
    auto adjusted = [&](int i) {
        float v = x[i];
        for (int j = 0; j < s_npen; ++j)
            if (s_pen_tok[j] == i) v = v > 0.f ? v / s_pen_fac[j] : v * s_pen_fac[j];
        return v;
    };
    // ---- 2. softmax numerators
    const float inv_t = 1.0f / fmaxf(R.temperature, 1e-5f);
    float mx = -INFINITY;
    for (int i = tid; i < V; i += kSampThreads) mx = fmaxf(mx, adjusted(i));
    mx = block_reduce_max(mx, s_redf);
    double zs = 0.0;
    for (int i = tid; i < V; i += kSampThreads) {
        const float e = __expf((adjusted(i) - mx) * inv_t);
        p[i] = e;
        zs += e;
    }
    const double Z = block_reduce_sum(zs, s_redd);
    __syncthreads();

    // ---- 3. top-k by radix select on the bit patterns
    if (R.top_k > 0 && R.top_k < V) {
        unsigned int prefix = 0;
        unsigned long long above = 0;   // elements strictly above the current prefix range
        for (int pass = 0; pass < 4; ++pass) {
            const int shift = 24 - 8 * pass;
            const unsigned int mask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
            for (int i = tid; i < 256; i += kSampThreads) s_hist[i] = 0;
            __syncthreads();
            for (int i = tid; i < V; i += kSampThreads) {
                const unsigned int b = __float_as_uint(p[i]);
                if ((b & mask) == (prefix & mask)) atomicAdd(&s_hist[(b >> shift) & 255], 1ULL);
            }
            __syncthreads();
            if (tid == 0) {
                unsigned long long cum = above;
                int d = 255;
                for (; d > 0; --d) {
                    if (cum + s_hist[d] >= (unsigned long long)R.top_k) break;
                    cum += s_hist[d];
                }
                s_prefix = prefix | ((unsigned int)d << shift);
                s_above = cum;
            }
            __syncthreads();
            prefix = s_prefix;
            above = s_above;
        }
        const float thr = __uint_as_float(prefix);
        for (int i = tid; i < V; i += kSampThreads)
            if (p[i] < thr) p[i] = 0.f;
        __syncthreads();
    }

    // ---- 4. top-p: nucleus boundary by radix descent over fixed-point masses
    if (R.top_p > 0.f && R.top_p < 1.f) {
        // masses are relative to the softmax total Z (the reference applies top-p to the un-renormalised top-k output)
        const double scale = 1099511627776.0 / Z;   // 2^40 / Z
        const unsigned long long target = (unsigned long long)((double)R.top_p * 1099511627776.0);
        unsigned int prefix = 0;
        unsigned long long above = 0;
        for (int pass = 0; pass < 4; ++pass) {
            const int shift = 24 - 8 * pass;
            const unsigned int mask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
            for (int i = tid; i < 256; i += kSampThreads) s_hist[i] = 0;
            __syncthreads();
            for (int i = tid; i < V; i += kSampThreads) {
                const unsigned int b = __float_as_uint(p[i]);
                if (b != 0 && (b & mask) == (prefix & mask))
                    atomicAdd(&s_hist[(b >> shift) & 255], (unsigned long long)((double)p[i] * scale));
            }
            __syncthreads();
            if (tid == 0) {
                unsigned long long cum = above;
                int d = 255;
                for (; d > 0; --d) {
                    if (cum + s_hist[d] > target) break;   // the boundary lies inside bin d
                    cum += s_hist[d];
                }
                s_prefix = prefix | ((unsigned int)d << shift);
                s_above = cum;
            }
            __syncthreads();
            prefix = s_prefix;
            above = s_above;
        }
        // values > boundary are kept; of the values == boundary keep as many as still fit (index order)
        const float bval = __uint_as_float(prefix);
        const unsigned long long each = (unsigned long long)((double)bval * scale);
        long long fit = (each > 0 && target > above) ? (long long)((target - above) / each) : 0;
        if (above == 0 && fit == 0) fit = 1;          // "always keep the first (largest) entry"
        // count the == boundary elements before each thread's range to keep exactly `fit` of them
        const int per = (V + kSampThreads - 1) / kSampThreads;
        const int lo = tid * per, hi = min(V, lo + per);
        int eq = 0;
        for (int i = lo; i < hi; ++i) eq += (__float_as_uint(p[i]) == prefix) ? 1 : 0;
        s_seg[tid] = (double)eq;
        __syncthreads();
        if (tid == 0) {
            double run = 0;
            for (int j = 0; j < kSampThreads; ++j) { const double c = s_seg[j]; s_seg[j] = run; run += c; }
        }
        __syncthreads();
        long long before = (long long)s_seg[tid];
        for (int i = lo; i < hi; ++i) {
            const unsigned int b = __float_as_uint(p[i]);
            if (b < prefix) p[i] = 0.f;
            else if (b == prefix) { if (before >= fit) p[i] = 0.f; ++before; }
        }
        __syncthreads();
    }

    // ---- 5. draw: first index whose inclusive prefix sum reaches u * total (index order)
    const int per = (V + kSampThreads - 1) / kSampThreads;
    const int lo = tid * per, hi = min(V, lo + per);
    double seg = 0.0;
    for (int i = lo; i < hi; ++i) seg += p[i];
    s_seg[tid] = seg;
    if (tid == 0) s_pick = -1;
    __syncthreads();
    if (tid == 0) {
        double run = 0;
        for (int j = 0; j < kSampThreads; ++j) { const double c = s_seg[j]; s_seg[j] = run; run += c; }
        s_redd[0] = run;
    }
    __syncthreads();
    const double total = s_redd[0];
    const double want = (double)u01 * total;
    const double start = s_seg[tid];
    if (want >= start && (want < start + seg || (tid == kSampThreads - 1))) {
        double run = start;
        int pick = -1;
        for (int i = lo; i < hi; ++i) {
            if (p[i] > 0.f) {
                run += p[i];
                pick = i;
                if (run > want) break;
            }
        }
        if (pick >= 0) s_pick = pick;
    }
    __syncthreads();
    if (tid == 0) {
        int pick = s_pick;
        if (pick < 0) {   // numerical corner (want == total): take the last kept element
            for (int i = V - 1; i >= 0; --i) if (p[i] > 0.f) { pick = i; break; }
        }
        pick = pick < 0 ? 0 : pick;
        out_ids[t] = pick;
        if (states) {     // the session's window of recent tokens and its random stream move on
            int n = R.n_recent;
            if (n >= kMaxRecent) {
                for (int j = 1; j < kMaxRecent; ++j) R.recent[j - 1] = R.recent[j];
                n = kMaxRecent - 1;
            }
            R.recent[n] = pick;
            R.n_recent = n + 1;
            states[ri].offset += 1;
        }
    }
}

}  // namespace b200

using namespace b200;

extern "C" int32_t b200_sample_row_bytes(void) { return (int32_t)sizeof(SampleRow); }

// rows: device array of T SampleRow records {temperature, top_p, penalty, top_k, n_recent, recent[50]};
// uniforms: T floats in [0,1); probs: T*V floats of scratch (holds the final unnormalised distribution on return).
extern "C" int b200_sample(const float* logits, int32_t ld, int32_t T, int32_t V, const void* rows, const float* uniforms,
                           float* probs, void* out_ids, int32_t pdl, void* stream_) {
    if (!logits || !rows || !uniforms || !probs || !out_ids) return fail(B200_EINVAL, "b200_sample: null pointer argument");
    if (T <= 0 || V <= 0) return fail(B200_EINVAL, "b200_sample: bad sizes T=%d V=%d", T, V);
    B200_CUDA_CHECK(launch(sample_kernel, dim3(T), dim3(kSampThreads), 0, static_cast<cudaStream_t>(stream_), pdl != 0,
                           logits, ld, V, static_cast<SampleRow*>(const_cast<void*>(rows)), uniforms, probs,
                           static_cast<long long*>(out_ids), static_cast<SampleState*>(nullptr),
                           static_cast<const int*>(nullptr)));
    return B200_OK;
}

extern "C" int32_t b200_sample_state_bytes(void) { return (int32_t)sizeof(SampleState); }

// Sampling with device-resident session state: rows / states are arrays indexed by row_index[t] (or t when row_index
// is NULL); each draw takes its uniform number from the session's Philox4x32-10 stream (seed, draws so far), appends
// the token to the row's recent-token window and advances the stream.  No host data per step: capturable in a graph.
extern "C" int b200_sample_stateful(const float* logits, int32_t ld, int32_t T, int32_t V, void* rows, void* states,
                                    const int32_t* row_index, float* probs, void* out_ids, int32_t pdl, void* stream_) {
    if (!logits || !rows || !states || !probs || !out_ids) return fail(B200_EINVAL, "b200_sample_stateful: null pointer argument");
    if (T <= 0 || V <= 0) return fail(B200_EINVAL, "b200_sample_stateful: bad sizes T=%d V=%d", T, V);
    B200_CUDA_CHECK(launch(sample_kernel, dim3(T), dim3(kSampThreads), 0, static_cast<cudaStream_t>(stream_), pdl != 0,
                           logits, ld, V, static_cast<SampleRow*>(rows), static_cast<const float*>(nullptr), probs,
                           static_cast<long long*>(out_ids), static_cast<SampleState*>(states), row_index));
    return B200_OK;
}
