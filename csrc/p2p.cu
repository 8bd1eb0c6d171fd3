// Stage-to-stage handoff inside one NVSwitch box without a communication library on the per-token path.
//
// A pipeline stage's CUDA graph ends with hop_send_kernel and the next stage's graph begins with
// hop_recv_kernel.  The sender writes the micro-batch's hidden state straight into a ring in the RECEIVER's
// memory with peer stores over NVLink; every 4-byte payload travels with the message number in one 8-byte
// word (the "LL" scheme of the tensor-parallel exchange in gemm_core.cuh), so the receiver needs no flag, no
// fence and no host: it polls its own memory until every word of the message carries the expected number.
// Message n of a link lives in ring entry n % depth; the pipeline never has more than `depth` messages of a
// link in flight (a session's next step needs its previous token, which needs the whole chain).
//
// Replaces the reference's per-hop path GPU -> CPU -> serialize -> libp2p -> deserialize -> GPU through the
// client (src/rpc_transport.py:736-766, src/rpc_handler.py:139-147).
#include "common.h"
#include "ptx.cuh"

namespace b200 {

struct HopParams {
    const uint32_t* x;      // send: [T][ldx32] 32-bit pairs of 16-bit elements
    uint32_t* x_out;        // recv
    int ldx32;
    const uint32_t* extra;  // send: [T][n_extra] 32-bit words (token ids, arg-max partials), or null
    uint32_t* extra_out;    // recv
    int n_extra;
    uint4* ring;            // ring base in the receiver's memory (peer-mapped on the sender)
    int depth, t_max, pairs_per_token;   // entry = t_max tokens x pairs_per_token 16-byte cells
    int T, W;               // tokens of this message, 32-bit hidden words per token (hidden / 2), W even
    int* ctr;               // local: messages sent / received on this link so far
    long long timeout_ns;
};

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t));
    return t;
}

__global__ void __launch_bounds__(1024) hop_send_kernel(const HopParams p) {
    __shared__ int s_seq;
    pdl_wait_prior_grid();   // x comes from the previous kernel of the stream
    pdl_launch_dependents();
    if (threadIdx.x == 0) s_seq = *p.ctr + 1;
    __syncthreads();
    const int seq = s_seq;
    const int ph = p.W >> 1, pe = (p.n_extra + 1) >> 1, P = ph + pe;
    uint4* entry = p.ring + (size_t)(seq % p.depth) * p.t_max * p.pairs_per_token;
    for (int idx = threadIdx.x; idx < p.T * P; idx += blockDim.x) {
        const int t = idx / P, c = idx - t * P;
        uint32_t a, b = 0;
        if (c < ph) {
            const uint2 v = *reinterpret_cast<const uint2*>(p.x + (size_t)t * p.ldx32 + 2 * c);
            a = v.x; b = v.y;
        } else {
            const int q = 2 * (c - ph);
            a = p.extra[(size_t)t * p.n_extra + q];
            if (q + 1 < p.n_extra) b = p.extra[(size_t)t * p.n_extra + q + 1];
        }
        asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %2};\n" ::"l"(entry + (size_t)t * p.pairs_per_token + c),
                     "r"(a), "r"(seq), "r"(b) : "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) *p.ctr = seq;
}

__global__ void __launch_bounds__(1024) hop_recv_kernel(const HopParams p) {
    __shared__ int s_seq;
    // Dependents may start at once: the next GEMM streams its weights while this kernel still polls
    // (it waits for this grid's completion before it touches activations).
    pdl_launch_dependents();
    pdl_wait_prior_grid();   // x_out may still be read by the previous kernel of the stream
    if (threadIdx.x == 0) s_seq = *p.ctr + 1;
    __syncthreads();
    const int seq = s_seq;
    const int ph = p.W >> 1, pe = (p.n_extra + 1) >> 1, P = ph + pe;
    const uint4* entry = p.ring + (size_t)(seq % p.depth) * p.t_max * p.pairs_per_token;
    unsigned long long t0 = 0;
    for (int idx = threadIdx.x; idx < p.T * P; idx += blockDim.x) {
        const int t = idx / P, c = idx - t * P;
        const uint4* src = entry + (size_t)t * p.pairs_per_token + c;
        uint32_t a, sa, b, sb;
        for (;;) {
            asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];\n"
                         : "=r"(a), "=r"(sa), "=r"(b), "=r"(sb) : "l"(src) : "memory");
            if ((int)sa == seq && (int)sb == seq) break;
            if (t0 == 0) t0 = globaltimer_ns();
            else if ((long long)(globaltimer_ns() - t0) > p.timeout_ns) {
                printf("b200: pipeline hop timed out waiting for message %d (thread %d saw %d/%d)\n", seq, threadIdx.x,
                       (int)sa, (int)sb);
                __trap();
            }
        }
        if (c < ph) {
            *reinterpret_cast<uint2*>(p.x_out + (size_t)t * p.ldx32 + 2 * c) = make_uint2(a, b);
        } else {
            const int q = 2 * (c - ph);
            p.extra_out[(size_t)t * p.n_extra + q] = a;
            if (q + 1 < p.n_extra) p.extra_out[(size_t)t * p.n_extra + q + 1] = b;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) *p.ctr = seq;
}

static int check_hop(const b200_hop_args* a, const char* who) {
    if (!a || !a->ring || !a->counter) return fail(B200_EINVAL, "%s: null ring / counter", who);
    if (a->T <= 0 || a->T > a->t_max) return fail(B200_EINVAL, "%s: T=%d outside [1, t_max=%d]", who, a->T, a->t_max);
    if (a->hidden < 0 || a->hidden % 4 != 0) return fail(B200_EINVAL, "%s: hidden=%d must be a multiple of 4", who, a->hidden);
    if (a->n_extra < 0 || a->depth < 2) return fail(B200_EINVAL, "%s: bad n_extra=%d / depth=%d", who, a->n_extra, a->depth);
    if (a->hidden / 4 + (a->n_extra + 1) / 2 > a->pairs_per_token)
        return fail(B200_ELIMIT, "%s: %d hidden + %d extra words exceed the ring's %d cells per token", who, a->hidden,
                    a->n_extra, a->pairs_per_token);
    if (a->hidden > 0 && (!a->x || a->ldx % 2 != 0 || (reinterpret_cast<uintptr_t>(a->x) & 7) != 0))
        return fail(B200_EINVAL, "%s: x must be 8-byte aligned with an even row stride", who);
    if (a->n_extra > 0 && !a->extra) return fail(B200_EINVAL, "%s: n_extra set without extra", who);
    return B200_OK;
}

static HopParams hop_params(const b200_hop_args* a) {
    HopParams p;
    p.x = static_cast<const uint32_t*>(a->x);
    p.x_out = static_cast<uint32_t*>(a->x);
    p.ldx32 = a->ldx / 2;
    p.extra = static_cast<const uint32_t*>(a->extra);
    p.extra_out = static_cast<uint32_t*>(a->extra);
    p.n_extra = a->n_extra;
    p.ring = static_cast<uint4*>(a->ring);
    p.depth = a->depth; p.t_max = a->t_max; p.pairs_per_token = a->pairs_per_token;
    p.T = a->T; p.W = a->hidden / 2;
    p.ctr = a->counter;
    p.timeout_ns = (a->timeout_ms > 0 ? (long long)a->timeout_ms : 30000LL) * 1000000LL;
    return p;
}

}  // namespace b200

using namespace b200;

extern "C" int64_t b200_hop_ring_bytes(int32_t depth, int32_t t_max, int32_t pairs_per_token) {
    return (int64_t)depth * t_max * pairs_per_token * 16;
}

extern "C" int b200_hop_send(const b200_hop_args* a, void* stream_) {
    const int rc = check_hop(a, "b200_hop_send");
    if (rc != B200_OK) return rc;
    B200_CUDA_CHECK(launch(hop_send_kernel, dim3(1), dim3(1024), 0, static_cast<cudaStream_t>(stream_), a->pdl != 0,
                           hop_params(a)));
    return B200_OK;
}

extern "C" int b200_hop_recv(const b200_hop_args* a, void* stream_) {
    const int rc = check_hop(a, "b200_hop_recv");
    if (rc != B200_OK) return rc;
    B200_CUDA_CHECK(launch(hop_recv_kernel, dim3(1), dim3(1024), 0, static_cast<cudaStream_t>(stream_), a->pdl != 0,
                           hop_params(a)));
    return B200_OK;
}

// Loads both kernels' code on the current device, so that their first launch may happen inside a stream capture.
extern "C" int b200_hop_init(void) {
    cudaFuncAttributes fa;
    B200_CUDA_CHECK(cudaFuncGetAttributes(&fa, hop_send_kernel));
    B200_CUDA_CHECK(cudaFuncGetAttributes(&fa, hop_recv_kernel));
    return B200_OK;
}
