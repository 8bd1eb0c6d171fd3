/*
 * b200_petals.h — C ABI of the sm_100a kernel library behind the pipeline stages.
 *
 * All entry points take raw device pointers and a CUDA stream (cudaStream_t cast
 * to void*), launch asynchronously on that stream, allocate nothing unless the
 * comment says so, and return 0 on success or a negative B200_E* code (the text
 * of the last error is available from b200_last_error()).
 *
 * Reference citations (jwkim-skku Mini-Petals, paths relative to the reference
 * root) name the torch op sequence each kernel replaces; the reference itself
 * contains no native code.
 */
#ifndef B200_PETALS_H_
#define B200_PETALS_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK 0
#define B200_EINVAL -1
#define B200_ECUDA -2
#define B200_ELIMIT -3

/* element types of activations / weights */
#define B200_F16 0
#define B200_BF16 1

const char* b200_last_error(void);
int b200_version(void);
/* Number of SMs of the current device (grid sizing for persistent kernels). */
int b200_sm_count(void);

/* ------------------------------------------------------------------ GEMM --
 * out[T,N] = epilogue( norm(x)[T,K] . W[N,K]^T )      (W = torch Linear weight)
 *
 * Swap-AB tcgen05 kernel: the weight is the 128-row M operand, streamed once
 * from HBM as pre-tiled, pre-swizzled 16 KiB blocks with bulk async copies;
 * tokens are the N operand (16..256 per tile), accumulators live in TMEM.
 * Work is split stream-K style over a persistent grid so that every SM streams
 * the same number of weight bytes whatever N and K are; partial tiles meet in
 * `scratch` and are summed in a fixed order (bit-reproducible).
 *
 * Opt-in fp8 weights (fp8 = 1): e4m3 weights with one scale per output row, activations quantised
 * per token on the fly, tcgen05.mma kind::f8f6f4, scales applied in the epilogue — half the weight
 * bytes per token (stands in for the reference's dead bitsandbytes NF4/INT8 path,
 * petals/server/server.py:88,189-191).
 *
 * Replaces q/k/v_proj, o_proj, gate/up/down_proj, lm_head and their
 * surrounding norm / activation / residual ops:
 * petals/llama/block.py:88-90,151,208-213,227-238; src/llama_partition.py:466-470.
 */
#define B200_EPI_NONE 0
#define B200_EPI_GELU_TANH 1 /* GPT-2 c_fc */
#define B200_EPI_SWIGLU 2    /* requires dual-packed gate/up weight */

#define B200_NORM_NONE 0
#define B200_NORM_RMS 1
#define B200_NORM_LAYER 2

/* Flag-chained decode step ("LL chain").  Between two kernel launches the GPU needs 3-4 us to notice that the
 * first grid is complete and to release the second one's dependency wait - five times per decoder block, with
 * the HBM idle.  With a chain the kernels of a decode step hand their activation vectors over as "LL words"
 * (two 16-bit values + a 32-bit tag in one 8-byte store, the scheme of the stream-K and tensor-parallel
 * exchanges): a consumer launched with programmatic stream serialization does NOT wait for the producer grid, it
 * polls the words it needs until they carry the tag (step number << 12 | id of the producing launch).  Words are
 * self-validating, so no fence, flag or grid dependency is needed, and a consumer starts on the first values while
 * the producer is still finishing others.  The caller numbers the producing launches of a step (ids 1..4095, unique
 * within the step), advances *seq once per step BEFORE the step's first launch (stream order), launches every kernel
 * of the step with pdl = 1 into one stream, and ends the step with an ordinary (grid-dependent) kernel.
 * LL buffers: [T][ld] 8-byte words, word i of a row = elements 2i, 2i+1; zero-initialised once.
 * Replaces nothing in the reference (its blocks are eager torch ops, petals/llama/block.py:200-240); it removes the
 * launch-boundary cost that CUDA graphs alone (petals/llama/cuda_graphs.py) leave in place. */
typedef struct b200_ll_chain {
    const int32_t* seq;   /* device: the step number (> 0); NULL = chain off, everything below ignored */
    const void* x_ll;     /* input vector as LL words (the kernel then never reads args->x) */
    const void* res_ll;   /* residual as LL words, when args->residual is set */
    void* out_ll;         /* output as LL words (16-bit outputs; args->out is still written when non-NULL) */
    int32_t ldx_ll, ldr_ll, ldo_ll;   /* row strides in words */
    int32_t x_op, res_op, out_op;     /* launch ids: who produced x / the residual, and this launch's own id */
    int32_t half;         /* 0/1: half of ll_buf for this launch's stream-K exchange; alternate between consecutive GEMMs */
    int32_t reserved;
} b200_ll_chain;

typedef struct b200_gemm_args {
    const void* w_packed; /* b200_pack layout, see ops/packing.py */
    const void* x;        /* [T, K] 16-bit, row stride ldx elements */
    void* out;            /* [T, N] 16-bit (or fp32 when out_fp32), row stride ldo */
    const void* residual; /* [T, N] 16-bit added after the epilogue, or NULL */
    const void* bias;     /* [N] 16-bit, or NULL */
    const void* norm_w;   /* [K] 16-bit norm gain (norm_mode != 0) */
    const void* norm_b;   /* [K] 16-bit LayerNorm beta, or NULL */
    float* scratch;       /* >= b200_gemm_scratch_bytes() bytes */
    int32_t* counters;    /* >= n_tiles * t_tiles ints, all zero; left all zero */
    int32_t T, N, K;      /* logical sizes; K is the packed (64-padded) K */
    int32_t ldx, ldo, ldr;
    int32_t n_tiles;      /* packed 128-row tiles (dual: pairs) */
    int32_t dual;         /* 1 = gate/up interleaved per tile */
    int32_t epilogue;
    int32_t norm_mode;
    float norm_eps;
    int32_t dtype;        /* B200_F16 / B200_BF16 */
    int32_t out_fp32;
    int32_t grid;         /* 0 = one CTA per SM */
    int32_t pdl;          /* 1 = launch with programmatic stream serialization */
    int32_t debug;        /* 0 in production; experiment switches, see csrc/gemm.cu */
    int32_t fp8;          /* 1 = e4m3 weights with one scale per output row (pack_weight_fp8, needs w_scale);
                             2 = OCP-MX: e4m3 weights with one UE8M0 scale per 32 elements along k packed behind every
                             k-block (pack_weight_mx), activations block-quantised on the fly,
                             tcgen05.mma.kind::mxf8f6f4.block_scale with the scales in tensor memory.  T <= 32, K % 128 == 0 */
    const float* w_scale; /* fp8: per-output-row dequantisation scale, [n_tiles*128] (dual: gate rows then up rows) */
    /* Optional L2 warm-up for the kernels that follow in the stream (b200_gemm only): once a CTA has issued
     * its last weight load it asks the L2 for its share of up to two byte ranges (cp.async.bulk.prefetch.L2),
     * typically the head of the next linear's packed weights, so the HBM keeps streaming through this
     * kernel's fix-up / epilogue and the next kernel's launch.  NULL / 0 = none.  Ranges must be 16-byte
     * aligned device memory that stays valid until the stream has passed this kernel. */
    const void* prefetch[2];
    int64_t prefetch_bytes[2];
    /* Optional ticket-free fix-up for decode sizes (b200_gemm with T <= 16; b200_gemm_allreduce / b200_gemm_gather and
     * fp8 weights with T <= 4):
     * b200_gemm_ll_bytes() bytes of device memory, zeroed ONCE by the caller and then owned by the library
     * (one buffer per stream of GEMM launches; it may be shared by all GEMMs of that stream).  When every
     * output tile is split over the same whole number of CTAs, the CTAs publish their partial sums as
     * (value, tag) 8-byte words and the tile's last CTA polls for them and clears them again - no fence, no
     * atomic ticket, no counter (3-way split: 2.3 -> 0.4 us, 8-way: 5.9 -> 1.6 us per launch at T = 1; with 2..16 tokens every CTA of the
     * tile reduces a share of its elements).  Same summation
     * order as the ticket path, i.e. bit-identical results.  NULL = ticket path. */
    void* ll_buf;
    /* Optional with norm_mode == B200_NORM_RMS: per-token 1/rms computed beforehand (b200_rms_rstd, same bits as the
     * in-kernel statistics); the kernel then skips its own statistics pass.  Pays off from ~8 tokens per step. */
    const float* rstd;
    /* Optional flag chain (decode sizes, 16-bit weights, b200_gemm only): see b200_ll_chain. */
    b200_ll_chain chain;
} b200_gemm_args;

int b200_gemm(const b200_gemm_args* args, void* stream);
/* Upper bound of the scratch needed by any b200_gemm call with T <= max_tokens. */
int64_t b200_gemm_scratch_bytes(int32_t max_tokens);
int64_t b200_gemm_ll_bytes(void);
/* Token-tile width the kernel will pick for T tokens (16,32,64,128,256). */
int32_t b200_gemm_token_tile(int32_t T, int32_t dual);

/* Prefill-size variant (T >= 64, no norm prologue, 16-bit weights; same argument struct, scratch and
 * counters unused): weights by cp.async.bulk, activations by a 2-D TMA tensor map, two TMEM accumulator
 * buffers so the epilogue of a tile overlaps the main loop of the next. */
int b200_gemm_prefill(const b200_gemm_args* args, void* stream);

/* Row-parallel linear of a tensor-parallel span with the all-reduce fused into the tile epilogue
 * (decode sizes, T <= 16): every rank runs this with its K shard of the weight; the CTA that finishes
 * an output tile pushes its fp32 partial into every rank's staging buffer with peer stores over
 * NVLink, waits for the other ranks' partials of the same tile (flags in local memory), adds them in
 * rank order and applies bias / residual once.  All ranks end with bit-identical outputs; tiles are
 * exchanged while other tiles are still in the tensor core.  The reference's dead TP path runs an
 * NCCL all-reduce after each row-parallel linear (petals/server/backend.py:134-136 via tensor_parallel).
 * stage[r] / flags[r] are rank r's buffers mapped into this process (e.g. torch symmetric memory):
 *   stage: 2 * world * 16 * n_max floats,  flags: 2 * world * tiles_max ints (zeroed once),
 *   seq / exit_counter: one int each, local, zeroed once.  All ranks must issue the same sequence of
 *   b200_gemm_allreduce calls on a workspace. */
typedef struct b200_tp_args {
    int32_t world, rank, n_max, tiles_max;
    void* stage[8];
    void* flags[8];
    int32_t* seq;
    int32_t* exit_counter;
    int32_t mode;   /* 0: the producer waits and reduces (all-reduce in one kernel); 1: the producer only pushes
                       its partial tiles and the next GEMM (b200_gemm_gather) waits for them and adds them up */
} b200_tp_args;
int b200_gemm_allreduce(const b200_gemm_args* args, const b200_tp_args* tp, void* stream);
/* Consumer of a mode-1 producer ("all-gather -> GEMM"): every CTA first rebuilds
 * x[t,:] = residual[t,:] + bias + sum over ranks of the pushed partial rows (rank order, bit-identical on
 * all ranks), stores it to args->x, then runs b200_gemm on it.  The NVLink latency of the exchange hides
 * behind the launch of this kernel instead of sitting at the producer's tail. */
int b200_gemm_gather(const b200_gemm_args* args, const b200_tp_args* tp, const void* residual, int32_t ldr,
                     const void* bias, void* stream);

/* Row-parallel linear of a tensor-parallel span at PREFILL sizes: all-reduce of the ranks' partial outputs through
 * the NVSwitch (NVLS) with the linear's epilogue in the same kernel: out = sum_r partial_r + bias + residual on every
 * rank.  Each rank reads its slice of the elements through the multicast address with multimem.ld_reduce (the switch
 * adds the ranks' values, fp32 accumulation), applies bias / residual, and writes the slice through the multicast
 * address of the output with multimem.st (replicated into every rank's memory): reduce-scatter by load, all-gather by
 * store, no rank polls its peers in turn.  The reference's dead TP path: NCCL all-reduce + separate torch adds
 * (petals/server/backend.py:134-136).
 *   mc_partial / mc_out  MULTICAST addresses (e.g. torch symmetric memory, handle.multicast_ptr + offset) of two
 *                        symmetric buffers; every rank's GEMM wrote its [T, N] partial into its own copy of the first.
 *   flags_local / flags_peer[r]  16 ints of symmetric memory per rank (zeroed once): this rank's copy and every rank's.
 *   seq, cta_done        one int each, local, zeroed once.  All ranks must issue the same sequence of calls. */
typedef struct b200_nvls_args {
    const void* mc_partial;
    void* mc_out;
    const void* residual;  /* local [T, ldr] 16-bit, or NULL */
    const void* bias;      /* [N] 16-bit, or NULL */
    int32_t T, N, ldp, ldo, ldr, dtype;
    int32_t world, rank;
    int32_t* flags_local;
    void* flags_peer[8];
    int32_t* seq;
    int32_t* cta_done;
    int32_t pdl, reserved;
} b200_nvls_args;
int b200_tp_allreduce_nvls(const b200_nvls_args* args, void* stream);

/* b200_gemm for callers that cannot allocate device memory (no counterpart in the reference, which
 * has no native code and lets torch allocate freely).  Same arguments and result as b200_gemm, but
 * args->scratch and args->counters are ignored: the workspace is a fixed ~5 MiB array inside the
 * library image, resolved per device on the first call made on that device (make that call before
 * starting a stream capture).  Limits: T <= 16 tokens, <= 2048 output tiles, <= 160 CTAs; beyond
 * them the call returns B200_ELIMIT.  Calls share the one array: keep them on one stream — kernels
 * from two calls that overlap on the GPU race on it.  args->pdl and args->ll_buf are ignored.
 * Performs no cudaMalloc/cudaFree. */
#define B200_GEMM_STATIC_MAX_TOKENS 16
int b200_gemm_static_scratch(const b200_gemm_args* args, void* stream);

/* ------------------------------------------------------------- attention --
 * Decode: for each of T independent sessions' new tokens, apply RoPE to q and k,
 * append k/v to that session's pre-allocated cache in place (no torch.cat
 * reallocation as in petals/llama/block.py:123-128), and attend over the cache
 * with a split-KV online softmax (GQA heads of one kv head share the K/V
 * stream; no repeat_kv copy, block.py:131-132).  Scores never touch HBM
 * (block.py:134-141 materialises [nh,T,S] in fp32).
 * kv_ptrs[t] = {K base, V base} of this layer for token t's session, each
 * laid out [n_kv_heads][s_max][head_dim]; pos[t] = tokens already cached.
 */
typedef struct b200_attn_decode_args {
    const void* qkv;        /* [T, (nh + 2 nkv) * d] : q | k | v, row stride ldqkv */
    void* out;              /* [T, nh * d], row stride ldo */
    const uint64_t* kv_ptrs;/* device [T][2] */
    const int32_t* pos;     /* device [T] */
    const float* inv_freq;  /* device [d/2] RoPE inverse frequencies, NULL = no RoPE (GPT-2) */
    float* scratch;         /* >= b200_attn_decode_scratch_bytes() */
    int32_t* counters;      /* T * n_kv_heads ints, zero on entry and on exit */
    int32_t T, n_heads, n_kv_heads, head_dim, s_max, splits, ldqkv, ldo;
    float scale;            /* 1/sqrt(d) */
    int32_t dtype, pdl;
    /* Optional: 2 * b200_attn_decode_scratch_bytes() bytes, zeroed ONCE by the caller, then owned by the library
     * (one per stream).  The KV splits of a head then publish their partial (o, m, l) as (value, tag) 8-byte
     * words and the last split polls, merges and clears them: no fence / atomic ticket in the merge.  Same
     * arithmetic and order as the ticket path.  NULL = ticket path (scratch + counters). */
    void* ll_scratch;
    /* Optional flag chain (see b200_ll_chain): qkv is read from chain.x_ll (tag of chain.x_op) instead of waiting for
     * the producer grid, the output also goes to chain.out_ll with chain.out_op; needs ll_scratch.  res_* unused. */
    b200_ll_chain chain;
} b200_attn_decode_args;
int b200_attn_decode(const b200_attn_decode_args* args, void* stream);
int64_t b200_attn_decode_scratch_bytes(int32_t T, int32_t n_heads, int32_t head_dim, int32_t splits);

/* Prefill / chunk: rotate q in place, rotate k and store k, v into the session cache at
 * positions [pos0, pos0 + T).  (HF rotary + cache append, block.py:96-128.) */
int b200_rope_append(void* qkv, void* kcache, void* vcache, const float* inv_freq, int32_t T, int32_t n_heads,
                     int32_t n_kv_heads, int32_t head_dim, int32_t s_max, int32_t ldqkv, int32_t pos0,
                     int32_t dtype, int32_t pdl, void* stream);
/* Causal attention of a T-token chunk (already appended) against cache rows [0, pos0 + T).
 * The reference never builds the causal mask on its live path; this does. */
int b200_attn_prefill(const void* q, void* out, const void* kcache, const void* vcache, int32_t T, int32_t n_heads,
                      int32_t n_kv_heads, int32_t head_dim, int32_t s_max, int32_t ldq, int32_t ldo, int32_t pos0,
                      float scale, int32_t dtype, int32_t pdl, void* stream);

/* Same contract on the tensor cores: 128-query x 128-key tiles, S = QK^T and O = PV by tcgen05.mma
 * (V consumed in place as an MN-major operand), online softmax between them, accumulators in TMEM. */
int b200_attn_prefill_tc(const void* q, void* out, const void* kcache, const void* vcache, int32_t T, int32_t n_heads,
                         int32_t n_kv_heads, int32_t head_dim, int32_t s_max, int32_t ldq, int32_t ldo, int32_t pos0,
                         float scale, int32_t dtype, int32_t pdl, void* stream);

/* ----------------------------------------------------------- elementwise --
 * Embedding gather (+ learned positions for GPT-2): src/llama_partition.py:107.
 * ids are int64 on the device (the last stage writes them there). */
int b200_embed(const void* ids, const void* pos, int32_t pos0, const void* wte, const void* wpe, void* out, int32_t T,
               int32_t H, int32_t vocab, int32_t max_pos, int32_t dtype, int32_t pdl, void* stream);
/* Head of a flag chain: the 16-bit rows x[T, H] (already complete in stream order) re-published as LL words tagged
 * (*seq << 12 | op), so that the chain's first GEMM can poll them like any other link of the chain. */
int b200_ll_publish(const void* x, int32_t ldx, void* out_ll, int32_t ld_ll, int32_t T, int32_t H, const int32_t* seq,
                    int32_t op, int32_t pdl, void* stream);
/* Standalone RMSNorm (layer_norm=0) / LayerNorm (1) for prefill-size inputs: block.py:208-213. */
/* rstd[t] = 1/sqrt(mean(x[t]^2) + eps), one block per token; see b200_gemm_args.rstd. */
int b200_rms_rstd(const void* x, float* rstd, int32_t T, int32_t K, int32_t ldx, float eps, int32_t dtype, int32_t pdl,
                  void* stream);
int b200_norm(const void* x, const void* w, const void* b, void* out, int32_t T, int32_t H, int32_t ldx, int32_t ldo,
              float eps, int32_t layer_norm, int32_t dtype, int32_t pdl, void* stream);
/* Greedy sampling: first index of the row maximum of fp32 logits (src/rpc_handler.py:331-335).
 * part_val/part_idx: T*chunks scratch; counters: T ints kept zero. */
int b200_argmax(const float* logits, int32_t ld, int32_t T, int32_t V, float* part_val, int32_t* part_idx,
                int32_t* counters, void* out_ids, int32_t chunks, int32_t pdl, void* stream);

/* Arg-max over one SHARD of the vocabulary (vocab-sharded LM head): indices are reported with `idx_offset` added,
 * `init_pair` (T x {float bits, int32 index}, or NULL) is a candidate from another shard that the shard's own
 * maximum has to beat (ties go to the lower index, as torch.argmax would on the whole row), and `out_pair`
 * (T x 2 words, or NULL) receives the winning (value, index) so that the next shard can continue.  out_ids may be NULL. */
int b200_argmax_partial(const float* logits, int32_t ld, int32_t T, int32_t V, float* part_val, int32_t* part_idx,
                        int32_t* counters, void* out_ids, int32_t chunks, int32_t idx_offset, const void* init_pair,
                        void* out_pair, int32_t pdl, void* stream);

/* ---------------------------------------------------------- pipeline hop --
 * Hidden-state handoff between two pipeline stages on different GPUs of one NVSwitch box, made of two small
 * kernels that are captured into the stages' CUDA graphs: b200_hop_send (last node of the producer's graph) writes
 * the micro-batch into a ring in the RECEIVER's memory with peer stores over NVLink, each 4-byte payload paired
 * with the message number in one 8-byte word; b200_hop_recv (first node of the consumer's graph) polls its own
 * memory until the whole message carries the expected number and unpacks it.  No host, no flag, no fence, no
 * library call on the per-token path.  Replaces the reference's hop through the client:
 * src/rpc_transport.py:736-766 (GPU -> CPU -> serialize -> libp2p -> deserialize -> GPU, twice per stage).
 *   ring     b200_hop_ring_bytes(depth, t_max, pairs_per_token) bytes in the receiver's memory, zeroed once; the
 *            sender passes the peer-mapped address (e.g. torch symmetric memory), the receiver its local one.
 *            Message n uses entry n % depth: depth must exceed the number of messages the link can have in flight.
 *   counter  one int per link END (sender and receiver each own one, both start at 0, local memory).
 *   x        [T, hidden] 16-bit rows (send: source, recv: destination), row stride ldx elements; hidden may be 0.
 *   extra    [T, n_extra] 32-bit words that travel with each token (token ids, arg-max partials), or NULL.
 * A receive that sees nothing for timeout_ms (0 = 30 s) prints a message and traps instead of hanging the GPU. */
typedef struct b200_hop_args {
    void* x;
    void* extra;
    void* ring;
    int32_t* counter;
    int32_t T, hidden, ldx, n_extra;
    int32_t depth, t_max, pairs_per_token;
    int32_t timeout_ms, pdl;
} b200_hop_args;
int64_t b200_hop_ring_bytes(int32_t depth, int32_t t_max, int32_t pairs_per_token);
int b200_hop_send(const b200_hop_args* args, void* stream);
int b200_hop_recv(const b200_hop_args* args, void* stream);
/* Loads the two kernels on the current device (call once before capturing them into a graph). */
int b200_hop_init(void);

/* Opt-in device-side event trace of the decode kernels (profiling aid, scripts/trace_decode.py): the GEMM and the
 * decode attention stamp %globaltimer at entry / after their dependency wait / first accumulator / exit for the first
 * and last CTA.  buf: 8 * (1 + 2 * cap) bytes of zeroed device memory ([0] = events written, then (ns, tag) pairs with
 * tag = kernel << 56 | aux << 32 | phase << 16 | last_cta; aux of a GEMM = n_tiles | k_blocks << 12); NULL = off.  Synchronous (cudaMemcpyToSymbol); the reference's only
 * timing aid is time.time() around forward calls (src/rpc_handler.py:190-249). */
int b200_trace_set(void* buf, int32_t cap);

/* Temperature / top-k / top-p / repetition-penalty sampling on the device with the reference's rules
 * (src/rpc_handler.py:327-403), one CTA per row, no sort (radix descent over float bit patterns).
 * rows: T records {float temperature, top_p, penalty; int32 top_k, n_recent; int32 recent[50]} on the
 * device (b200_sample_row_bytes() each); uniforms: T floats in [0,1) (e.g. torch.rand on the device,
 * philox); probs: T*V floats of scratch, holds the final un-normalised distribution on return. */
int32_t b200_sample_row_bytes(void);
int b200_sample(const float* logits, int32_t ld, int32_t T, int32_t V, const void* rows, const float* uniforms,
                float* probs, void* out_ids, int32_t pdl, void* stream);
/* The same with the sessions' sampling state on the device: rows[row_index[t]] (row_index NULL: rows[t]) holds the
 * parameters and the window of recent tokens, states[..] = {uint64 seed, uint64 draws so far}
 * (b200_sample_state_bytes() each).  The uniform number of a draw is the first word of
 * Philox4x32-10(key = seed, counter = draws so far); after the draw the kernel appends the token to the window and
 * advances the stream.  Nothing comes from the host per step, so the call can sit in a captured graph, and a session's
 * tokens depend only on its seed (the reference draws from torch's global generator on whichever server samples,
 * src/rpc_handler.py:396-403). */
int32_t b200_sample_state_bytes(void);
int b200_sample_stateful(const float* logits, int32_t ld, int32_t T, int32_t V, void* rows, void* states,
                         const int32_t* row_index, float* probs, void* out_ids, int32_t pdl, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_PETALS_H_ */
