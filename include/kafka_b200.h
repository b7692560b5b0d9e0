/* kafka_b200.h — C ABI of libkafka_b200.so, the sm_100a completion engine that sits under the
 * reference's LLMProvider plugin API.
 *
 * What this boundary replaces.  In egrigokhan/kafka-llm-service the provider plugin is a Python ABC
 * (src/llm/base.py:67 LLMProvider; stream_completion src/llm/base.py:166-219; completion
 * src/llm/base.py:222-262) whose only implementation hands the arithmetic to a remote gateway
 * (src/llm/portkey.py:379 and :616, `await client.chat.completions.create(**params)`).  There is no
 * FFI in the reference; the functions below are what a local provider binds (via ctypes, see
 * INTEGRATION.md) in place of that network hop.  Each entry point cites the reference call site whose
 * work it takes over.
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / C++ types.
 *   - every device pointer and every stream is allocated by the caller (PyTorch in this repo) and
 *     BORROWED; the library allocates nothing on the device.
 *   - every call only enqueues work on the given stream (no hidden synchronisation).
 *   - return 0 on success; non-zero on failure, message via kb_last_error() (thread-local).
 *   - one engine per device, driven by one host thread at a time.
 *   - there is no CPU fallback in this library.
 */
#ifndef KAFKA_B200_H
#define KAFKA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KB_ABI_VERSION 2
#define KB_BLOCK_TOKENS 64 /* tokens per KV block (one TMA bulk tile = 64 tokens x head_dim) */
#define KB_MAX_SPLITS 16   /* upper bound on split-KV partitions of one sequence in decode attention */

#define KB_OK 0
#define KB_ERR_INVALID 1
#define KB_ERR_CUDA 2

typedef void* kb_stream; /* a cudaStream_t */

typedef struct kb_engine kb_engine;

/* Llama-family architecture (the shapes SURVEY.md §8d lists; nothing like it exists in the reference,
 * whose `model` is just a string forwarded to the gateway, src/llm/portkey.py:267). */
typedef struct kb_model_cfg {
    int32_t hidden;     /* model width, multiple of 64 */
    int32_t n_layers;
    int32_t n_heads;    /* query heads */
    int32_t n_kv_heads; /* n_heads / n_kv_heads in {1,2,4,8,16} */
    int32_t head_dim;   /* 64 or 128 */
    int32_t ffn;        /* MLP inner width, multiple of 64 */
    int32_t vocab;      /* multiple of 8 */
    int32_t max_pos;    /* context limit; beyond it the host raises a context-length error */
    float rope_theta;
    float rms_eps;
} kb_model_cfg;

/* Device pointers, row-major [out_features, in_features] like torch.nn.Linear.weight.  The five GEMM
 * operands (wqkv, wo, w_gate_up, w_down, lm_head) are FP16 (see KB_GEMM_F16); norms and embed are bf16.
 *   wqkv      : [(n_heads + 2 n_kv_heads) * head_dim, hidden]   rows = q heads, k heads, v heads
 *   wo        : [hidden, n_heads * head_dim]
 *   w_gate_up : [2 * ffn, hidden]   row 2j = gate_j, row 2j+1 = up_j (interleaved for the SwiGLU epilogue)
 *   w_down    : [hidden, ffn]
 */
typedef struct kb_layer_weights {
    const void* attn_norm; /* [hidden] bf16 */
    const void* wqkv;
    const void* wo;
    const void* mlp_norm;  /* [hidden] bf16 */
    const void* w_gate_up;
    const void* w_down;
} kb_layer_weights;

typedef struct kb_weight_ptrs {
    const void* embed;      /* [vocab, hidden] bf16 */
    const void* final_norm; /* [hidden] bf16 */
    const void* lm_head;    /* [vocab, hidden] fp16 */
    const kb_layer_weights* layers; /* host array of n_layers entries */
} kb_weight_ptrs;

typedef struct kb_sampling {
    float temperature; /* <= 0 : greedy argmax (ties -> lowest id) */
    uint64_t seed;     /* Gumbel-max sampling stream when temperature > 0 */
    uint64_t step;     /* caller-incremented counter mixed into the stream */
    int32_t banned_token; /* token id never produced (e.g. EOS masked while benchmarking), -1 = none */
    /* Optional restriction of the candidate set to ids in [allowed_lo, allowed_hi) plus up to two
     * extra ids (e.g. newline and <|eot_id|>); allowed_hi <= allowed_lo means "whole vocabulary". */
    int32_t allowed_lo, allowed_hi;
    int32_t extra_a, extra_b; /* -1 = unused */
    /* Optional per-sequence overrides for a continuous batch whose requests differ (device arrays of
     * n_seqs entries, NULL = use the scalars above): temperature and a seed XOR-ed into `seed`. */
    const float* seq_temperature;
    const uint64_t* seq_seed;
    /* Nucleus sampling (the `top_p` the reference's provider API documents, src/llm/portkey.py:246): with 0 < top_p < 1
     * and temperature > 0 only the most probable candidates whose probabilities (softmax of logit / T over the candidate
     * set) add up to top_p are sampled from. 0 or >= 1 = off. seq_top_p: optional per-sequence device array. Needs the
     * whole vocabulary on one GPU (refused with a vocabulary-sharded lm_head). */
    float top_p;
    const float* seq_top_p;
} kb_sampling;

/* ---- engine lifetime ------------------------------------------------------------------------- */

/* Takes over: provider construction, src/kafka/v1.py:177-181 (PortkeyLLMProvider(...)). */
int kb_engine_create(int device, const kb_model_cfg* cfg, kb_engine** out);
void kb_engine_destroy(kb_engine* e);

/* Takes over: nothing in the reference (weights live behind the gateway). Pointers are borrowed. */
int kb_bind_weights(kb_engine* e, const kb_weight_ptrs* w);

/* KV pools, layout [layer][block][kv_head][dim plane = dim / 64][token_in_block][64 dims] FP16 (values
 * beyond +-65504 are stored saturated, see KB_GEMM_F16) with the
 * 16-byte chunks of each 128-byte row XOR-swizzled by (token & 7) — every plane of a (block, head) tile is
 * a canonical SWIZZLE_128B K-major matrix.  k_pool and v_pool each hold
 * n_layers * n_blocks * n_kv_heads * KB_BLOCK_TOKENS * head_dim fp16.
 * Takes over: the thread history the reference re-sends in full on every call
 * (src/agents/base.py:222-223) — here it stays resident in HBM. */
int kb_bind_kv_pool(kb_engine* e, void* k_pool, void* v_pool, int32_t n_blocks);

/* Activation workspace for up to max_tokens packed tokens / max_seqs sequences per call. */
size_t kb_workspace_bytes(const kb_model_cfg* cfg, int32_t max_tokens, int32_t max_seqs);
int kb_bind_workspace(kb_engine* e, void* ws, size_t bytes, int32_t max_tokens, int32_t max_seqs);

/* ---- tensor parallelism (BASELINE config 5: Llama-3-70B over 8 GPUs) ----------------------------
 * Megatron-style sharding, one process and one engine per GPU: the engine is created with the LOCAL
 * shapes of its rank (n_heads / tp, n_kv_heads / tp, ffn / tp; hidden and vocab unchanged), wqkv and
 * w_gate_up hold this rank's rows, wo and w_down this rank's K-columns, embed / norms / lm_head are
 * replicated.  The two row-sharded projections of every layer produce partial sums that must be added
 * over the ranks; the engine calls `fn(ctx, buf, n, stream)` to do that in place on `buf` (n fp32,
 * device) in stream order.  Rank 0 has already folded the residual into its partial, so the reduced
 * buffer is the new residual stream on every rank.  In this repo the callback is NCCL over NVLink
 * through torch.distributed (kafka_b200/tp.py); no counterpart exists in the reference.
 * Every rank must make the same kb_prefill / kb_decode calls with the same arguments. */
typedef int (*kb_allreduce_fn)(void* ctx, float* buf, size_t n, kb_stream stream);
int kb_set_tensor_parallel(kb_engine* e, int32_t tp_rank, int32_t tp_size, kb_allreduce_fn fn, void* ctx);

/* The same sums WITHOUT a library collective: one kernel per all-reduce that works directly on peer
 * memory over NVLink / NVSwitch (csrc/kb_collective.cu).  The caller provides a SYMMETRIC region — the
 * same allocation size on every rank, mapped into every rank's address space (CUDA IPC / fabric
 * handles; kafka_b200/tp.py uses torch's symmetric-memory rendezvous for the mapping only):
 *     peer_base[r]   this process's mapping of rank r's region (peer_base[own rank] = the local one)
 *     mc_base        multicast (NVLS) mapping of the region, or NULL -> plain peer loads and stores
 *     flags_offset   byte offset of KB_TP_FLAG_BYTES of barrier flags, zeroed on every rank before the
 *                    first call (and a host barrier after zeroing)
 * kb_set_tensor_parallel_peers moves the engine's residual stream and partial-sum buffer (2 x
 * align256(max_tokens * hidden * 4) bytes from offset 0) into the region and replaces the callback;
 * call it after kb_bind_workspace, on every rank.  kb_allreduce_peers is the bare collective (in
 * place, n fp32 at offset_bytes; `epoch` must be odd, equal on all ranks and grow by 2 per call).
 * A rank whose peers do not arrive within the region's timeout (default 30 s) traps instead of hanging.
 * No reference counterpart. */
#define KB_TP_MAX_RANKS 8
#define KB_TP_MAX_BLOCKS 128 /* grid bound of the collectives: all blocks must be co-resident (148 SMs) */
#define KB_TP_FLAG_BYTES 2048
typedef struct kb_tp_peers {
    void* peer_base[KB_TP_MAX_RANKS];
    void* mc_base;
    size_t bytes;
    size_t flags_offset;
    /* HOST word that holds the next barrier epoch of this region (odd, starts at 1, +2 per collective), shared by
     * everything that launches collectives on the region — engines attached to it and direct kb_allreduce_peers
     * callers — so that an epoch is never reused however often engines are re-attached. NULL: the engine keeps a
     * private counter (one engine per region only). Never dereferenced on the device. */
    uint32_t* epoch_counter;
    /* SM cycles a rank waits for its peers before it gives up (prints the late rank and traps, which ends the
     * CUDA context of this process). 0 = $KB_TP_TIMEOUT_S seconds (default 30) at ~2 GHz. The collectives assume
     * nothing else occupies the GPUs: all blocks of all ranks must be running at the same time. */
    unsigned long long timeout_cycles;
} kb_tp_peers;
int kb_set_tensor_parallel_peers(kb_engine* e, int32_t tp_rank, int32_t tp_size, const kb_tp_peers* peers,
                                 int32_t blocks /* grid of the collective, 0 = default */,
                                 int32_t wire_f16 /* 0: fp32 partial sums, rank 0 folds the residual in;
                                                     1: fp16 partial sums (KB_EPI_F16), added in fp32 by the
                                                        switch, sum added to the residual by the next RMSNorm
                                                     2: fp16 partial sums, and the collective, the residual add
                                                        and the RMSNorm are ONE kernel per projection
                                                        (kb_tp_reduce_norm_gather): rank r keeps rows
                                                        [r*ceil(T/size), ...) of the residual stream, norms them
                                                        once and multicasts the normed fp16 rows; hidden <= 8192.
                                                        Region use: x fp32 | partial fp16 | normed fp16 */);
/* Vocabulary-sharded lm_head under tensor parallelism ("local top-1 + max-reduce"): the lm_head
 * bound with kb_bind_weights holds only rows [vocab_offset, vocab_offset + vocab_local) of the vocabulary (both
 * multiples of 8). kb_prefill / kb_decode then compute this rank's slice of the logits ([n_seqs, vocab_local]; a
 * caller's logits_out receives that slice), pick the local winner per sequence (same Gumbel noise, banned id and
 * lowest-id tie rule as the unsharded sampler, all on GLOBAL ids) and compare the ranks' (value, id) pairs through
 * the peer region: every rank ends with the token the unsharded sampler would have picked, 8 bytes per (rank,
 * sequence) cross NVLink, and each GPU holds 1/size of the matrix. `pairs_offset`: 2 x max_seqs x 8 bytes of the
 * region's payload, 256-byte aligned, behind the two residual buffers. Needs kb_set_tensor_parallel_peers first;
 * vocab_local = 0 goes back to a replicated lm_head.  No reference counterpart. */
int kb_set_lm_head_shard(kb_engine* e, int32_t vocab_offset, int32_t vocab_local, size_t pairs_offset);
/* The two steps on their own (unit tests). kb_sample_slice: `logits_slice` = [n_seqs, vocab_local] fp32 holding ids
 * [vocab_offset, +vocab_local) of a `vocab`-wide vocabulary; writes per sequence {fp32 winning value, int32 global id}
 * (8 bytes; id 0x7fffffff and value -inf when the slice holds no candidate) to out_pairs. kb_tp_argmax_pairs: every
 * rank has such pairs at `pairs_offset` of ITS part of the region; out_tokens[s] = id of the largest value over the
 * ranks (ties: lowest id). `epoch` as for kb_allreduce_peers (advance by 2). */
int kb_sample_slice(const float* logits_slice, int32_t n_seqs, int32_t vocab, int32_t vocab_offset, int32_t vocab_local,
                    const kb_sampling* sampling, void* out_pairs, kb_stream stream);
int kb_tp_argmax_pairs(const kb_tp_peers* peers, int32_t rank, int32_t world, size_t pairs_offset, int32_t n_seqs,
                       uint32_t epoch, int32_t* out_tokens, kb_stream stream);
/* The fused kernel of mode 2 on its own (unit tests): reduces n_rows x hidden fp16 partial sums at
 * partial_offset of the region over all ranks, adds the sums of this rank's rows to x (local fp32,
 * [n_rows, hidden]), writes fp16(rmsnorm(x) * norm_w) of those rows to xn_offset on EVERY rank.
 * `epoch` as for kb_allreduce_peers.  No reference counterpart. */
int kb_tp_reduce_norm_gather(const kb_tp_peers* peers, int32_t rank, int32_t world, size_t partial_offset,
                             size_t xn_offset, float* x, const void* norm_w_bf16, int32_t n_rows, int32_t hidden,
                             float eps, uint32_t epoch, int32_t blocks, kb_stream stream);
int kb_allreduce_peers(const kb_tp_peers* peers, int32_t rank, int32_t world, size_t offset_bytes, size_t n,
                       int32_t f16 /* 0: n fp32; 1: n fp16, added in fp32 */, uint32_t epoch, int32_t blocks,
                       kb_stream stream);

/* ---- the hot path ------------------------------------------------------------------------------
 * Both calls run embed -> n_layers x (RMSNorm, QKV GEMM, RoPE + KV append, paged attention,
 * O GEMM + residual, RMSNorm, gate/up GEMM + SwiGLU, down GEMM + residual) -> final RMSNorm of the
 * last position of each sequence -> lm_head GEMM -> sampling.
 *
 * Takes over: `response = await client.chat.completions.create(**params)` + the chunk loop,
 * src/llm/portkey.py:379 and :425-475 (streaming) / :616 (non-streaming).
 *
 * kb_prefill: incremental prefill of packed new tokens of n_seqs sequences.
 *   tokens       [n_tokens] int32 device   new token ids, sequences back to back
 *   cu_seqlens   [n_seqs+1] int32 device   prefix sums of new-token counts
 *   ctx_lens     [n_seqs]   int32 device   tokens already resident in KV per sequence
 *   block_tables [n_seqs, max_blocks] int32 device   KV block ids per sequence
 *   max_new      largest new-token count of any sequence (host knows it)
 *   logits_last  fp32 [n_seqs, vocab] device or NULL
 *   out_tokens   int32 [n_seqs] device or NULL  (sampled from the last position)
 */
int kb_prefill(kb_engine* e, const int32_t* tokens, const int32_t* cu_seqlens, const int32_t* ctx_lens,
               const int32_t* block_tables, int32_t n_seqs, int32_t n_tokens, int32_t max_blocks,
               int32_t max_new, const kb_sampling* sampling, float* logits_last, int32_t* out_tokens,
               kb_stream stream);

/* kb_decode: one token per sequence.  last_tokens may be the previous call's out_tokens (device
 * feedback, no host round trip).  If advance_ctx != 0 the call finishes by ctx_lens[i] += 1 so the
 * same arguments can be replayed for the next step.  max_ctx = largest ctx_lens[i] the host knows of
 * (only used to decide how many split-KV partitions to launch; 0 = assume max_blocks full). */
int kb_decode(kb_engine* e, const int32_t* last_tokens, int32_t* ctx_lens, const int32_t* block_tables,
              int32_t batch, int32_t max_blocks, int32_t max_ctx, const kb_sampling* sampling,
              int32_t* out_tokens, float* logits_dbg, int32_t advance_ctx, kb_stream stream);

/* ---- per-kernel entry points (unit parity tests; same kernels the hot path launches) ----------- */

/* y[t,:] = x[t,:] * rsqrt(mean(x^2) + eps) * w,  x fp32 [n, hidden], y fp16 (out_f16 != 0) or bf16 */
int kb_rmsnorm(const float* x, const void* w_bf16, void* y, int32_t n, int32_t hidden, float eps,
               int32_t out_f16, kb_stream stream);

/* x[t,:] = float(embed[token[t], :]) */
int kb_embed(const int32_t* tokens, const void* embed_bf16, float* x, int32_t n, int32_t hidden,
             int32_t vocab, kb_stream stream);

/* D = A[M,K] * W[N,K]^T on tcgen05 tensor cores, fp32 accumulation in TMEM.
 * epilogue: 0 = bf16 out[M,N]; 1 = fp32 out[M,N]; 2 = fp32 out[M,N] = resid[M,N] + D (out may alias
 * resid); 3 = SwiGLU on interleaved columns, bf16 out[M,N/2] = silu(D[:,2j]) * D[:,2j+1]. */
#define KB_EPI_BF16 0
#define KB_EPI_F32 1
#define KB_EPI_RESID_F32 2
#define KB_EPI_SWIGLU_BF16 3
#define KB_EPI_SWIGLU_F16 4 /* same, fp16 out (the hot path's activation format, see KB_GEMM_A_F16) */
#define KB_EPI_F16 5        /* out[M,N] fp16 (tensor-parallel partial sums sent as fp16, see kb_tp_peers) */
/* OR into `epilogue`: BOTH operands are fp16 instead of bf16 (tcgen05 kind::f16 requires A and B to
 * share one format; mixing traps).  The hot path keeps its GEMM activations (normed hidden states,
 * attention output, MLP inner) in fp16 — 10 mantissa bits against bf16's 7 — because their rounding is
 * the largest error term of a bf16 pipeline, and therefore binds the projection weights
 * as fp16 re-encodings of the bf16 values (exact for |w| >= 2^-14).  q/k/v, the KV cache and the
 * attention probabilities are fp16 too (round 2: at Llama-3-8B's layer geometry their bf16 rounding was the
 * largest remaining term, 4.5e-3 relative on the attention output — profiles/r02_stage_errors.md);
 * accumulation and the residual stream are fp32 throughout.  Every fp32 -> fp16 conversion SATURATES at
 * +-65504 instead of producing inf. */
#define KB_GEMM_F16 0x100
int kb_gemm(const void* a_bf16, const void* w_bf16, void* out, const float* resid, int32_t M, int32_t N,
            int32_t K, int32_t epilogue, kb_stream stream);
/* Same, with the tile forced: tile_n in {64, 128, 256} = one CTA per 128 x tile_n tile; a negative
 * multiple of 16 in [-256, -32] = a CTA pair (tcgen05 cta_group::2) per 256 x |tile_n| tile; {1, 2, 4} =
 * the small-M kernel (M <= 128: operands swapped so the 128 MMA rows are weight rows, that many CTAs
 * per tile sharing K through a cluster and adding their partial sums in rank order); -(1024 * s + bn) with
 * s in {2, 4} = the CTA-pair kernel with split-K: every 256 x bn tile is computed by s work units, each over 1/s
 * of K, that add their partial sums to `out` one after the other in a fixed order (only for the in-place
 * residual epilogue, out == resid; the heuristic picks it for the decode step's O and down projections);
 * 0 = heuristic
 * (small-M kernel for M <= 32, and up to 64 for N >= 16384 or K >= 8192; else the cheapest tile under a per-SM ingress cost model). For tuning
 * and for tests that must cover every instantiation. */
int kb_gemm_ex(const void* a_bf16, const void* w_bf16, void* out, const float* resid, int32_t M, int32_t N,
               int32_t K, int32_t epilogue, int32_t tile_n, kb_stream stream);

/* Rotate q and k (HF rotate-half convention), write q in place and k,v into the paged pools.
 *   qkv [n_tokens, (n_heads + 2 n_kv_heads) * head_dim] fp16;  token t of sequence s sits at position
 *   ctx_lens[s] + (t - cu_seqlens[s]).  k_pool_layer / v_pool_layer point at ONE layer's pool. */
int kb_rope_append(void* qkv_f16, const int32_t* cu_seqlens, const int32_t* ctx_lens,
                   const int32_t* block_tables, void* k_pool_layer, void* v_pool_layer, int32_t n_seqs,
                   int32_t n_tokens, int32_t max_blocks, int32_t n_heads, int32_t n_kv_heads,
                   int32_t head_dim, float rope_theta, kb_stream stream);
/* The same with the head_dim / 2 rotary frequencies given as a device table instead of theta^(-2i/head_dim) — scaled
 * RoPE (Llama-3.1's "llama3" rule and the like are applied by the host to the table; kafka_b200.model_config.inv_freq).
 * kb_set_rope_inv_freq makes the engine's kb_prefill / kb_decode use such a table (the pointer must stay valid; NULL
 * returns to the theta rule).  No reference counterpart. */
int kb_rope_append_freqs(void* qkv_f16, const int32_t* cu_seqlens, const int32_t* ctx_lens, const int32_t* block_tables,
                         void* k_pool_layer, void* v_pool_layer, int32_t n_seqs, int32_t n_tokens, int32_t max_blocks,
                         int32_t n_heads, int32_t n_kv_heads, int32_t head_dim, const float* inv_freq, kb_stream stream);
int kb_set_rope_inv_freq(kb_engine* e, const float* inv_freq);

/* Paged attention, one query token per sequence, over kv_lens[s] + kv_len_bias cached tokens.
 *   q   [batch, q_stride] fp16 (first n_heads*head_dim columns are the heads); the pools are fp16;
 *   out [batch, n_heads*head_dim], fp16 when out_f16 != 0 (the hot path), else bf16
 *   max_kv_len : host-side bound on kv_lens[s] + kv_len_bias (0 = max_blocks * KB_BLOCK_TOKENS)
 *   scratch    : optional device buffer for split-KV partials (kb_attn_decode_scratch_bytes); NULL = no
 *                cross-CTA split (small batches then use fewer SMs).
 * The pools must hold finite values everywhere (zero-initialise them once): masked tail positions of
 * the last block are multiplied by an exact 0. */
size_t kb_attn_decode_scratch_bytes(int32_t batch, int32_t n_heads, int32_t head_dim);
int kb_attn_decode(const void* q_f16, int32_t q_stride, const void* k_pool_layer, const void* v_pool_layer,
                   const int32_t* kv_lens, int32_t kv_len_bias, const int32_t* block_tables, void* out16,
                   int32_t batch, int32_t max_blocks, int32_t n_heads, int32_t n_kv_heads, int32_t head_dim,
                   int32_t max_kv_len, void* scratch, size_t scratch_bytes, int32_t out_f16, kb_stream stream);

/* Decode attention for callers that cannot allocate device memory.  No counterpart in the reference
 * (nor in its plugin API): it exists for hosts of this library that are handed ready-made input and
 * output buffers but have no allocator of their own — a foreign-language binding working on IPC or
 * pool memory owned by someone else, a process whose HBM is carved up in advance, or a latency-bound
 * loop that cannot afford cudaMalloc's device-wide synchronisation — and therefore cannot supply the
 * split-KV scratch kb_attn_decode asks for.  Same arguments and result as kb_attn_decode; the partials
 * live in a KB_STATIC_SCRATCH_BYTES array inside the library image (held by every GPU that loads the
 * library, until exit).  Limits and rules:
 *   - a shape whose partials exceed KB_STATIC_SCRATCH_BYTES is refused with KB_ERR_INVALID and a
 *     message; it is not quietly run with fewer splits;
 *   - the first call on each device resolves the array's device address (cudaGetSymbolAddress);
 *     make that call with the device current and before any stream capture;
 *   - no cudaMalloc / cudaFree / allocating routine is called;
 *   - all calls on a device share the array (its contents are whatever the previous call left; every
 *     call rewrites all it reads, so no clearing is needed), hence calls that can overlap on the GPU
 *     race with each other and must be serialised by the caller. */
#define KB_STATIC_SCRATCH_BYTES (2 * 1024 * 1024)
size_t kb_static_scratch_capacity(void);
int kb_attn_decode_static_scratch(const void* q_f16, int32_t q_stride, const void* k_pool_layer,
                                  const void* v_pool_layer, const int32_t* kv_lens, int32_t kv_len_bias,
                                  const int32_t* block_tables, void* out16, int32_t batch,
                                  int32_t max_blocks, int32_t n_heads, int32_t n_kv_heads, int32_t head_dim,
                                  int32_t max_kv_len, int32_t out_f16, kb_stream stream);

/* Paged causal attention for packed new tokens (incremental prefill): query i of sequence s attends
 * to the ctx_lens[s] cached tokens plus new tokens 0..i. K/V of the new tokens must already be
 * appended (kb_rope_append). */
int kb_attn_prefill(const void* q_f16, int32_t q_stride, const void* k_pool_layer, const void* v_pool_layer,
                    const int32_t* cu_seqlens, const int32_t* ctx_lens, const int32_t* block_tables,
                    void* out16, int32_t n_seqs, int32_t n_tokens, int32_t max_new, int32_t max_blocks,
                    int32_t n_heads, int32_t n_kv_heads, int32_t head_dim, int32_t out_f16, kb_stream stream);

/* out_tokens[s] = argmax_j (logits[s,j] / T + gumbel(seed, step, s, j)); greedy when T <= 0. */
int kb_sample(const float* logits, int32_t n_seqs, int32_t vocab, const kb_sampling* sampling,
              int32_t* out_tokens, kb_stream stream);

/* Device-side hand-over of sampled ids between two decode steps, so that step N+1 can be enqueued
 * while step N is still running: dst[i] = prev_out[src_index[i]] where 0 <= src_index[i] < n_prev
 * (sequence i was in step N's batch at that slot); other dst[i] keep what the host staged (sequences
 * that just left prefill).  Reference counterpart: none — the reference awaits one HTTP response per
 * call (/root/reference/src/llm/portkey.py:379) and has no step pipeline. */
int kb_select_tokens(int32_t* dst, const int32_t* prev_out, const int32_t* src_index, int32_t n,
                     int32_t n_prev, kb_stream stream);

/* ---- diagnostics ------------------------------------------------------------------------------- */
/* Per-kernel timing of the hot path with CUDA events recorded on the call's own stream between
 * launches (bench.py's roofline numbers).  kb_profile_begin arms it; every kb_prefill / kb_decode
 * until kb_profile_end adds its intervals; kb_profile_end synchronises on the last event, returns
 * the summed milliseconds and launch counts per category and disarms.  Events are host objects;
 * no device memory is allocated. */
#define KB_PROF_EMBED 0
#define KB_PROF_RMSNORM 1
#define KB_PROF_GEMM_QKV 2
#define KB_PROF_ROPE_APPEND 3
#define KB_PROF_ATTENTION 4
#define KB_PROF_GEMM_O 5
#define KB_PROF_GEMM_GATE_UP 6
#define KB_PROF_GEMM_DOWN 7
#define KB_PROF_LM_HEAD 8
#define KB_PROF_SAMPLE 9
#define KB_PROF_ALLREDUCE 10
#define KB_PROF_CATS 11
int kb_profile_begin(kb_engine* e);
int kb_profile_end(kb_engine* e, float* ms_by_cat, int32_t* launches_by_cat);

/* Read-only bandwidth probe: streams `bytes` of `buf` (16-byte aligned device memory, caller-owned, not
 * modified) through every SM and XORs what it read into `sink` (>= 4 * 148 * 4 bytes). bench.py times it over
 * a slice of the KV pool much larger than L2 to report what a purely reading kernel reaches on this GPU (the
 * driver's HBM peak is a copy measurement). Reference counterpart: none. */
int kb_hbm_read_probe(const void* buf, size_t bytes, void* sink, size_t sink_bytes, kb_stream stream);

const char* kb_last_error(void);
int kb_abi_version(void);
/* Number of kernels this library has enqueued since load (bench.py's gpu_launches). */
unsigned long long kb_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* KAFKA_B200_H */
