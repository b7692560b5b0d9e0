// kb_engine.cu — the per-GPU engine behind the C ABI: borrowed weights / KV pool / workspace and the
// forward pass (prefill and decode) that chains the kernels of kb_gemm.cu, kb_attn.cu and
// kb_elementwise.cu on one stream.
//
// Takes over `await client.chat.completions.create(**params)` (src/llm/portkey.py:379, :616).
#include "kb_common.cuh"

#include <cstddef>

#include <vector>
#include <new>
#include <string.h>

namespace kb {

thread_local char g_err[512] = {0};
unsigned long long g_launches = 0;

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// implemented in the other translation units
int gemm(const void* a, const void* w, void* out, const float* resid, int M, int N, int K, int epi,
         cudaStream_t stream);
int embed(const int* tokens, const void* embed_w, float* x, int n, int hidden, int vocab, cudaStream_t stream);
int rmsnorm(float* x, const void* w, void* y, const int* last_rows, const void* delta_f16, int n, int hidden, float eps,
            int out_f16, cudaStream_t stream);
int rope_append(void* qkv, const int* cu, const int* ctx, const int* bt, void* kp, void* vp, int n_seqs,
                int n_tokens, int max_blocks, int n_heads, int n_kv_heads, int head_dim, float theta,
                const float* inv_freq, cudaStream_t stream);
int sample(const float* logits, int n_seqs, int vocab, const kb_sampling* sp, int* out_tokens, int* ctx_advance,
           cudaStream_t stream);
int sample_slice(const float* logits, int n_seqs, int vocab, int v0, int n_local, const kb_sampling* sp, int* out_tokens,
                 int* ctx_advance, void* pair_out, cudaStream_t stream);
int tp_argmax(const kb_tp_peers* P, int rank, int world, size_t pairs_off, int n_seqs, uint32_t epoch, int* out_tokens,
              int* ctx_advance, cudaStream_t stream);
int attn_decode(const void* q, int q_stride, const void* kp, const void* vp, const int* kv_lens, int bias,
                const int* bt, void* out, int batch, int max_blocks, int n_heads, int n_kv_heads, int head_dim,
                int max_kv_len, void* scratch, size_t scratch_bytes, int out_f16, cudaStream_t stream);
int attn_prefill(const void* q, int q_stride, const void* kp, const void* vp, const int* cu, const int* ctx,
                 const int* bt, void* out, int n_seqs, int n_tokens, int max_new, int max_blocks, int n_heads,
                 int n_kv_heads, int head_dim, int out_f16, cudaStream_t stream);
size_t attn_decode_scratch_bytes(int batch, int n_heads, int head_dim, int n_splits);
int tp_reduce_norm_gather(const kb_tp_peers* P, int rank, int world, size_t partial_off, size_t xn_off, float* x,
                          const void* norm_w, int T, int hidden, float eps, uint32_t epoch, int blocks,
                          cudaStream_t stream);
int gather_last_rows(const void* src_f16, const int* cu_seqlens, void* dst_f16, int n_seqs, int hidden,
                     cudaStream_t stream);
int allreduce_peers(const kb_tp_peers* P, int rank, int world, size_t offset_bytes, size_t n, int f16, uint32_t epoch,
                    int blocks, cudaStream_t stream);

static inline size_t align256(size_t v) { return (v + 255) & ~size_t(255); }

// layout the ctypes binding mirrors (kafka_b200/cabi.py: Sampling; the ABI tests check the size from that side)
static_assert(sizeof(kb_sampling) == 80 && offsetof(kb_sampling, top_p) == 64 && offsetof(kb_sampling, seq_top_p) == 72,
              "kb_sampling layout changed: update the binding and KB_ABI_VERSION");

struct Workspace {
    float* x = nullptr;           // [T, hidden] fp32 residual stream
    float* xr = nullptr;          // [T, hidden] fp32 partial sums of a row-sharded projection (tensor parallel)
    // every 16-bit activation is FP16
    __half* xn = nullptr;         // [T, hidden] normed activations (GEMM A operand)
    __half* qkv = nullptr;        // [T, qkv_dim]
    __half* attn = nullptr;       // [T, n_heads*head_dim]
    __half* h = nullptr;          // [T, ffn]
    __half* xl = nullptr;         // [S, hidden] final-normed last rows
    float* logits = nullptr;      // [S, vocab]
    void* split = nullptr;        // split-KV partials
    size_t split_bytes = 0;
    size_t total = 0;
};

static void layout_workspace(const kb_model_cfg& c, int T, int S, uint8_t* base, Workspace* w) {
    size_t off = 0;
    auto take = [&](size_t bytes) {
        uint8_t* p = base ? base + off : nullptr;
        off += align256(bytes);
        return p;
    };
    const size_t qkv_dim = (size_t)(c.n_heads + 2 * c.n_kv_heads) * c.head_dim;
    w->x = reinterpret_cast<float*>(take((size_t)T * c.hidden * 4));
    w->xr = reinterpret_cast<float*>(take((size_t)T * c.hidden * 4));
    w->xn = reinterpret_cast<__half*>(take((size_t)T * c.hidden * 2));
    w->qkv = reinterpret_cast<__half*>(take((size_t)T * qkv_dim * 2));
    w->attn = reinterpret_cast<__half*>(take((size_t)T * c.n_heads * c.head_dim * 2));
    w->h = reinterpret_cast<__half*>(take((size_t)T * c.ffn * 2));
    w->xl = reinterpret_cast<__half*>(take((size_t)S * c.hidden * 2));
    w->logits = reinterpret_cast<float*>(take((size_t)S * c.vocab * 4));
    w->split_bytes = attn_decode_scratch_bytes(S, c.n_heads, c.head_dim, KB_MAX_SPLITS);
    w->split = take(w->split_bytes);
    w->total = off;
}

static int check_cfg(const kb_model_cfg* c) {
    KB_CHECK(c != nullptr, "null model cfg");
    KB_CHECK(c->hidden > 0 && c->hidden % 64 == 0, "hidden=%d must be a positive multiple of 64", c->hidden);
    KB_CHECK(c->ffn > 0 && c->ffn % 64 == 0, "ffn=%d must be a positive multiple of 64", c->ffn);
    KB_CHECK(c->head_dim == 64 || c->head_dim == 128, "head_dim=%d must be 64 or 128", c->head_dim);
    KB_CHECK(c->n_layers > 0 && c->n_heads > 0 && c->n_kv_heads > 0, "bad layer/head counts");
    KB_CHECK(c->n_heads % c->n_kv_heads == 0, "n_heads %% n_kv_heads != 0");
    const int g = c->n_heads / c->n_kv_heads;
    KB_CHECK(g == 1 || g == 2 || g == 4 || g == 8, "GQA group %d not in {1,2,4,8}", g);
    KB_CHECK((c->n_heads * c->head_dim) % 64 == 0, "n_heads*head_dim must be a multiple of 64");
    KB_CHECK(c->vocab > 0 && c->vocab % 8 == 0, "vocab=%d must be a positive multiple of 8", c->vocab);
    KB_CHECK(c->rope_theta > 1.f && c->rms_eps > 0.f, "bad rope_theta / rms_eps");
    return KB_OK;
}

}  // namespace kb

struct kb_engine {
    int device = 0;
    kb_model_cfg cfg{};
    std::vector<kb_layer_weights> layers;
    const void* embed = nullptr;
    const void* final_norm = nullptr;
    const void* lm_head = nullptr;
    bool have_weights = false;
    __half* k_pool = nullptr;     // fp16 pools
    __half* v_pool = nullptr;
    int n_blocks = 0;
    kb::Workspace ws;
    int max_tokens = 0, max_seqs = 0;
    // tensor parallel (kb_set_tensor_parallel): cfg holds the LOCAL head / ffn counts of this rank
    int tp_rank = 0, tp_size = 1;
    kb_allreduce_fn allreduce = nullptr;
    void* allreduce_ctx = nullptr;
    // ... or the in-kernel all-reduce over peer memory (kb_set_tensor_parallel_peers)
    bool use_peers = false;
    kb_tp_peers peers{};
    int peer_blocks = 0;
    int peer_wire = 0;            // 0 fp32 partial sums | 1 fp16, added to the residual by the next RMSNorm |
                                  // 2 fp16, collective + residual add + RMSNorm fused (kb_tp_reduce_norm_gather)
    uint32_t peer_epoch = 1;      // only used when the region carries no shared counter (kb_tp_peers.epoch_counter)
    // vocabulary-sharded lm_head (kb_set_lm_head_shard): the bound lm_head holds rows [lm_v0, lm_v0 + lm_local)
    int lm_v0 = 0, lm_local = 0;  // lm_local == 0: the whole vocabulary (replicated)
    size_t lm_pairs_off = 0;      // region offset of 2 x max_seqs x 8 bytes: this rank's (value, id) pairs
    int lm_parity = 0;
    const float* rope_inv_freq = nullptr;   // head_dim / 2 rotary frequencies (kb_set_rope_inv_freq), else theta^(-2i/hd)
    // profiling: one event before every kernel (+ one closing event per call)
    bool prof_on = false;
    std::vector<cudaEvent_t> prof_pool;  // reusable events
    size_t prof_used = 0;
    std::vector<std::pair<int, std::pair<cudaEvent_t, cudaEvent_t>>> prof_iv;  // (cat, (start, stop))
    cudaEvent_t prof_last = nullptr;
    int prof_last_cat = -1;
};

namespace kb {

// Epoch of the next collective on the engine's region: taken from the region's shared host counter when there
// is one (so that every engine attached to the region, and direct kb_allreduce_peers callers, continue ONE
// sequence — a reused epoch would let the flag barriers pass without waiting), else from the engine.
static inline uint32_t next_epoch(kb_engine* e) {
    uint32_t* c = e->peers.epoch_counter ? e->peers.epoch_counter : &e->peer_epoch;
    const uint32_t v = *c;
    *c = v + 2;
    return v;
}

// One forward pass over n_tokens packed tokens of n_seqs sequences.
//   decode == true : one token per sequence, cu_seqlens == nullptr
static int forward(kb_engine* e, const int* tokens, const int* cu_seqlens, int* ctx_lens, const int* block_tables,
                   int n_seqs, int n_tokens, int max_blocks, int max_new, int max_ctx, bool decode,
                   const kb_sampling* sampling, float* logits_out, int* out_tokens, bool advance_ctx,
                   cudaStream_t stream) {
    KB_CHECK(e != nullptr, "null engine");
    KB_CHECK(e->have_weights, "weights not bound (kb_bind_weights)");
    KB_CHECK(e->k_pool && e->v_pool, "KV pool not bound (kb_bind_kv_pool)");
    KB_CHECK(e->ws.x != nullptr, "workspace not bound (kb_bind_workspace)");
    KB_CHECK(n_seqs > 0 && n_tokens >= n_seqs, "bad batch: n_seqs=%d n_tokens=%d", n_seqs, n_tokens);
    KB_CHECK(n_tokens <= e->max_tokens, "n_tokens=%d exceeds workspace max_tokens=%d", n_tokens, e->max_tokens);
    KB_CHECK(n_seqs <= e->max_seqs, "n_seqs=%d exceeds workspace max_seqs=%d", n_seqs, e->max_seqs);
    KB_CHECK(tokens && ctx_lens && block_tables, "null token / ctx_lens / block_tables pointer");
    KB_CHECK(decode || cu_seqlens != nullptr, "prefill needs cu_seqlens");
    KB_CHECK(out_tokens == nullptr || sampling != nullptr, "out_tokens requested without sampling params");
    KB_CHECK(e->tp_size <= 1 || e->allreduce != nullptr || e->use_peers,
             "tensor parallel engine without an all-reduce (kb_set_tensor_parallel / _peers)");
    int dev = -1;
    KB_CUDA(cudaGetDevice(&dev));
    KB_CHECK(dev == e->device, "engine belongs to device %d but current device is %d", e->device, dev);

    // profiling marks: an event before each kernel closes the previous interval and opens the next
    auto mark = [&](int cat) {
        if (!e->prof_on) return;
        if (e->prof_used == e->prof_pool.size()) {
            cudaEvent_t ev;
            if (cudaEventCreate(&ev) != cudaSuccess) return;
            e->prof_pool.push_back(ev);
        }
        cudaEvent_t ev = e->prof_pool[e->prof_used++];
        cudaEventRecord(ev, stream);
        if (e->prof_last_cat >= 0) e->prof_iv.push_back({e->prof_last_cat, {e->prof_last, ev}});
        e->prof_last = ev;
        e->prof_last_cat = cat;
    };
    const kb_model_cfg& c = e->cfg;
    Workspace& w = e->ws;
    const int qdim = c.n_heads * c.head_dim;
    const int qkv_dim = (c.n_heads + 2 * c.n_kv_heads) * c.head_dim;
    const size_t layer_stride = (size_t)e->n_blocks * c.n_kv_heads * KB_BLOCK_TOKENS * c.head_dim;
    int rc;
    float* x = w.x;    // residual stream; with tensor parallelism x and xr trade places after every all-reduce
    float* xr = w.xr;
    const void* delta = nullptr;   // fp16 wire: reduced projection output not yet added to x (next rmsnorm does)

    // x += a[T,K] * wt[hidden,K]^T.  Tensor parallel: this rank holds a K-slice of the projection, the
    // partial sums go through the all-reduce. fp32 wire: rank 0 folds the residual into its partial, so
    // the reduced buffer IS the new residual stream (no separate add). fp16 wire (half the NVLink
    // bytes): the partials are stored as fp16, summed in fp32 inside the switch, and the fp16 sum is
    // added to the fp32 residual by the RMSNorm that follows.
    auto project_into_residual = [&](const void* a, const void* wt, int K) -> int {
        if (e->tp_size <= 1)
            return gemm(a, wt, x, x, n_tokens, c.hidden, K, KB_EPI_RESID_F32 | KB_GEMM_F16, stream);
        if (e->use_peers && e->peer_wire == 1) {
            int r = gemm(a, wt, w.xr, nullptr, n_tokens, c.hidden, K, KB_EPI_F16 | KB_GEMM_F16, stream);
            if (r) return r;
            mark(KB_PROF_ALLREDUCE);
            const size_t off = reinterpret_cast<char*>(w.xr) - static_cast<char*>(e->peers.peer_base[e->tp_rank]);
            r = allreduce_peers(&e->peers, e->tp_rank, e->tp_size, off, (size_t)n_tokens * c.hidden, 1, next_epoch(e),
                                e->peer_blocks, stream);
            delta = w.xr;
            return r;
        }
        int r = (e->tp_rank == 0)
                    ? gemm(a, wt, xr, x, n_tokens, c.hidden, K, KB_EPI_RESID_F32 | KB_GEMM_F16, stream)
                    : gemm(a, wt, xr, nullptr, n_tokens, c.hidden, K, KB_EPI_F32 | KB_GEMM_F16, stream);
        if (r) return r;
        mark(KB_PROF_ALLREDUCE);
        if (e->use_peers) {
            const size_t off = reinterpret_cast<char*>(xr) - static_cast<char*>(e->peers.peer_base[e->tp_rank]);
            r = allreduce_peers(&e->peers, e->tp_rank, e->tp_size, off, (size_t)n_tokens * c.hidden, 0, next_epoch(e),
                                e->peer_blocks, stream);
            if (r) return r;
        } else if (e->allreduce(e->allreduce_ctx, xr, (size_t)n_tokens * c.hidden, stream) != 0) {
            set_error("tensor-parallel all-reduce callback failed");
            return KB_ERR_INVALID;
        }
        float* t = x; x = xr; xr = t;
        return KB_OK;
    };

    // Fused tensor-parallel mode: after each row-sharded projection ONE kernel reduces the fp16 partial
    // sums, adds them to this rank's rows of the residual stream, norms those rows with the NEXT norm's
    // weights and multicasts the normed rows — the following GEMM reads them from `xn`.
    const bool fused = e->tp_size > 1 && e->use_peers && e->peer_wire == 2;
    char* region = fused ? static_cast<char*>(e->peers.peer_base[e->tp_rank]) : nullptr;
    const size_t part_off = fused ? static_cast<size_t>(reinterpret_cast<char*>(w.xr) - region) : 0;
    const size_t xn_off = part_off + align256((size_t)e->max_tokens * c.hidden * 2);
    void* xn = fused ? static_cast<void*>(region + xn_off) : static_cast<void*>(w.xn);
    auto project_reduce_norm = [&](const void* a, const void* wt, int K, const void* next_norm) -> int {
        int r = gemm(a, wt, w.xr, nullptr, n_tokens, c.hidden, K, KB_EPI_F16 | KB_GEMM_F16, stream);
        if (r) return r;
        mark(KB_PROF_ALLREDUCE);
        r = tp_reduce_norm_gather(&e->peers, e->tp_rank, e->tp_size, part_off, xn_off, x, next_norm, n_tokens, c.hidden,
                                  c.rms_eps, next_epoch(e), e->peer_blocks, stream);
        return r;
    };

    mark(KB_PROF_EMBED);
    if ((rc = embed(tokens, e->embed, x, n_tokens, c.hidden, c.vocab, stream))) return rc;
    for (int L = 0; L < c.n_layers; ++L) {
        const kb_layer_weights& lw = e->layers[L];
        __half* kp = e->k_pool + (size_t)L * layer_stride;
        __half* vp = e->v_pool + (size_t)L * layer_stride;
        if (!fused || L == 0) {
            mark(KB_PROF_RMSNORM);
            if ((rc = rmsnorm(x, lw.attn_norm, xn, nullptr, delta, n_tokens, c.hidden, c.rms_eps, 1, stream))) return rc;
            delta = nullptr;
        }
        mark(KB_PROF_GEMM_QKV);
        if ((rc = gemm(xn, lw.wqkv, w.qkv, nullptr, n_tokens, qkv_dim, c.hidden, KB_EPI_F16 | KB_GEMM_F16, stream)))
            return rc;
        mark(KB_PROF_ROPE_APPEND);
        if ((rc = rope_append(w.qkv, cu_seqlens, ctx_lens, block_tables, kp, vp, n_seqs, n_tokens, max_blocks,
                              c.n_heads, c.n_kv_heads, c.head_dim, c.rope_theta, e->rope_inv_freq, stream)))
            return rc;
        mark(KB_PROF_ATTENTION);
        if (decode) {
            rc = attn_decode(w.qkv, qkv_dim, kp, vp, ctx_lens, /*bias=*/1, block_tables, w.attn, n_seqs, max_blocks,
                             c.n_heads, c.n_kv_heads, c.head_dim, max_ctx > 0 ? max_ctx + 1 : 0, w.split,
                             w.split_bytes, /*out_f16=*/1, stream);
        } else {
            rc = attn_prefill(w.qkv, qkv_dim, kp, vp, cu_seqlens, ctx_lens, block_tables, w.attn, n_seqs, n_tokens,
                              max_new, max_blocks, c.n_heads, c.n_kv_heads, c.head_dim, /*out_f16=*/1, stream);
        }
        if (rc) return rc;
        mark(KB_PROF_GEMM_O);
        if (fused) {
            if ((rc = project_reduce_norm(w.attn, lw.wo, qdim, lw.mlp_norm))) return rc;
        } else {
            if ((rc = project_into_residual(w.attn, lw.wo, qdim))) return rc;
            mark(KB_PROF_RMSNORM);
            if ((rc = rmsnorm(x, lw.mlp_norm, xn, nullptr, delta, n_tokens, c.hidden, c.rms_eps, 1, stream))) return rc;
            delta = nullptr;
        }
        mark(KB_PROF_GEMM_GATE_UP);
        if ((rc = gemm(xn, lw.w_gate_up, w.h, nullptr, n_tokens, 2 * c.ffn, c.hidden,
                       KB_EPI_SWIGLU_F16 | KB_GEMM_F16, stream)))
            return rc;
        mark(KB_PROF_GEMM_DOWN);
        if (fused) {
            const void* next_norm = (L + 1 < c.n_layers) ? e->layers[L + 1].attn_norm : e->final_norm;
            if ((rc = project_reduce_norm(w.h, lw.w_down, c.ffn, next_norm))) return rc;
        } else if ((rc = project_into_residual(w.h, lw.w_down, c.ffn))) {
            return rc;
        }
    }
    if (logits_out == nullptr && out_tokens == nullptr) {  // KV fill only
        mark(-1);
        return KB_OK;
    }
    mark(KB_PROF_RMSNORM);
    // final norm on the last row of every sequence, lm_head, sampling
    const void* xl = w.xl;
    if (fused) {   // every row is already final-normed in xn: decode uses it as is, prefill picks the last rows
        if (decode) xl = xn;
        else if ((rc = gather_last_rows(xn, cu_seqlens, w.xl, n_seqs, c.hidden, stream))) return rc;
    } else if ((rc = rmsnorm(x, e->final_norm, w.xl, decode ? nullptr : cu_seqlens, delta, n_seqs, c.hidden, c.rms_eps,
                             1, stream))) {
        return rc;
    }
    float* lg = logits_out ? logits_out : w.logits;
    mark(KB_PROF_LM_HEAD);
    if (e->lm_local > 0) {
        // vocabulary-sharded lm_head: logits of this rank's slice only ([n_seqs, lm_local]; a caller's logits_out
        // receives that slice), local winner per sequence, then the ranks' winners are compared over the region
        KB_CHECK(e->use_peers && e->tp_size > 1, "sharded lm_head without the peer region (kb_set_tensor_parallel_peers)");
        if ((rc = gemm(xl, e->lm_head, lg, nullptr, n_seqs, e->lm_local, c.hidden, KB_EPI_F32 | KB_GEMM_F16, stream)))
            return rc;
        if (out_tokens) {
            mark(KB_PROF_SAMPLE);
            const size_t off = e->lm_pairs_off + (size_t)e->lm_parity * e->max_seqs * 8;
            e->lm_parity ^= 1;
            void* pairs = static_cast<char*>(e->peers.peer_base[e->tp_rank]) + off;
            if ((rc = sample_slice(lg, n_seqs, c.vocab, e->lm_v0, e->lm_local, sampling, nullptr, nullptr, pairs, stream)))
                return rc;
            if ((rc = tp_argmax(&e->peers, e->tp_rank, e->tp_size, off, n_seqs, next_epoch(e), out_tokens,
                                advance_ctx ? ctx_lens : nullptr, stream)))
                return rc;
        }
        mark(-1);
        return KB_OK;
    }
    if ((rc = gemm(xl, e->lm_head, lg, nullptr, n_seqs, c.vocab, c.hidden, KB_EPI_F32 | KB_GEMM_F16, stream)))
        return rc;
    if (out_tokens) {
        mark(KB_PROF_SAMPLE);
        if ((rc = sample(lg, n_seqs, c.vocab, sampling, out_tokens, advance_ctx ? ctx_lens : nullptr, stream)))
            return rc;
    }
    mark(-1);
    return KB_OK;
}

}  // namespace kb

// ---------------------------------------------------------------------------------------------
// extern "C"
// ---------------------------------------------------------------------------------------------
extern "C" const char* kb_last_error(void) { return kb::g_err; }
extern "C" int kb_abi_version(void) { return KB_ABI_VERSION; }
extern "C" unsigned long long kb_launch_count(void) { return kb::g_launches; }

extern "C" int kb_engine_create(int device, const kb_model_cfg* cfg, kb_engine** out) {
    KB_CHECK(out != nullptr, "kb_engine_create: null out");
    *out = nullptr;
    int rc = kb::check_cfg(cfg);
    if (rc) return rc;
    int count = 0;
    cudaError_t ce = cudaGetDeviceCount(&count);
    if (ce != cudaSuccess || count <= 0) {
        kb::set_error("kb_engine_create: no CUDA device available (%s); this library has no CPU fallback",
                      cudaGetErrorString(ce));
        return KB_ERR_CUDA;
    }
    KB_CHECK(device >= 0 && device < count, "kb_engine_create: device %d out of range (%d devices)", device, count);
    cudaDeviceProp prop;
    KB_CUDA(cudaGetDeviceProperties(&prop, device));
    KB_CHECK(prop.major == 10, "kb_engine_create: device %d is sm_%d%d; this library is built for sm_100a only", device,
             prop.major, prop.minor);
    kb_engine* e = new (std::nothrow) kb_engine();
    KB_CHECK(e != nullptr, "kb_engine_create: out of host memory");
    e->device = device;
    e->cfg = *cfg;
    *out = e;
    return KB_OK;
}

extern "C" void kb_engine_destroy(kb_engine* e) {
    if (!e) return;
    for (cudaEvent_t ev : e->prof_pool) cudaEventDestroy(ev);
    delete e;
}

extern "C" int kb_set_rope_inv_freq(kb_engine* e, const float* inv_freq) {
    KB_CHECK(e != nullptr, "kb_set_rope_inv_freq: null engine");
    e->rope_inv_freq = inv_freq;     // head_dim / 2 floats on the engine's device, owned by the caller; NULL = theta rule
    return KB_OK;
}

extern "C" int kb_set_lm_head_shard(kb_engine* e, int32_t vocab_offset, int32_t vocab_local, size_t pairs_offset) {
    using namespace kb;
    KB_CHECK(e != nullptr, "kb_set_lm_head_shard: null engine");
    if (vocab_local <= 0) {          // back to a replicated lm_head
        e->lm_v0 = e->lm_local = 0;
        return KB_OK;
    }
    KB_CHECK(e->use_peers && e->tp_size > 1, "kb_set_lm_head_shard: configure the peer region first "
             "(kb_set_tensor_parallel_peers)");
    KB_CHECK(vocab_offset >= 0 && vocab_offset % 8 == 0 && vocab_local % 8 == 0 &&
                 vocab_offset + vocab_local <= e->cfg.vocab,
             "kb_set_lm_head_shard: slice [%d, +%d) of %d (offset and size multiples of 8)", vocab_offset, vocab_local,
             e->cfg.vocab);
    const size_t each = align256((size_t)e->max_tokens * e->cfg.hidden * 4);
    const size_t need = 2 * (size_t)e->max_seqs * 8;
    KB_CHECK(pairs_offset % 256 == 0 && pairs_offset >= 2 * each && pairs_offset + need <= e->peers.flags_offset,
             "kb_set_lm_head_shard: %zu bytes of pairs at %zu overlap the residual buffers (2 x %zu) or the flags (%zu)",
             need, pairs_offset, each, e->peers.flags_offset);
    e->lm_v0 = vocab_offset;
    e->lm_local = vocab_local;
    e->lm_pairs_off = pairs_offset;
    e->lm_parity = 0;
    return KB_OK;
}

extern "C" int kb_set_tensor_parallel_peers(kb_engine* e, int32_t tp_rank, int32_t tp_size, const kb_tp_peers* peers,
                                            int32_t blocks, int32_t wire_f16) {
    using namespace kb;
    KB_CHECK(e != nullptr && peers != nullptr, "kb_set_tensor_parallel_peers: null argument");
    KB_CHECK(tp_size >= 2 && tp_size <= KB_TP_MAX_RANKS && tp_rank >= 0 && tp_rank < tp_size,
             "kb_set_tensor_parallel_peers: bad rank %d of %d (2..%d ranks)", tp_rank, tp_size, KB_TP_MAX_RANKS);
    KB_CHECK(e->ws.x != nullptr, "kb_set_tensor_parallel_peers: bind the workspace first");
    const size_t each = align256((size_t)e->max_tokens * e->cfg.hidden * 4);
    KB_CHECK(peers->flags_offset >= 2 * each && peers->flags_offset % 256 == 0 &&
                 peers->flags_offset + KB_TP_FLAG_BYTES <= peers->bytes,
             "kb_set_tensor_parallel_peers: region of %zu bytes (flags at %zu) cannot hold 2 x %zu bytes + flags",
             peers->bytes, peers->flags_offset, each);
    for (int r = 0; r < tp_size; ++r)
        KB_CHECK(peers->peer_base[r] != nullptr, "kb_set_tensor_parallel_peers: no mapping of rank %d", r);
    e->tp_rank = tp_rank;
    e->tp_size = tp_size;
    e->peers = *peers;
    e->peer_blocks = blocks;
    KB_CHECK(wire_f16 >= 0 && wire_f16 <= 2, "kb_set_tensor_parallel_peers: wire mode %d not in {0,1,2}", wire_f16);
    KB_CHECK(wire_f16 != 2 || e->cfg.hidden <= 8192, "kb_set_tensor_parallel_peers: fused mode needs hidden <= 8192");
    e->peer_wire = wire_f16;
    // the epoch sequence is never reset: re-configuring (blocks, wire format, back from the callback) must
    // continue the flag sequence of the region (next_epoch)
    e->use_peers = true;
    e->ws.x = reinterpret_cast<float*>(peers->peer_base[tp_rank]);
    e->ws.xr = reinterpret_cast<float*>(static_cast<char*>(peers->peer_base[tp_rank]) + each);
    return KB_OK;
}

extern "C" int kb_set_tensor_parallel(kb_engine* e, int32_t tp_rank, int32_t tp_size, kb_allreduce_fn fn, void* ctx) {
    KB_CHECK(e != nullptr, "kb_set_tensor_parallel: null engine");
    KB_CHECK(tp_size >= 1 && tp_rank >= 0 && tp_rank < tp_size, "kb_set_tensor_parallel: bad rank %d of %d", tp_rank,
             tp_size);
    KB_CHECK(tp_size == 1 || fn != nullptr, "kb_set_tensor_parallel: tp_size > 1 needs an all-reduce callback");
    e->tp_rank = tp_rank;
    e->tp_size = tp_size;
    e->allreduce = fn;
    e->use_peers = false;
    e->allreduce_ctx = ctx;
    return KB_OK;
}

extern "C" int kb_profile_begin(kb_engine* e) {
    KB_CHECK(e != nullptr, "kb_profile_begin: null engine");
    e->prof_on = true;
    e->prof_used = 0;
    e->prof_iv.clear();
    e->prof_last = nullptr;
    e->prof_last_cat = -1;
    return KB_OK;
}

extern "C" int kb_profile_end(kb_engine* e, float* ms_by_cat, int32_t* launches_by_cat) {
    KB_CHECK(e != nullptr && ms_by_cat != nullptr && launches_by_cat != nullptr, "kb_profile_end: null argument");
    e->prof_on = false;
    for (int i = 0; i < KB_PROF_CATS; ++i) { ms_by_cat[i] = 0.f; launches_by_cat[i] = 0; }
    if (e->prof_last) KB_CUDA(cudaEventSynchronize(e->prof_last));
    for (auto& iv : e->prof_iv) {
        float ms = 0.f;
        KB_CUDA(cudaEventElapsedTime(&ms, iv.second.first, iv.second.second));
        if (iv.first >= 0 && iv.first < KB_PROF_CATS) {
            ms_by_cat[iv.first] += ms;
            launches_by_cat[iv.first] += 1;
        }
    }
    e->prof_iv.clear();
    e->prof_used = 0;
    return KB_OK;
}

extern "C" int kb_bind_weights(kb_engine* e, const kb_weight_ptrs* w) {
    KB_CHECK(e && w && w->layers, "kb_bind_weights: null argument");
    KB_CHECK(w->embed && w->final_norm && w->lm_head, "kb_bind_weights: null embed / final_norm / lm_head");
    e->layers.assign(w->layers, w->layers + e->cfg.n_layers);
    for (int i = 0; i < e->cfg.n_layers; ++i) {
        const kb_layer_weights& l = e->layers[i];
        KB_CHECK(l.attn_norm && l.wqkv && l.wo && l.mlp_norm && l.w_gate_up && l.w_down,
                 "kb_bind_weights: null pointer in layer %d", i);
    }
    e->embed = w->embed;
    e->final_norm = w->final_norm;
    e->lm_head = w->lm_head;
    e->have_weights = true;
    return KB_OK;
}

extern "C" int kb_bind_kv_pool(kb_engine* e, void* k_pool, void* v_pool, int32_t n_blocks) {
    KB_CHECK(e && k_pool && v_pool && n_blocks > 0, "kb_bind_kv_pool: bad arguments");
    KB_CHECK((reinterpret_cast<uintptr_t>(k_pool) & 127) == 0 && (reinterpret_cast<uintptr_t>(v_pool) & 127) == 0,
             "kb_bind_kv_pool: pools must be 128-byte aligned");
    e->k_pool = reinterpret_cast<__half*>(k_pool);
    e->v_pool = reinterpret_cast<__half*>(v_pool);
    e->n_blocks = n_blocks;
    return KB_OK;
}

extern "C" size_t kb_workspace_bytes(const kb_model_cfg* cfg, int32_t max_tokens, int32_t max_seqs) {
    if (kb::check_cfg(cfg) || max_tokens <= 0 || max_seqs <= 0) return 0;
    kb::Workspace w;
    kb::layout_workspace(*cfg, max_tokens, max_seqs, nullptr, &w);
    return w.total;
}

extern "C" int kb_bind_workspace(kb_engine* e, void* ws, size_t bytes, int32_t max_tokens, int32_t max_seqs) {
    KB_CHECK(e && ws && max_tokens > 0 && max_seqs > 0, "kb_bind_workspace: bad arguments");
    KB_CHECK((reinterpret_cast<uintptr_t>(ws) & 255) == 0, "kb_bind_workspace: workspace must be 256-byte aligned");
    kb::Workspace w;
    kb::layout_workspace(e->cfg, max_tokens, max_seqs, reinterpret_cast<uint8_t*>(ws), &w);
    KB_CHECK(w.total <= bytes, "kb_bind_workspace: need %zu bytes, got %zu", w.total, bytes);
    e->ws = w;
    e->max_tokens = max_tokens;
    e->max_seqs = max_seqs;
    return KB_OK;
}

extern "C" int kb_prefill(kb_engine* e, const int32_t* tokens, const int32_t* cu_seqlens, const int32_t* ctx_lens,
                          const int32_t* block_tables, int32_t n_seqs, int32_t n_tokens, int32_t max_blocks,
                          int32_t max_new, const kb_sampling* sampling, float* logits_last, int32_t* out_tokens,
                          kb_stream stream) {
    return kb::forward(e, tokens, cu_seqlens, const_cast<int32_t*>(ctx_lens), block_tables, n_seqs, n_tokens,
                       max_blocks, max_new, 0, false, sampling, logits_last, out_tokens, false,
                       reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int kb_decode(kb_engine* e, const int32_t* last_tokens, int32_t* ctx_lens, const int32_t* block_tables,
                         int32_t batch, int32_t max_blocks, int32_t max_ctx, const kb_sampling* sampling,
                         int32_t* out_tokens, float* logits_dbg, int32_t advance_ctx, kb_stream stream) {
    return kb::forward(e, last_tokens, nullptr, ctx_lens, block_tables, batch, batch, max_blocks, 1, max_ctx, true,
                       sampling, logits_dbg, out_tokens, advance_ctx != 0, reinterpret_cast<cudaStream_t>(stream));
}
