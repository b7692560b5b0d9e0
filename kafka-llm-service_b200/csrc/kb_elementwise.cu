// kb_elementwise.cu — the HBM-bound glue kernels of the completion path (sm_100a):
// token embedding gather, RMSNorm, RoPE + paged KV append, sampling.
//
// Takes over pieces of the remote completion (src/llm/portkey.py:379); the reference has no local
// arithmetic.  All kernels are pure streaming: one pass over their inputs with 128-bit accesses.
#include "kb_common.cuh"

#include <stdlib.h>

namespace kb {

// ---------------------------------------------------------------------------------------------
// embed: x[t,:] = float(E[token[t],:])        bytes/token = hidden * (2 + 4)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
embed_kernel(const int* __restrict__ tokens, const __nv_bfloat16* __restrict__ embed, float* __restrict__ x,
             int hidden, int vocab) {
    pdl_wait();
    pdl_launch();
    const int t = blockIdx.x;
    int tok = tokens[t];
    tok = tok < 0 ? 0 : (tok >= vocab ? vocab - 1 : tok);
    const uint4* src = reinterpret_cast<const uint4*>(embed + (size_t)tok * hidden);
    float4* dst = reinterpret_cast<float4*>(x + (size_t)t * hidden);
    for (int i = threadIdx.x; i < hidden / 8; i += blockDim.x) {
        const uint4 v = src[i];
        const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&v);
        const float2 a = __bfloat1622float2(p[0]), b = __bfloat1622float2(p[1]);
        const float2 c = __bfloat1622float2(p[2]), d = __bfloat1622float2(p[3]);
        dst[2 * i] = make_float4(a.x, a.y, b.x, b.y);
        dst[2 * i + 1] = make_float4(c.x, c.y, d.x, d.y);
    }
}

// ---------------------------------------------------------------------------------------------
// rmsnorm: y[i,:] = act16(x[row(i),:] * rsqrt(mean(x^2) + eps) * w)   (act16 = fp16 or bf16)
//   row(i) = i, or cu_seqlens[i+1]-1 when last_rows != nullptr (final norm of each sequence's last token)
//   bytes/row = hidden * (4 + 2) (+ w, L2-resident)
//   One 256-thread block per row. A warp per row (the row in 32 float4 registers per lane, no block barrier)
//   was measured slower: 8.6 vs 5.5 us per launch at 512 x 4096 (tools/elementwise_bench.py).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
rmsnorm_kernel(float* __restrict__ x, const __nv_bfloat16* __restrict__ w, uint16_t* __restrict__ y,
               const int* __restrict__ last_rows, const __half* __restrict__ delta, int hidden, float eps,
               int out_f16) {
    pdl_wait();
    pdl_launch();
    const int i = blockIdx.x;
    const int row = last_rows ? last_rows[i + 1] - 1 : i;
    float4* xr = reinterpret_cast<float4*>(x + (size_t)row * hidden);
    const int n4 = hidden / 4;
    __shared__ float red[8];
    // the row stays in registers between the two passes (hidden <= 8192: 8 float4 per thread); wider rows
    // are re-read from L2
    constexpr int KEEP = 8;
    const bool keep = n4 <= KEEP * 256;
    float4 v[KEEP];
    float ss = 0.f;
    const uint2* dr = delta ? reinterpret_cast<const uint2*>(delta + (size_t)row * hidden) : nullptr;
#pragma unroll
    for (int k = 0; k < KEEP; ++k) {
        const int j = threadIdx.x + k * 256;
        if (j < n4) {
            float4 t = xr[j];
            if (dr != nullptr) {
                // tensor parallel, fp16 wire: the reduced projection output has not been added to the residual
                // stream yet — do it here (x += delta, written back) instead of in a kernel of its own
                const uint2 d = dr[j];
                const float2 d0 = __half22float2(*reinterpret_cast<const __half2*>(&d.x));
                const float2 d1 = __half22float2(*reinterpret_cast<const __half2*>(&d.y));
                t.x += d0.x; t.y += d0.y; t.z += d1.x; t.w += d1.y;
                xr[j] = t;
            }
            v[k] = t;
            ss += t.x * t.x + t.y * t.y + t.z * t.z + t.w * t.w;
        }
    }
    for (int j = threadIdx.x + KEEP * 256; j < n4; j += 256) {     // hidden > 8192 only
        float4 t = xr[j];
        if (dr != nullptr) {
            const uint2 d = dr[j];
            const float2 d0 = __half22float2(*reinterpret_cast<const __half2*>(&d.x));
            const float2 d1 = __half22float2(*reinterpret_cast<const __half2*>(&d.y));
            t.x += d0.x; t.y += d0.y; t.z += d1.x; t.w += d1.y;
            xr[j] = t;
        }
        ss += t.x * t.x + t.y * t.y + t.z * t.z + t.w * t.w;
    }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) tot += red[k];
    const float inv = rsqrtf(tot / (float)hidden + eps);
    const uint2* wr = reinterpret_cast<const uint2*>(w);
    uint2* yr = reinterpret_cast<uint2*>(y + (size_t)i * hidden);
    auto emit = [&](int j, const float4& t) {
        const uint2 wv = wr[j];
        const float2 w0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&wv.x));
        const float2 w1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&wv.y));
        uint2 o;
        o.x = pack_act(t.x * inv * w0.x, t.y * inv * w0.y, out_f16);
        o.y = pack_act(t.z * inv * w1.x, t.w * inv * w1.y, out_f16);
        yr[j] = o;
    };
#pragma unroll
    for (int k = 0; k < KEEP; ++k) {
        const int j = threadIdx.x + k * 256;
        if (j < n4) emit(j, v[k]);
    }
    if (!keep)
        for (int j = threadIdx.x + KEEP * 256; j < n4; j += 256) emit(j, xr[j]);
}

// ---------------------------------------------------------------------------------------------
// rope + KV append.  One block per token.
//   q heads : rotated in place in qkv
//   k heads : rotated, written to k_pool (swizzled paged layout)
//   v heads : copied to v_pool
// qkv, and the pools, are FP16 (the QKV GEMM's KB_EPI_F16 epilogue): q and k are rounded to 16
// bits exactly twice (GEMM output, rotation), each time with 10 mantissa bits.
// HF "rotate_half" convention: pairs (i, i + hd/2), angle = pos * theta^(-2i/hd), fp32 math.
//   bytes/token = (n_heads*2 + n_kv_heads*2*2) * head_dim * 2  (read+write q, read k/v, write pools)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
rope_append_kernel(__half* __restrict__ qkv, const int* __restrict__ cu_seqlens,
                   const int* __restrict__ ctx_lens, const int* __restrict__ block_tables,
                   __half* __restrict__ kpool, __half* __restrict__ vpool, int n_seqs, int max_blocks,
                   int n_heads, int n_kv_heads, int head_dim, float log2_theta, const float* __restrict__ inv_freq_table) {
    pdl_wait();
    pdl_launch();
    const int tkn = blockIdx.x;
    __shared__ float s_cos[64], s_sin[64];
    __shared__ int s_seq, s_pos;
    if (threadIdx.x == 0) {
        int seq, pos;
        if (cu_seqlens == nullptr) {
            seq = tkn;
            pos = ctx_lens[seq];
        } else {
            int lo = 0, hi = n_seqs;  // find seq with cu[seq] <= tkn < cu[seq+1]
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if (cu_seqlens[mid] <= tkn) lo = mid; else hi = mid;
            }
            seq = lo;
            pos = ctx_lens[seq] + (tkn - cu_seqlens[seq]);
        }
        s_seq = seq;
        s_pos = pos;
    }
    __syncthreads();
    const int seq = s_seq, pos = s_pos;
    const int half = head_dim >> 1;
    if ((int)threadIdx.x < half) {
        // inv_freq = theta^(-2i/hd), or the caller's table (scaled RoPE: Llama-3.1's rule is applied on the host)
        const float inv_freq = inv_freq_table ? inv_freq_table[threadIdx.x]
                                              : exp2f(-log2_theta * (float)(2 * threadIdx.x) / (float)head_dim);
        float sn, cs;
        sincosf((float)pos * inv_freq, &sn, &cs);
        s_cos[threadIdx.x] = cs;
        s_sin[threadIdx.x] = sn;
    }
    __syncthreads();

    const int qkv_dim = (n_heads + 2 * n_kv_heads) * head_dim;
    __half* row = qkv + (size_t)tkn * qkv_dim;
    const int blk = pos / KB_BLOCK_TOKENS, tib = pos % KB_BLOCK_TOKENS;
    const int phys = block_tables[(size_t)seq * max_blocks + blk];
    const int cph = head_dim >> 4;  // 16-byte chunks per half head

    // rotate q and k heads: work item = (head, chunk of the lower half)
    const int n_rot = (n_heads + n_kv_heads) * cph;
    for (int it = threadIdx.x; it < n_rot; it += blockDim.x) {
        const int head = it / cph, c = it % cph;
        __half* hp = row + (size_t)head * head_dim;
        uint4 lo4 = *reinterpret_cast<const uint4*>(hp + c * 8);
        uint4 hi4 = *reinterpret_cast<const uint4*>(hp + half + c * 8);
        uint32_t* lo2 = reinterpret_cast<uint32_t*>(&lo4);
        uint32_t* hi2 = reinterpret_cast<uint32_t*>(&hi4);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&lo2[j]));
            const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&hi2[j]));
            const int i0 = c * 8 + 2 * j;
            const float c0 = s_cos[i0], n0 = s_sin[i0], c1 = s_cos[i0 + 1], n1 = s_sin[i0 + 1];
            lo2[j] = pack_f16(a.x * c0 - b.x * n0, a.y * c1 - b.y * n1);
            hi2[j] = pack_f16(b.x * c0 + a.x * n0, b.y * c1 + a.y * n1);
        }
        if (head < n_heads) {
            *reinterpret_cast<uint4*>(hp + c * 8) = lo4;
            *reinterpret_cast<uint4*>(hp + half + c * 8) = hi4;
        } else {
            const int kvh = head - n_heads;
            __half* tile = kpool + ((size_t)phys * n_kv_heads + kvh) * KB_BLOCK_TOKENS * head_dim;
            *reinterpret_cast<uint4*>(tile + kv_swizzled_elem(tib, c * 8, head_dim)) = lo4;
            *reinterpret_cast<uint4*>(tile + kv_swizzled_elem(tib, half + c * 8, head_dim)) = hi4;
        }
    }
    // copy v heads
    const int n_v = n_kv_heads * (head_dim >> 3);
    const __half* vsrc = row + (size_t)(n_heads + n_kv_heads) * head_dim;
    for (int it = threadIdx.x; it < n_v; it += blockDim.x) {
        const int kvh = it / (head_dim >> 3), c = it % (head_dim >> 3);
        const uint4 v = *reinterpret_cast<const uint4*>(vsrc + (size_t)kvh * head_dim + c * 8);
        __half* tile = vpool + ((size_t)phys * n_kv_heads + kvh) * KB_BLOCK_TOKENS * head_dim;
        *reinterpret_cast<uint4*>(tile + kv_swizzled_elem(tib, c * 8, head_dim)) = v;
    }
}

// ---------------------------------------------------------------------------------------------
// sampling: argmax_j (logit_j / T + gumbel_j);  greedy when T <= 0 (ties -> lowest id).
//   bytes/sequence = vocab * 4
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
    x += 0x9e3779b97f4a7c15ull;
    x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
    return x ^ (x >> 31);
}

// Nucleus (top-p) threshold of one row, for the block that samples it. Candidates are visited through `scan`
// (the same visitor the sampling pass uses); v = logit / T. Returns the smallest v that stays a candidate: the
// candidates with the largest probabilities whose mass reaches `top_p` of the total (the usual "smallest set
// whose cumulative probability is >= top_p"; candidates within ~3e-5 in log-probability of the cut are kept too).
// Masses are summed in 40-bit fixed point with integer atomics, so the threshold does not depend on the order in
// which threads arrive — sampling with a seed stays reproducible.
constexpr int NUCLEUS_BINS = 1024;
constexpr float NUCLEUS_SPAN = 32.f;      // log-probability range resolved below the row maximum

template <typename Scan>
__device__ float nucleus_threshold(Scan& scan, float top_p, int banned, float inv_t) {
    // four private copies (one per warp quarter): a row's candidates crowd into the few bins near the maximum, and
    // 64-bit shared-memory atomics on one copy serialised there (512 x 128 256 logits, tools/elementwise_bench.py:
    // 843 us with one copy and a serial walk over the bins, 528 us with four copies and the parallel walk below;
    // plain temperature sampling of the same rows: 145 us, greedy: 65 us)
    constexpr int COPIES = 4;
    __shared__ unsigned long long hist_all[COPIES][NUCLEUS_BINS];
    unsigned long long* hist = hist_all[0];
    unsigned long long* mine = hist_all[(threadIdx.x >> 5) & (COPIES - 1)];
    __shared__ float s_red[32];
    __shared__ float s_lo;
    __shared__ unsigned long long s_above, s_target;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    // row maximum of v over the candidates
    float m = -INFINITY;
    scan([&](int j, float raw) { if (j != banned) m = fmaxf(m, raw * inv_t); });
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) s_red[warp] = m;
    __syncthreads();
    m = -INFINITY;
    for (int k = 0; k < nw; ++k) m = fmaxf(m, s_red[k]);
    __syncthreads();
    if (!(m > -INFINITY)) return -INFINITY;                   // no candidate at all
    auto mass = [](float u) { return (unsigned long long)(__expf(u) * 1099511627776.f); };   // 2^40 fixed point, u <= 0
    // level 0: bins over [-SPAN, 0] (anything lower falls into bin 0, u = 0 into the top bin); level 1: the bin in
    // which the cumulative mass (from the top) crosses the target, subdivided again. A candidate belongs to the
    // crossing bin iff its LEVEL-0 index says so — the same arithmetic both times, so no candidate is lost or
    // counted twice at a bin edge.
    const float w0 = NUCLEUS_SPAN / NUCLEUS_BINS, w1 = w0 / NUCLEUS_BINS;
    auto bin0 = [&](float u) {
        int b = (int)floorf((u + NUCLEUS_SPAN) / w0);
        return b < 0 ? 0 : (b >= NUCLEUS_BINS ? NUCLEUS_BINS - 1 : b);
    };
    __shared__ int s_bin;
    unsigned long long above = 0, target = 0;
    float lo_u = -NUCLEUS_SPAN;
    int cross = 0;
    for (int level = 0; level < 2; ++level) {
        for (int b = threadIdx.x; b < COPIES * NUCLEUS_BINS; b += blockDim.x) hist_all[0][b] = 0;
        __syncthreads();
        scan([&](int j, float raw) {
            if (j == banned) return;
            const float u = raw * inv_t - m;
            int b = bin0(u);
            if (level == 1) {
                if (b != cross) return;
                b = (int)floorf((u - lo_u) / w1);
                b = b < 0 ? 0 : (b >= NUCLEUS_BINS ? NUCLEUS_BINS - 1 : b);
            }
            atomicAdd(&mine[b], mass(u));
        });
        __syncthreads();
        for (int b = threadIdx.x; b < NUCLEUS_BINS; b += blockDim.x) {
            unsigned long long t = hist_all[0][b];
#pragma unroll
            for (int c = 1; c < COPIES; ++c) t += hist_all[c][b];
            hist[b] = t;
        }
        __syncthreads();
        if (blockDim.x == NUCLEUS_BINS) {
            // one thread per bin: inclusive sums from the most probable bin downwards (warp shuffles, then across the
            // 32 warps), the crossing bin = the FIRST bin from the top whose running sum
            // reaches the target. (A serial walk by one thread cost ~50 us per row.)
            const int r = NUCLEUS_BINS - 1 - (int)threadIdx.x;            // thread 0 owns the top bin
            const unsigned long long h = hist[r];
            unsigned long long run = h;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned long long up = __shfl_up_sync(0xffffffffu, run, o);
                if (lane >= o) run += up;
            }
            __shared__ unsigned long long s_warp[32];
            if (lane == 31) s_warp[warp] = run;
            __syncthreads();
            if (warp == 0) {
                unsigned long long w = s_warp[lane];
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const unsigned long long up = __shfl_up_sync(0xffffffffu, w, o);
                    if (lane >= o) w += up;
                }
                s_warp[lane] = w;
            }
            __syncthreads();
            if (warp > 0) run += s_warp[warp - 1];                         // mass of bins r .. top, this level
            if (threadIdx.x == 0) {
                if (level == 0) {
                    target = (unsigned long long)((double)s_warp[31] * (double)top_p);
                    if (target < 1) target = 1;
                    s_target = target;
                }
                s_bin = 0;                                                  // default: never reached -> lowest bin
                s_above = above + s_warp[31] - hist[0];
            }
            __syncthreads();
            target = s_target;
            // exactly one thread sees the crossing: its inclusive sum reaches the target, the exclusive one does not
            if (above + run >= target && above + run - h < target && r > 0) {
                s_bin = r;
                s_above = above + run - h;
            }
            __syncthreads();
            if (threadIdx.x == 0) s_lo = lo_u + (float)s_bin * (level == 0 ? w0 : w1);
        } else if (threadIdx.x == 0) {
            if (level == 0) {
                unsigned long long z = 0;
                for (int b = 0; b < NUCLEUS_BINS; ++b) z += hist[b];
                target = (unsigned long long)((double)z * (double)top_p);
                if (target < 1) target = 1;
            }
            unsigned long long acc = above;
            int b = NUCLEUS_BINS - 1;
            for (; b > 0; --b) {                               // from the most probable bin downwards
                if (acc + hist[b] >= target) break;
                acc += hist[b];
            }
            s_above = acc;                                     // mass strictly above bin b
            s_target = target;
            s_bin = b;
            s_lo = lo_u + (float)b * (level == 0 ? w0 : w1);
        }
        __syncthreads();
        above = s_above;
        target = s_target;
        lo_u = s_lo;
        cross = s_bin;
        __syncthreads();
    }
    return m + lo_u;       // lower edge of the crossing sub-bin: that whole sub-bin stays in
}

__global__ void __launch_bounds__(1024)
sample_kernel(const float* __restrict__ logits, int vocab, int v0, float temperature, uint64_t seed, uint64_t step,
              int banned, int lo, int hi, int extra_a, int extra_b, const float* __restrict__ seq_temperature,
              const uint64_t* __restrict__ seq_seed, float top_p, const float* __restrict__ seq_top_p,
              int* __restrict__ out_tokens, int* __restrict__ ctx_lens_advance, uint2* __restrict__ pair_out) {
    // `logits` holds `vocab` columns per row, the ids [v0, v0 + vocab) of the vocabulary (v0 = 0: all of it;
    // a vocabulary-sharded lm_head under tensor parallelism: this rank's slice, v0 a multiple of 4). Every id
    // below is a GLOBAL id — the Gumbel noise, the banned id and the tie rule (lowest id) are those of the
    // unsharded kernel, so the best of the ranks' local winners IS the unsharded winner. pair_out != nullptr:
    // write (winning value, id) for that exchange instead of the token (kb_collective.cu, tp_argmax_kernel).
    pdl_wait();
    pdl_launch();
    const int seq = blockIdx.x;
    const float* lg = logits + (size_t)seq * vocab - v0;
    if (seq_temperature) temperature = seq_temperature[seq];
    if (seq_seed) seed ^= seq_seed[seq];
    if (seq_top_p) top_p = seq_top_p[seq];
    const bool greedy = !(temperature > 0.f);
    const float inv_t = greedy ? 1.f : 1.f / temperature;
    const uint64_t base = splitmix64(seed ^ splitmix64(step * 0x100000001b3ull + (uint64_t)seq));
    // candidates: ids in [lo, hi) and the (at most two) extra ids. The 16-byte aligned middle [v_lo, v_hi) of the
    // range is read as float4, four loads in flight per thread (a full-vocabulary row is 0.5 MB: with one 4-byte
    // load per thread in flight the kernel ran at 1.6 TB/s); the ragged ends and the extras go one by one.
    int v_lo = lo, v_hi = lo;
    if ((vocab & 3) == 0 && ((hi & ~3) > ((lo + 3) & ~3))) {
        v_lo = (lo + 3) & ~3;
        v_hi = hi & ~3;
    }
    auto scan = [&](auto&& fn) {
        const float4* p4 = reinterpret_cast<const float4*>(lg + v_lo);
        const int n4 = (v_hi - v_lo) >> 2;
        const int bd = blockDim.x;
        for (int i0 = threadIdx.x; i0 < n4; i0 += 4 * bd) {
            float4 t[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (i0 + k * bd < n4) t[k] = p4[i0 + k * bd];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (i0 + k * bd < n4) {
                    const int j = v_lo + 4 * (i0 + k * bd);
                    fn(j, t[k].x);
                    fn(j + 1, t[k].y);
                    fn(j + 2, t[k].z);
                    fn(j + 3, t[k].w);
                }
            }
        }
        const int n_head = v_lo - lo, n_tail = hi - v_hi;        // v_lo == v_hi == lo: everything is "tail"
        const int n_edge = n_head + n_tail + (extra_a >= 0 ? 1 : 0) + (extra_b >= 0 ? 1 : 0);
        for (int c = threadIdx.x; c < n_edge; c += blockDim.x) {
            int j;
            if (c < n_head) j = lo + c;
            else if (c < n_head + n_tail) j = v_hi + (c - n_head);
            else if (c == n_head + n_tail && extra_a >= 0) j = extra_a;
            else j = extra_b;
            fn(j, lg[j]);
        }
    };
    // nucleus sampling: candidates below the top-p cut are dropped before the noise is added (block-uniform branch)
    float cut = -INFINITY;
    if (!greedy && top_p > 0.f && top_p < 1.f) cut = nucleus_threshold(scan, top_p, banned, inv_t);
    float best = -INFINITY;
    int best_i = 0x7fffffff;
    scan([&](int j, float raw) {
        float v = raw * inv_t;
        if (v < cut || j == banned) {
            v = -INFINITY;
        } else if (!greedy) {
            const uint64_t r = splitmix64(base + (uint64_t)j);
            const float u = ((float)(r >> 40) + 0.5f) * (1.0f / 16777216.0f);  // (0,1)
            v += -__logf(-__logf(u));
        }
        if (v > best || (v == best && j < best_i)) { best = v; best_i = j; }
    });
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
        if (ov > best || (ov == best && oi < best_i)) { best = ov; best_i = oi; }
    }
    __shared__ float sv[32];
    __shared__ int si[32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { sv[warp] = best; si[warp] = best_i; }
    __syncthreads();
    if (warp == 0) {
        const int nw = blockDim.x >> 5;
        best = lane < nw ? sv[lane] : -INFINITY;
        best_i = lane < nw ? si[lane] : 0x7fffffff;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
            if (ov > best || (ov == best && oi < best_i)) { best = ov; best_i = oi; }
        }
        if (lane == 0) {
            if (pair_out != nullptr) {
                pair_out[seq] = make_uint2(__float_as_uint(best), (uint32_t)best_i);
            } else {
                out_tokens[seq] = best_i == 0x7fffffff ? 0 : best_i;
                if (ctx_lens_advance) ctx_lens_advance[seq] += 1;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// select_tokens: next step's input ids without a host round trip. dst[i] = prev_out[src[i]] where
// src[i] >= 0 (the id sampled by the step still in flight), else dst[i] keeps the id the host staged.
// ---------------------------------------------------------------------------------------------
__global__ void select_tokens_kernel(int* __restrict__ dst, const int* __restrict__ prev_out,
                                     const int* __restrict__ src, int n, int n_prev) {
    pdl_wait();
    pdl_launch();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int s = src[i];
    if (s >= 0 && s < n_prev) dst[i] = prev_out[s];
}

// ---------------------------------------------------------------------------------------------
// hbm_read_probe: a read-only stream over a caller buffer (16-byte loads, eight in flight per thread, persistent
// grid), one 4-byte result per block so the loads cannot be optimised away. bench.py times it over a slice of
// the KV pool far larger than L2 to know what a kernel that ONLY reads can reach on this GPU — the driver's
// measured HBM peak is a copy (half reads, half writes), which the decode-attention kernel exceeds.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512)
hbm_read_probe_kernel(const uint4* __restrict__ src, size_t n16, uint32_t* __restrict__ sink) {
    uint32_t acc = 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 7 * stride < n16; i += 8 * stride) {
        uint4 v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = __ldcs(src + i + k * stride);      // streaming: no reuse
#pragma unroll
        for (int k = 0; k < 8; ++k) acc ^= v[k].x ^ v[k].y ^ v[k].z ^ v[k].w;
    }
    for (; i < n16; i += stride) {
        const uint4 v = __ldcs(src + i);
        acc ^= v.x ^ v.y ^ v.z ^ v.w;
    }
    acc ^= __shfl_xor_sync(0xffffffffu, acc, 16);
    if ((threadIdx.x & 31) == 0) atomicXor(&sink[blockIdx.x], acc);
}

int hbm_read_probe(const void* buf, size_t bytes, void* sink, size_t sink_bytes, cudaStream_t stream) {
    KB_CHECK(buf && sink && bytes >= 16 && (reinterpret_cast<uintptr_t>(buf) & 15) == 0,
             "kb_hbm_read_probe: need a 16-byte aligned buffer of at least 16 bytes");
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int blocks = sms * 4;
    KB_CHECK(sink_bytes >= (size_t)blocks * 4, "kb_hbm_read_probe: sink needs %d bytes", blocks * 4);
    hbm_read_probe_kernel<<<blocks, 512, 0, stream>>>(reinterpret_cast<const uint4*>(buf), bytes / 16,
                                                      reinterpret_cast<uint32_t*>(sink));
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

// ---------------------------------------------------------------------------------------------
// host launchers
// ---------------------------------------------------------------------------------------------
int select_tokens(int* dst, const int* prev_out, const int* src, int n, int n_prev, cudaStream_t stream) {
    KB_CHECK(n > 0 && n_prev > 0 && dst && prev_out && src, "kb_select_tokens: bad arguments");
    KB_CUDA(launch_kernel(select_tokens_kernel, dim3((n + 255) / 256), dim3(256), 0, stream, dst, prev_out, src, n, n_prev));
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

int embed(const int* tokens, const void* embed_w, float* x, int n, int hidden, int vocab, cudaStream_t stream) {
    KB_CHECK(n > 0 && hidden % 8 == 0, "kb_embed: bad shape n=%d hidden=%d", n, hidden);
    KB_CUDA(launch_kernel(embed_kernel, dim3(n), dim3(256), 0, stream, tokens, (const __nv_bfloat16*)embed_w, x, hidden, vocab));
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

int rmsnorm(float* x, const void* w, void* y, const int* last_rows, const void* delta_f16, int n, int hidden, float eps,
            int out_f16, cudaStream_t stream) {
    KB_CHECK(n > 0 && hidden % 4 == 0, "kb_rmsnorm: bad shape n=%d hidden=%d", n, hidden);
    KB_CUDA(launch_kernel(rmsnorm_kernel, dim3(n), dim3(256), 0, stream, x, (const __nv_bfloat16*)w, (uint16_t*)y, last_rows,
                                          (const __half*)delta_f16, hidden, eps, out_f16));
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

int rope_append(void* qkv, const int* cu, const int* ctx, const int* bt, void* kp, void* vp, int n_seqs,
                int n_tokens, int max_blocks, int n_heads, int n_kv_heads, int head_dim, float theta,
                const float* inv_freq, cudaStream_t stream) {
    KB_CHECK(n_tokens > 0 && n_seqs > 0, "kb_rope_append: empty batch");
    KB_CHECK(head_dim == 64 || head_dim == 128, "kb_rope_append: head_dim must be 64 or 128 (got %d)", head_dim);
    KB_CHECK(cu != nullptr || n_tokens == n_seqs, "kb_rope_append: cu_seqlens == NULL means one token per sequence");
    KB_CUDA(launch_kernel(rope_append_kernel, dim3(n_tokens), dim3(128), 0, stream, (__half*)qkv, cu, ctx, bt, (__half*)kp,
                                                      (__half*)vp, n_seqs, max_blocks, n_heads, n_kv_heads,
                                                      head_dim, log2f(theta), inv_freq));
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

// Slice form: `logits` holds the `n_local` vocabulary ids from `v0` on (of `vocab` in all); pair_out as in the kernel.
int sample_slice(const float* logits, int n_seqs, int vocab, int v0, int n_local, const kb_sampling* sp, int* out_tokens,
                 int* ctx_advance, void* pair_out, cudaStream_t stream) {
    KB_CHECK(n_seqs > 0 && vocab > 0 && sp != nullptr && (out_tokens != nullptr || pair_out != nullptr),
             "kb_sample: bad arguments");
    KB_CHECK(v0 >= 0 && n_local > 0 && v0 + n_local <= vocab && v0 % 4 == 0, "kb_sample: bad vocabulary slice [%d, +%d) of %d",
             v0, n_local, vocab);
    // nucleus sampling needs the probability mass of the WHOLE vocabulary: a rank that holds a slice cannot cut it
    KB_CHECK(n_local == vocab || (!(sp->top_p > 0.f && sp->top_p < 1.f) && sp->seq_top_p == nullptr),
             "kb_sample: top_p needs the whole vocabulary on one GPU (replicated lm_head)");
    int lo = 0, hi = vocab, ea = -1, eb = -1;
    if (sp->allowed_hi > sp->allowed_lo) {
        lo = sp->allowed_lo < 0 ? 0 : sp->allowed_lo;
        hi = sp->allowed_hi > vocab ? vocab : sp->allowed_hi;
        ea = (sp->extra_a >= 0 && sp->extra_a < vocab && (sp->extra_a < lo || sp->extra_a >= hi)) ? sp->extra_a : -1;
        eb = (sp->extra_b >= 0 && sp->extra_b < vocab && (sp->extra_b < lo || sp->extra_b >= hi) &&
              sp->extra_b != ea) ? sp->extra_b : -1;
        KB_CHECK(hi > lo, "kb_sample: empty allowed range");
    }
    // this slice's part of the candidates (possibly none: the kernel then reports -inf / no id)
    const int v1 = v0 + n_local;
    if (lo < v0) lo = v0;
    if (hi > v1) hi = v1;
    if (hi < lo) hi = lo;
    if (ea >= 0 && (ea < v0 || ea >= v1)) ea = -1;
    if (eb >= 0 && (eb < v0 || eb >= v1)) eb = -1;
    KB_CUDA(launch_kernel(sample_kernel, dim3(n_seqs), dim3(1024), 0, stream, logits, n_local, v0, sp->temperature, sp->seed,
                          sp->step, sp->banned_token, lo, hi, ea, eb, sp->seq_temperature, sp->seq_seed, sp->top_p,
                          sp->seq_top_p, out_tokens, ctx_advance, static_cast<uint2*>(pair_out)));
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

int sample(const float* logits, int n_seqs, int vocab, const kb_sampling* sp, int* out_tokens, int* ctx_advance,
           cudaStream_t stream) {
    KB_CHECK(out_tokens != nullptr, "kb_sample: bad arguments");
    return sample_slice(logits, n_seqs, vocab, 0, vocab, sp, out_tokens, ctx_advance, nullptr, stream);
}

}  // namespace kb

extern "C" int kb_embed(const int32_t* tokens, const void* embed_bf16, float* x, int32_t n, int32_t hidden,
                        int32_t vocab, kb_stream stream) {
    return kb::embed(tokens, embed_bf16, x, n, hidden, vocab, reinterpret_cast<cudaStream_t>(stream));
}
extern "C" int kb_rmsnorm(const float* x, const void* w_bf16, void* y, int32_t n, int32_t hidden, float eps,
                          int32_t out_f16, kb_stream stream) {
    return kb::rmsnorm(const_cast<float*>(x), w_bf16, y, nullptr, nullptr, n, hidden, eps, out_f16,
                       reinterpret_cast<cudaStream_t>(stream));
}
extern "C" int kb_rope_append(void* qkv_f16, const int32_t* cu_seqlens, const int32_t* ctx_lens,
                              const int32_t* block_tables, void* k_pool_layer, void* v_pool_layer, int32_t n_seqs,
                              int32_t n_tokens, int32_t max_blocks, int32_t n_heads, int32_t n_kv_heads,
                              int32_t head_dim, float rope_theta, kb_stream stream) {
    return kb::rope_append(qkv_f16, cu_seqlens, ctx_lens, block_tables, k_pool_layer, v_pool_layer, n_seqs, n_tokens,
                           max_blocks, n_heads, n_kv_heads, head_dim, rope_theta, nullptr,
                           reinterpret_cast<cudaStream_t>(stream));
}
extern "C" int kb_rope_append_freqs(void* qkv_f16, const int32_t* cu_seqlens, const int32_t* ctx_lens,
                                    const int32_t* block_tables, void* k_pool_layer, void* v_pool_layer, int32_t n_seqs,
                                    int32_t n_tokens, int32_t max_blocks, int32_t n_heads, int32_t n_kv_heads,
                                    int32_t head_dim, const float* inv_freq, kb_stream stream) {
    KB_CHECK(inv_freq != nullptr, "kb_rope_append_freqs: null frequency table");
    return kb::rope_append(qkv_f16, cu_seqlens, ctx_lens, block_tables, k_pool_layer, v_pool_layer, n_seqs, n_tokens,
                           max_blocks, n_heads, n_kv_heads, head_dim, 10000.f, inv_freq,
                           reinterpret_cast<cudaStream_t>(stream));
}
extern "C" int kb_sample(const float* logits, int32_t n_seqs, int32_t vocab, const kb_sampling* sampling,
                         int32_t* out_tokens, kb_stream stream) {
    return kb::sample(logits, n_seqs, vocab, sampling, out_tokens, nullptr, reinterpret_cast<cudaStream_t>(stream));
}
extern "C" int kb_sample_slice(const float* logits_slice, int32_t n_seqs, int32_t vocab, int32_t vocab_offset,
                               int32_t vocab_local, const kb_sampling* sampling, void* out_pairs, kb_stream stream) {
    return kb::sample_slice(logits_slice, n_seqs, vocab, vocab_offset, vocab_local, sampling, nullptr, nullptr, out_pairs,
                            reinterpret_cast<cudaStream_t>(stream));
}
extern "C" int kb_hbm_read_probe(const void* buf, size_t bytes, void* sink, size_t sink_bytes, kb_stream stream) {
    return kb::hbm_read_probe(buf, bytes, sink, sink_bytes, reinterpret_cast<cudaStream_t>(stream));
}
extern "C" int kb_select_tokens(int32_t* dst, const int32_t* prev_out, const int32_t* src_index, int32_t n,
                                int32_t n_prev, kb_stream stream) {
    return kb::select_tokens(dst, prev_out, src_index, n, n_prev, reinterpret_cast<cudaStream_t>(stream));
}
