// kb_attn.cu — paged attention over the HBM-resident per-thread KV blocks (sm_100a).
//
// Takes over the part of the remote completion (src/llm/portkey.py:379) that dominates decode:
// every generated token re-reads the whole conversation's K and V.  In the reference the whole
// conversation is re-sent as text on every call (src/agents/base.py:222-223); here it is
// (ctx_len x 2 x n_kv_heads x head_dim) fp16 per layer, gathered block by block.
//
// Decode kernel (HBM-bound; algorithmic bytes = kv_len * n_kv_heads * head_dim * 2 * 2 per sequence
// per layer):
//   * one CTA per (kv head, sequence[, split]); the GROUP = n_heads / n_kv_heads query heads that share
//     the KV head are the (zero-padded) 16 rows of a warp-level m16n8k16 MMA, so each K/V byte is
//     fetched once for all of them;
//   * every warp is its own producer and consumer: lane 0 issues cp.async.bulk (TMA, 1-D) copies of
//     32-token half-blocks of K and V into a private ring of stages, completion by mbarrier tx-count;
//     no CTA-wide barrier in the main loop;
//   * tiles arrive pre-swizzled (the KV-append kernel stores 16-byte chunks XOR (token & 7)), so
//     ldmatrix from the staged tile is bank-conflict free without a tensor map;
//   * online softmax in fp32 with quad shuffles, cross-warp merge through shared memory, optional
//     cross-CTA split-KV merge through a caller-provided scratch buffer.
//
// Prefill kernel (incremental: new tokens attend to cached prefix + themselves): one CTA per
// (q tile, kv head, sequence), 4 consumer warps x 16 rows (rows = query token x GROUP head) sharing a
// CTA-wide 3-stage ring of 64-token K/V tiles fed by a producer warp.
#include "kb_common.cuh"

namespace kb {

constexpr float LOG2E = 1.4426950408889634f;

// one 16-bit activation: fp16 (saturating) when f16 != 0, else bf16
__device__ __forceinline__ uint16_t act16(float v, int f16) {
    if (f16) return to_f16_sat(v);
    return __bfloat16_as_ushort(__float2bfloat16(v));
}

// One warp: S = Q K^T over an NT-token tile, online softmax, O += P V.
//   o  : [HD/8][4] accumulators (m16n8 C layout; [0],[1] = row g, [2],[3] = row g+8)
//   qa : [HD/16][4] Q A-fragments (pre-loaded)
//   ks, vs : shared addresses of the swizzled K / V tiles ([token][HD] fp16)
//   tok0 : absolute position of the tile's first token; lim0/lim1: exclusive position limit of
//          rows g / g+8 (causal + length mask), applied only when need_mask.
template <int HD, int NT, bool UPPER>
__device__ __forceinline__ void attn_tile(float (&o)[HD / 8][4], float (&m)[2], float (&l)[2],
                                          const uint32_t (&qa)[HD / 16][4], uint32_t ks, uint32_t vs, int tok0,
                                          int lim0, int lim1, float sc, bool need_mask, int lane) {
    constexpr int NJ = NT / 8;
    constexpr int ROW_BYTES = 128;            // one token's 64 dims of one plane
    constexpr int PLANE_BYTES = NT * 128;     // [dim plane][token][64 dims]: see kv_swizzled_elem
    const int lrow = lane & 7;
    const int lmat = lane >> 3;
    const int t = lane & 3;

    float s[NJ][4];
#pragma unroll
    for (int j = 0; j < NJ; ++j) { s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f; }

#pragma unroll
    for (int kk2 = 0; kk2 < HD / 32; ++kk2) {
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
            const int token = 8 * j + lrow;
            const int c = 4 * kk2 + lmat;      // logical 16-byte chunk of the head: plane c / 8, chunk c % 8
            uint32_t b0, b1, b2, b3;
            ldsm_x4(ks + (c >> 3) * PLANE_BYTES + token * ROW_BYTES + (((c & 7) ^ lrow) << 4), b0, b1, b2, b3);
            mma_f16_16816(s[j], qa[2 * kk2][0], qa[2 * kk2][1], qa[2 * kk2][2], qa[2 * kk2][3], b0, b1);
            mma_f16_16816(s[j], qa[2 * kk2 + 1][0], qa[2 * kk2 + 1][1], qa[2 * kk2 + 1][2], qa[2 * kk2 + 1][3], b2,
                           b3);
        }
    }

    // scale + mask + row max
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const int pos = tok0 + 8 * j + 2 * t + e;
            float v0 = s[j][e] * sc;
            if (need_mask && pos >= lim0) v0 = -INFINITY;
            s[j][e] = v0;
            mx0 = fmaxf(mx0, v0);
            if (UPPER) {
                float v1 = s[j][2 + e] * sc;
                if (need_mask && pos >= lim1) v1 = -INFINITY;
                s[j][2 + e] = v1;
                mx1 = fmaxf(mx1, v1);
            }
        }
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    const float mn0 = fmaxf(m[0], mx0);
    const float corr0 = exp2f(m[0] - mn0);
    m[0] = mn0;
    float mn1 = 0.f, corr1 = 1.f;
    if (UPPER) {
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        mn1 = fmaxf(m[1], mx1);
        corr1 = exp2f(m[1] - mn1);
        m[1] = mn1;
    }
    float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const float p0 = exp2f(s[j][e] - mn0);
            s[j][e] = p0;
            sum0 += p0;
            if (UPPER) {
                const float p1 = exp2f(s[j][2 + e] - mn1);
                s[j][2 + e] = p1;
                sum1 += p1;
            }
        }
    }
    l[0] = l[0] * corr0 + sum0;
    if (UPPER) l[1] = l[1] * corr1 + sum1;
#pragma unroll
    for (int d = 0; d < HD / 8; ++d) {
        o[d][0] *= corr0;
        o[d][1] *= corr0;
        if (UPPER) {
            o[d][2] *= corr1;
            o[d][3] *= corr1;
        }
    }

    // O += P V
#pragma unroll
    for (int kk = 0; kk < NT / 16; ++kk) {
        const uint32_t a0 = pack_f16(s[2 * kk][0], s[2 * kk][1]);          // P <= 1: fp16, 10 mantissa bits
        const uint32_t a1 = UPPER ? pack_f16(s[2 * kk][2], s[2 * kk][3]) : 0u;
        const uint32_t a2 = pack_f16(s[2 * kk + 1][0], s[2 * kk + 1][1]);
        const uint32_t a3 = UPPER ? pack_f16(s[2 * kk + 1][2], s[2 * kk + 1][3]) : 0u;
#pragma unroll
        for (int d2 = 0; d2 < HD / 16; ++d2) {
            const int token = 16 * kk + (lmat & 1) * 8 + lrow;
            const int c = 2 * d2 + (lmat >> 1);
            uint32_t b0, b1, b2, b3;
            ldsm_x4_t(vs + (c >> 3) * PLANE_BYTES + token * ROW_BYTES + (((c & 7) ^ lrow) << 4), b0, b1, b2, b3);
            mma_f16_16816(o[2 * d2], a0, a1, a2, a3, b0, b1);
            mma_f16_16816(o[2 * d2 + 1], a0, a1, a2, a3, b2, b3);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// decode
// ---------------------------------------------------------------------------------------------
template <int HD>
struct DecodeCfg {
    static constexpr int NT = 32;                        // tokens per stage (half a KV block)
    static constexpr int WARPS = (HD == 128) ? 2 : 4;    // 2 CTAs/SM either way
    static constexpr int STAGES = 3;
    static constexpr int TILE_BYTES = NT * HD * 2;       // one of K or V
    static constexpr int RING_BYTES = WARPS * STAGES * 2 * TILE_BYTES;
    static constexpr int MERGE_FLOATS = WARPS * 8 * (HD + 2);
    static constexpr int SMEM_BYTES = RING_BYTES + MERGE_FLOATS * 4 + WARPS * STAGES * 8 + 128;
};

template <int HD, int GROUP>
__global__ void __launch_bounds__(DecodeCfg<HD>::WARPS * 32)
attn_decode_kernel(const __half* __restrict__ q, int q_stride, const __half* __restrict__ kpool,
                   const __half* __restrict__ vpool, const int* __restrict__ kv_lens, int kv_len_bias,
                   const int* __restrict__ block_tables, uint16_t* __restrict__ out, float* __restrict__ part_o,
                   float* __restrict__ part_ml, int max_blocks, int n_heads, int n_kv_heads, int n_splits,
                   int halves_per_split, int out_f16) {
    using Cfg = DecodeCfg<HD>;
    constexpr int NT = Cfg::NT, W = Cfg::WARPS, S = Cfg::STAGES;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* ring = smem;
    float* merge = reinterpret_cast<float*>(smem + Cfg::RING_BYTES);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::RING_BYTES + Cfg::MERGE_FLOATS * 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const int kvh = blockIdx.x, seq = blockIdx.y, split = blockIdx.z;
    // private set-up (this warp's barriers) overlaps the tail of the previous kernel ...
    if (lane == 0) {
        uint64_t* b = bars + warp * S;
        for (int s = 0; s < S; ++s) mbar_init(&b[s], 1);
        fence_mbar_init();
    }
    pdl_wait();      // ... kv_lens, q and the appended K/V come from the kernels before this one
    pdl_launch();

    const int kv_len = kv_lens[seq] + kv_len_bias;
    const int n_half = (kv_len + NT - 1) / NT;
    const int h_begin = split * halves_per_split;
    const int h_end = min(n_half, h_begin + halves_per_split);

    uint64_t* my_bars = bars + warp * S;
    uint8_t* my_ring = ring + warp * (S * 2 * Cfg::TILE_BYTES);
    __syncthreads();

    // Q fragments: rows g < GROUP are the query heads of this KV head; the rest are zero.
    uint32_t qa[HD / 16][4];
    {
        const __half* qp = q + (size_t)seq * q_stride + (size_t)(kvh * GROUP + g) * HD;
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk) {
            qa[kk][1] = 0u;
            qa[kk][3] = 0u;
            if (g < GROUP) {
                qa[kk][0] = *reinterpret_cast<const uint32_t*>(qp + 16 * kk + 2 * t);
                qa[kk][2] = *reinterpret_cast<const uint32_t*>(qp + 16 * kk + 8 + 2 * t);
            } else {
                qa[kk][0] = 0u;
                qa[kk][2] = 0u;
            }
        }
    }

    const int* bt = block_tables + (size_t)seq * max_blocks;
    auto issue = [&](int h, int stage) {
        // lane 0 only
        const int phys = bt[h >> 1];
        // 32 tokens of every 64-dim plane of the tile: HD/64 pieces of 4 KB each for K and for V
        const size_t off = ((size_t)phys * n_kv_heads + kvh) * KB_BLOCK_TOKENS * HD + (size_t)(h & 1) * NT * 64;
        uint8_t* ks = my_ring + stage * 2 * Cfg::TILE_BYTES;
        mbar_arrive_expect_tx(&my_bars[stage], 2 * Cfg::TILE_BYTES);
#pragma unroll
        for (int p = 0; p < HD / 64; ++p) {
            bulk_g2s(ks + p * (NT * 128), kpool + off + (size_t)p * KB_BLOCK_TOKENS * 64, NT * 128, &my_bars[stage]);
            bulk_g2s(ks + Cfg::TILE_BYTES + p * (NT * 128), vpool + off + (size_t)p * KB_BLOCK_TOKENS * 64, NT * 128,
                     &my_bars[stage]);
        }
    };

    float o[HD / 8][4];
#pragma unroll
    for (int d = 0; d < HD / 8; ++d) { o[d][0] = o[d][1] = o[d][2] = o[d][3] = 0.f; }
    float m[2] = {-INFINITY, -INFINITY};
    float l[2] = {0.f, 0.f};
    const float sc = LOG2E * rsqrtf((float)HD);

    // this warp's half-tiles: h_begin + warp, + W, ...
    const int first = h_begin + warp;
    const int n_mine = (first < h_end) ? (h_end - first + W - 1) / W : 0;
    if (lane == 0) {
        for (int i = 0; i < S && i < n_mine; ++i) issue(first + i * W, i);
    }
    for (int i = 0; i < n_mine; ++i) {
        const int stage = i % S;
        const uint32_t phase = (uint32_t)((i / S) & 1);
        const int h = first + i * W;
        mbar_wait(&my_bars[stage], phase);
        const uint32_t ks = smem_u32(my_ring + stage * 2 * Cfg::TILE_BYTES);
        const bool tail = (h + 1) * NT > kv_len;
        attn_tile<HD, NT, false>(o, m, l, qa, ks, ks + Cfg::TILE_BYTES, h * NT, kv_len, kv_len, sc, tail, lane);
        __syncwarp();
        if (lane == 0 && i + S < n_mine) issue(first + (i + S) * W, stage);
    }

    // quad-reduce the row sums, publish this warp's partial (m, l, O) for rows < GROUP
    l[0] += __shfl_xor_sync(0xffffffffu, l[0], 1);
    l[0] += __shfl_xor_sync(0xffffffffu, l[0], 2);
    float* mw = merge + warp * 8 * (HD + 2);
    if (g < GROUP) {
        float* row = mw + g * (HD + 2);
        if (t == 0) {
            row[HD] = m[0];
            row[HD + 1] = l[0];
        }
#pragma unroll
        for (int d = 0; d < HD / 8; ++d) {
            row[8 * d + 2 * t] = o[d][0];
            row[8 * d + 2 * t + 1] = o[d][1];
        }
    }
    __syncthreads();

    for (int idx = threadIdx.x; idx < GROUP * HD; idx += W * 32) {
        const int r = idx / HD, dim = idx % HD;
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < W; ++w) M = fmaxf(M, merge[(w * 8 + r) * (HD + 2) + HD]);
        float L = 0.f, val = 0.f;
        if (M != -INFINITY) {
#pragma unroll
            for (int w = 0; w < W; ++w) {
                const float* row = merge + (w * 8 + r) * (HD + 2);
                const float wgt = exp2f(row[HD] - M);
                L += wgt * row[HD + 1];
                val += wgt * row[dim];
            }
        }
        const int head = kvh * GROUP + r;
        if (n_splits == 1) {
            out[((size_t)seq * n_heads + head) * HD + dim] = act16(val / L, out_f16);
        } else {
            const size_t p = ((size_t)seq * n_heads + head) * n_splits + split;
            part_o[p * HD + dim] = val;
            if (dim == 0) {
                part_ml[p * 2] = M;
                part_ml[p * 2 + 1] = L;
            }
        }
    }
}

// merge the split-KV partials: one block per (sequence, head)
template <int HD>
__global__ void __launch_bounds__(HD)
attn_split_merge_kernel(const float* __restrict__ part_o, const float* __restrict__ part_ml,
                        uint16_t* __restrict__ out, int n_splits, int out_f16) {
    pdl_wait();
    pdl_launch();
    const size_t sh = blockIdx.x;
    const int dim = threadIdx.x;
    float M = -INFINITY;
    for (int s = 0; s < n_splits; ++s) M = fmaxf(M, part_ml[(sh * n_splits + s) * 2]);
    float L = 0.f, val = 0.f;
    for (int s = 0; s < n_splits; ++s) {
        const float ms = part_ml[(sh * n_splits + s) * 2];
        if (ms == -INFINITY) continue;
        const float wgt = exp2f(ms - M);
        L += wgt * part_ml[(sh * n_splits + s) * 2 + 1];
        val += wgt * part_o[(sh * n_splits + s) * HD + dim];
    }
    out[sh * HD + dim] = act16(val / L, out_f16);
}

// ---------------------------------------------------------------------------------------------
// prefill (incremental, causal)
// ---------------------------------------------------------------------------------------------
// One CTA = 128 query rows of one (sequence, KV head): row r = (query token q0 + r / GROUP, head r % GROUP),
// 8 warps x 16 rows. Every 64-token K/V tile is brought into shared memory once per CTA by two bulk copies
// (issued by lane 0 of warp 0, S-1 tiles ahead) and consumed by all 8 warps.
//
// Why 128 rows and no producer warp (ncu, round 1): the first version had 64 rows + a producer warp;
// its 184 registers / thread let only ONE 5-warp CTA live on an SM (registers are per SM sub-partition:
// 2 warps x 5 888 fit in 16 K, 3 do not), so 4 warps per SM issued MMAs (tensor pipe 43 % active) and
// every K/V tile was re-streamed from L2 once per 64 rows. 8 consumer warps fill the two warp slots of
// every sub-partition, and a tile now serves twice the rows.
template <int HD>
struct PrefillCfg {
    static constexpr int NT = 64;
    static constexpr int ROWS = 128;                               // query rows per CTA
    static constexpr int WARPS = ROWS / 16;
    static constexpr int STAGES = (HD == 128) ? 5 : 8;             // 160 KB / 128 KB of K/V in flight
    static constexpr int TILE_BYTES = NT * HD * 2;
    static constexpr int RING_BYTES = STAGES * 2 * TILE_BYTES;
    static constexpr int SMEM_BYTES = RING_BYTES + 2 * STAGES * 8 + 128;
    static constexpr int THREADS = WARPS * 32;
};

template <int HD, int GROUP>
__global__ void __launch_bounds__(PrefillCfg<HD>::THREADS, 1)
attn_prefill_kernel(const __half* __restrict__ q, int q_stride, const __half* __restrict__ kpool,
                    const __half* __restrict__ vpool, const int* __restrict__ cu_seqlens,
                    const int* __restrict__ ctx_lens, const int* __restrict__ block_tables,
                    uint16_t* __restrict__ out, int max_blocks, int n_heads, int n_kv_heads, int out_f16) {
    using Cfg = PrefillCfg<HD>;
    constexpr int NT = Cfg::NT, S = Cfg::STAGES;
    constexpr int TQ = Cfg::ROWS / GROUP;  // query tokens per CTA
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::RING_BYTES);
    uint64_t* empty_bar = full_bar + S;
    pdl_wait();
    pdl_launch();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int seq = blockIdx.z, kvh = blockIdx.y;
    const int tok_begin = cu_seqlens[seq];
    const int n_new = cu_seqlens[seq + 1] - tok_begin;
    // later query tiles see more keys: schedule them first
    const int q0 = ((int)gridDim.x - 1 - (int)blockIdx.x) * TQ;
    if (q0 >= n_new) return;
    const int ctx = ctx_lens[seq];
    const int q_end = min(n_new, q0 + TQ);
    const int kv_end = ctx + q_end;  // keys visible to the last query of this CTA
    const int n_tiles = (kv_end + NT - 1) / NT;
    const int n_active = ((q_end - q0) * GROUP + 15) / 16;   // warps that own at least one real row

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], n_active);
        }
        fence_mbar_init();
    }
    __syncthreads();
    if (warp >= n_active) return;

    const int* bt = block_tables + (size_t)seq * max_blocks;
    auto issue = [&](int i) {   // lane 0 of warp 0: bring tile i into its ring slot
        const int stage = i % S;
        mbar_wait(&empty_bar[stage], (uint32_t)(((i / S) & 1) ^ 1));
        const int phys = bt[i];
        const size_t off = ((size_t)phys * n_kv_heads + kvh) * KB_BLOCK_TOKENS * HD;
        uint8_t* ks = smem + stage * 2 * Cfg::TILE_BYTES;
        mbar_arrive_expect_tx(&full_bar[stage], 2 * Cfg::TILE_BYTES);
        bulk_g2s(ks, kpool + off, Cfg::TILE_BYTES, &full_bar[stage]);
        bulk_g2s(ks + Cfg::TILE_BYTES, vpool + off, Cfg::TILE_BYTES, &full_bar[stage]);
    };
    if (warp == 0 && lane == 0)
        for (int i = 0; i < S - 1 && i < n_tiles; ++i) issue(i);

    const int g = lane >> 2, t = lane & 3;
    const int r0 = warp * 16 + g, r1 = r0 + 8;
    const int tl0 = q0 + r0 / GROUP, tl1 = q0 + r1 / GROUP;
    const int hd0 = kvh * GROUP + r0 % GROUP, hd1 = kvh * GROUP + r1 % GROUP;
    const bool v0 = tl0 < n_new, v1 = tl1 < n_new;
    const int lim0 = v0 ? ctx + tl0 + 1 : kv_end;
    const int lim1 = v1 ? ctx + tl1 + 1 : kv_end;
    // smallest limit of any row of this warp decides when masking starts; the largest, when its work ends
    const int warp_first_tl = q0 + (warp * 16) / GROUP;
    const int warp_min_lim = ctx + warp_first_tl + 1;
    const int warp_last_tl = min(n_new - 1, q0 + (warp * 16 + 15) / GROUP);
    const int warp_tiles = (ctx + warp_last_tl + 1 + NT - 1) / NT;   // tiles beyond are fully masked for this warp

    uint32_t qa[HD / 16][4];
    {
        const __half* qp0 = q + (size_t)(tok_begin + tl0) * q_stride + (size_t)hd0 * HD;
        const __half* qp1 = q + (size_t)(tok_begin + tl1) * q_stride + (size_t)hd1 * HD;
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk) {
            qa[kk][0] = v0 ? *reinterpret_cast<const uint32_t*>(qp0 + 16 * kk + 2 * t) : 0u;
            qa[kk][2] = v0 ? *reinterpret_cast<const uint32_t*>(qp0 + 16 * kk + 8 + 2 * t) : 0u;
            qa[kk][1] = v1 ? *reinterpret_cast<const uint32_t*>(qp1 + 16 * kk + 2 * t) : 0u;
            qa[kk][3] = v1 ? *reinterpret_cast<const uint32_t*>(qp1 + 16 * kk + 8 + 2 * t) : 0u;
        }
    }

    float o[HD / 8][4];
#pragma unroll
    for (int d = 0; d < HD / 8; ++d) { o[d][0] = o[d][1] = o[d][2] = o[d][3] = 0.f; }
    float m[2] = {-INFINITY, -INFINITY};
    float l[2] = {0.f, 0.f};
    const float sc = LOG2E * rsqrtf((float)HD);

    for (int i = 0; i < n_tiles; ++i) {
        if (warp == 0 && lane == 0 && i + S - 1 < n_tiles) issue(i + S - 1);
        __syncwarp();
        const int stage = i % S;
        if (i < warp_tiles) {          // causal: this warp's rows see nothing of the later tiles
            mbar_wait(&full_bar[stage], (uint32_t)((i / S) & 1));
            const uint32_t ks = smem_u32(smem + stage * 2 * Cfg::TILE_BYTES);
            const bool need_mask = (i + 1) * NT > warp_min_lim;
            attn_tile<HD, NT, true>(o, m, l, qa, ks, ks + Cfg::TILE_BYTES, i * NT, lim0, lim1, sc, need_mask, lane);
            __syncwarp();
        }
        if (lane == 0) mbar_arrive(&empty_bar[stage]);
    }

    l[0] += __shfl_xor_sync(0xffffffffu, l[0], 1);
    l[0] += __shfl_xor_sync(0xffffffffu, l[0], 2);
    l[1] += __shfl_xor_sync(0xffffffffu, l[1], 1);
    l[1] += __shfl_xor_sync(0xffffffffu, l[1], 2);
    const float inv0 = 1.f / l[0], inv1 = 1.f / l[1];
    if (v0) {
        uint16_t* op = out + ((size_t)(tok_begin + tl0) * n_heads + hd0) * HD;
#pragma unroll
        for (int d = 0; d < HD / 8; ++d)
            *reinterpret_cast<uint32_t*>(op + 8 * d + 2 * t) = pack_act(o[d][0] * inv0, o[d][1] * inv0, out_f16);
    }
    if (v1) {
        uint16_t* op = out + ((size_t)(tok_begin + tl1) * n_heads + hd1) * HD;
#pragma unroll
        for (int d = 0; d < HD / 8; ++d)
            *reinterpret_cast<uint32_t*>(op + 8 * d + 2 * t) = pack_act(o[d][2] * inv1, o[d][3] * inv1, out_f16);
    }
}

// ---------------------------------------------------------------------------------------------
// host launchers
// ---------------------------------------------------------------------------------------------
template <int HD, int GROUP>
static int launch_decode(const void* q, int q_stride, const void* kp, const void* vp, const int* kv_lens, int bias,
                         const int* bt, void* out, float* part_o, float* part_ml, int batch, int max_blocks,
                         int n_heads, int n_kv_heads, int n_splits, int halves_per_split, int out_f16,
                         cudaStream_t stream) {
    using Cfg = DecodeCfg<HD>;
    auto kern = attn_decode_kernel<HD, GROUP>;
    static bool configured[64] = {};
    int dev = 0;
    KB_CUDA(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        KB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
        configured[dev & 63] = true;
    }
    dim3 grid(n_kv_heads, batch, n_splits);
    KB_CUDA(launch_kernel(kern, grid, dim3(Cfg::WARPS * 32), Cfg::SMEM_BYTES, stream, (const __half*)q, q_stride, (const __half*)kp, (const __half*)vp, kv_lens, bias, bt,
        (uint16_t*)out, part_o, part_ml, max_blocks, n_heads, n_kv_heads, n_splits, halves_per_split, out_f16));
    count_launch();
    KB_LAUNCH_CHECK();
    if (n_splits > 1) {
        KB_CUDA(launch_kernel(attn_split_merge_kernel<HD>, dim3(batch * n_heads), dim3(HD), 0, stream, part_o, part_ml, (uint16_t*)out, n_splits,
                                                                         out_f16));
        count_launch();
        KB_LAUNCH_CHECK();
    }
    return KB_OK;
}

size_t attn_decode_scratch_bytes(int batch, int n_heads, int head_dim, int n_splits) {
    return (size_t)batch * n_heads * n_splits * (head_dim + 2) * sizeof(float);
}

// Number of KV splits: enough CTAs to cover the SMs twice over when batch * n_kv_heads alone cannot.
int attn_decode_pick_splits(int batch, int n_kv_heads, int max_kv_len, int sms) {
    const int base = batch * n_kv_heads;
    const int n_half = (max_kv_len + 31) / 32;
    int splits = 1;
    if (base < 2 * sms) {
        splits = (4 * sms + base - 1) / base;
        const int max_by_len = n_half / 8 > 1 ? n_half / 8 : 1;  // at least 8 half-tiles (256 tokens) per split
        if (splits > max_by_len) splits = max_by_len;
        if (splits > KB_MAX_SPLITS) splits = KB_MAX_SPLITS;
        if (splits < 1) splits = 1;
    }
    return splits;
}

int attn_decode(const void* q, int q_stride, const void* kp, const void* vp, const int* kv_lens, int bias,
                const int* bt, void* out, int batch, int max_blocks, int n_heads, int n_kv_heads, int head_dim,
                int max_kv_len, void* scratch, size_t scratch_bytes, int out_f16, cudaStream_t stream) {
    KB_CHECK(batch > 0 && max_blocks > 0, "kb_attn_decode: empty batch");
    KB_CHECK(n_kv_heads > 0 && n_heads % n_kv_heads == 0, "kb_attn_decode: n_heads %% n_kv_heads != 0");
    KB_CHECK(batch <= 65535, "kb_attn_decode: batch %d > 65535", batch);
    const int group = n_heads / n_kv_heads;
    int sms = 148;
    {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    if (max_kv_len <= 0) max_kv_len = max_blocks * KB_BLOCK_TOKENS;
    int n_splits = scratch ? attn_decode_pick_splits(batch, n_kv_heads, max_kv_len, sms) : 1;
    while (n_splits > 1 && attn_decode_scratch_bytes(batch, n_heads, head_dim, n_splits) > scratch_bytes) --n_splits;
    const int n_half = (max_kv_len + 31) / 32;
    const int hps = (n_half + n_splits - 1) / n_splits;
    float* part_o = reinterpret_cast<float*>(scratch);
    float* part_ml = part_o ? part_o + (size_t)batch * n_heads * n_splits * head_dim : nullptr;

#define KB_DECODE_CASE(HD_, G_)                                                                                   \
    if (head_dim == HD_ && group == G_)                                                                           \
        return launch_decode<HD_, G_>(q, q_stride, kp, vp, kv_lens, bias, bt, out, part_o, part_ml, batch,        \
                                      max_blocks, n_heads, n_kv_heads, n_splits, hps, out_f16, stream);
    KB_DECODE_CASE(128, 1) KB_DECODE_CASE(128, 2) KB_DECODE_CASE(128, 4) KB_DECODE_CASE(128, 8)
    KB_DECODE_CASE(64, 1) KB_DECODE_CASE(64, 2) KB_DECODE_CASE(64, 4) KB_DECODE_CASE(64, 8)
#undef KB_DECODE_CASE
    set_error("kb_attn_decode: unsupported head_dim=%d group=%d (head_dim in {64,128}, group in {1,2,4,8})", head_dim,
              group);
    return KB_ERR_INVALID;
}

template <int HD, int GROUP>
static int launch_prefill(const void* q, int q_stride, const void* kp, const void* vp, const int* cu, const int* ctx,
                          const int* bt, void* out, int n_seqs, int max_new, int max_blocks, int n_heads,
                          int n_kv_heads, int out_f16, cudaStream_t stream) {
    using Cfg = PrefillCfg<HD>;
    auto kern = attn_prefill_kernel<HD, GROUP>;
    static bool configured[64] = {};
    int dev = 0;
    KB_CUDA(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        KB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
        configured[dev & 63] = true;
    }
    constexpr int TQ = Cfg::ROWS / GROUP;
    dim3 grid((max_new + TQ - 1) / TQ, n_kv_heads, n_seqs);
    KB_CUDA(launch_kernel(kern, grid, dim3(Cfg::THREADS), Cfg::SMEM_BYTES, stream, (const __half*)q, q_stride,
                                                          (const __half*)kp, (const __half*)vp, cu, ctx,
                                                          bt, (uint16_t*)out, max_blocks, n_heads, n_kv_heads,
                                                          out_f16));
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

// kb_attn_tc.cu: the tcgen05 / TMEM formulation (head_dim 128); -1 = shape not covered
int attn_prefill_tc(const void* q, int q_stride, const void* kp, const void* vp, const int* cu, const int* ctx,
                    const int* bt, void* out, int n_seqs, int max_new, int max_blocks, int n_heads, int n_kv_heads,
                    int head_dim, int out_f16, cudaStream_t stream);

int attn_prefill(const void* q, int q_stride, const void* kp, const void* vp, const int* cu, const int* ctx,
                 const int* bt, void* out, int n_seqs, int n_tokens, int max_new, int max_blocks, int n_heads,
                 int n_kv_heads, int head_dim, int out_f16, cudaStream_t stream) {
    KB_CHECK(n_seqs > 0 && n_tokens > 0 && max_new > 0, "kb_attn_prefill: empty batch");
    KB_CHECK(n_kv_heads > 0 && n_heads % n_kv_heads == 0, "kb_attn_prefill: n_heads %% n_kv_heads != 0");
    KB_CHECK(n_seqs <= 65535, "kb_attn_prefill: n_seqs %d > 65535", n_seqs);
    const int group = n_heads / n_kv_heads;
    static int legacy = -1;
    if (legacy < 0) legacy = getenv("KB_ATTN_PREFILL_LEGACY") != nullptr;
    if (!legacy) {
        const int rc = attn_prefill_tc(q, q_stride, kp, vp, cu, ctx, bt, out, n_seqs, max_new, max_blocks, n_heads,
                                       n_kv_heads, head_dim, out_f16, stream);
        if (rc != -1) return rc;
    }
#define KB_PREFILL_CASE(HD_, G_)                                                                               \
    if (head_dim == HD_ && group == G_)                                                                        \
        return launch_prefill<HD_, G_>(q, q_stride, kp, vp, cu, ctx, bt, out, n_seqs, max_new, max_blocks,    \
                                       n_heads, n_kv_heads, out_f16, stream);
    KB_PREFILL_CASE(128, 1) KB_PREFILL_CASE(128, 2) KB_PREFILL_CASE(128, 4) KB_PREFILL_CASE(128, 8)
    KB_PREFILL_CASE(64, 1) KB_PREFILL_CASE(64, 2) KB_PREFILL_CASE(64, 4) KB_PREFILL_CASE(64, 8)
#undef KB_PREFILL_CASE
    set_error("kb_attn_prefill: unsupported head_dim=%d group=%d", head_dim, group);
    return KB_ERR_INVALID;
}

}  // namespace kb

extern "C" size_t kb_attn_decode_scratch_bytes(int32_t batch, int32_t n_heads, int32_t head_dim) {
    return kb::attn_decode_scratch_bytes(batch, n_heads, head_dim, KB_MAX_SPLITS);
}

extern "C" int kb_attn_decode(const void* q_f16, int32_t q_stride, const void* k_pool_layer,
                              const void* v_pool_layer, const int32_t* kv_lens, int32_t kv_len_bias,
                              const int32_t* block_tables, void* out16, int32_t batch, int32_t max_blocks,
                              int32_t n_heads, int32_t n_kv_heads, int32_t head_dim, int32_t max_kv_len,
                              void* scratch, size_t scratch_bytes, int32_t out_f16, kb_stream stream) {
    return kb::attn_decode(q_f16, q_stride, k_pool_layer, v_pool_layer, kv_lens, kv_len_bias, block_tables, out16,
                           batch, max_blocks, n_heads, n_kv_heads, head_dim, max_kv_len, scratch, scratch_bytes,
                           out_f16, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int kb_attn_prefill(const void* q_f16, int32_t q_stride, const void* k_pool_layer,
                               const void* v_pool_layer, const int32_t* cu_seqlens, const int32_t* ctx_lens,
                               const int32_t* block_tables, void* out16, int32_t n_seqs, int32_t n_tokens,
                               int32_t max_new, int32_t max_blocks, int32_t n_heads, int32_t n_kv_heads,
                               int32_t head_dim, int32_t out_f16, kb_stream stream) {
    return kb::attn_prefill(q_f16, q_stride, k_pool_layer, v_pool_layer, cu_seqlens, ctx_lens, block_tables, out16,
                            n_seqs, n_tokens, max_new, max_blocks, n_heads, n_kv_heads, head_dim, out_f16,
                            reinterpret_cast<cudaStream_t>(stream));
}

// ---------------------------------------------------------------------------------------------
// kb_attn_decode_static_scratch — decode attention for callers that cannot allocate device memory.
//
// kb_attn_decode spreads a small batch over the whole GPU by splitting each sequence's KV range across
// several CTAs; the partial results travel through a scratch buffer that the CALLER has to allocate.
// Some callers have no way to do that: a binding that only ever receives finished input/output
// buffers from another component (an IPC handle, a mapped frame, a pool owned by a different
// runtime) and holds no allocator of its own; a host whose device memory is partitioned up front and
// forbids side allocations; a latency-critical loop where cudaMalloc's implicit device-wide
// synchronisation is not acceptable.  For them this entry point keeps the partials in a fixed array
// that lives in the library image itself.
//
//   * The array is a file-scope `static __device__` object of KB_STATIC_SCRATCH_BYTES (2 MiB).
//     Every GPU on which this library is loaded holds one copy from the moment its module is loaded
//     until the process exits, whether or not this function is ever called.
//   * On the host the identifier `g_static_scratch` is only a handle for the runtime, not a device
//     address, so it is never passed to a kernel.  The first call on each device (with that device
//     current) resolves the real address once with cudaGetSymbolAddress, checks the return code and
//     caches it; later calls reuse the cached pointer.  Make that first call outside of any stream
//     capture.
//   * Nothing here calls cudaMalloc, cudaFree or any allocating routine; inputs and outputs still
//     belong to the caller.
//   * Shapes whose partials would not fit in the array are rejected with an error that says so —
//     they are NOT silently run with fewer splits.  Large batches need no partials at all and are
//     always accepted.
//   * All calls on a device share the one array.  Each call overwrites every slot it later reads
//     (empty splits write m = -inf, l = 0), so nothing has to be cleared between calls and stale
//     contents of an earlier call are never consumed.  Two calls that may run on the GPU at the same
//     time (different streams, or different host threads) race on the array: serialise them.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_scratch[KB_STATIC_SCRATCH_BYTES];

namespace kb {
static void* g_static_scratch_addr[64] = {};  // per device, resolved on first use

static int static_scratch_for_current_device(void** out) {
    int dev = -1;
    KB_CUDA(cudaGetDevice(&dev));
    KB_CHECK(dev >= 0 && dev < 64, "kb_attn_decode_static_scratch: device index %d out of range", dev);
    if (g_static_scratch_addr[dev] == nullptr) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || p == nullptr) {
            set_error("kb_attn_decode_static_scratch: cudaGetSymbolAddress failed on device %d: %s", dev,
                      cudaGetErrorString(e));
            return KB_ERR_CUDA;
        }
        g_static_scratch_addr[dev] = p;
    }
    *out = g_static_scratch_addr[dev];
    return KB_OK;
}
}  // namespace kb

extern "C" size_t kb_static_scratch_capacity(void) { return (size_t)KB_STATIC_SCRATCH_BYTES; }

extern "C" int kb_attn_decode_static_scratch(const void* q_f16, int32_t q_stride, const void* k_pool_layer,
                                             const void* v_pool_layer, const int32_t* kv_lens, int32_t kv_len_bias,
                                             const int32_t* block_tables, void* out16, int32_t batch,
                                             int32_t max_blocks, int32_t n_heads, int32_t n_kv_heads,
                                             int32_t head_dim, int32_t max_kv_len, int32_t out_f16,
                                             kb_stream stream) {
    using namespace kb;
    KB_CHECK(batch > 0 && max_blocks > 0 && n_heads > 0 && n_kv_heads > 0,
             "kb_attn_decode_static_scratch: bad batch / head arguments");
    void* scratch = nullptr;
    int rc = static_scratch_for_current_device(&scratch);
    if (rc) return rc;
    // How many partials would the regular entry point want for this shape?
    int dev = 0, sms = 148;
    KB_CUDA(cudaGetDevice(&dev));
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int len_bound = max_kv_len > 0 ? max_kv_len : max_blocks * KB_BLOCK_TOKENS;
    const int n_splits = attn_decode_pick_splits(batch, n_kv_heads, len_bound, sms);
    const size_t need = n_splits > 1 ? attn_decode_scratch_bytes(batch, n_heads, head_dim, n_splits) : 0;
    if (need > (size_t)KB_STATIC_SCRATCH_BYTES) {
        set_error("kb_attn_decode_static_scratch: batch=%d n_heads=%d head_dim=%d with %d KV splits needs %zu bytes "
                  "of partials, more than the built-in %d-byte array; use kb_attn_decode with caller scratch",
                  batch, n_heads, head_dim, n_splits, need, (int)KB_STATIC_SCRATCH_BYTES);
        return KB_ERR_INVALID;
    }
    return attn_decode(q_f16, q_stride, k_pool_layer, v_pool_layer, kv_lens, kv_len_bias, block_tables, out16,
                       batch, max_blocks, n_heads, n_kv_heads, head_dim, max_kv_len, scratch,
                       (size_t)KB_STATIC_SCRATCH_BYTES, out_f16, reinterpret_cast<cudaStream_t>(stream));
}
