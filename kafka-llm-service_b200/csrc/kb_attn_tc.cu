// kb_attn_tc.cu — causal prefill attention on the 5th-gen tensor cores (tcgen05 + TMEM), head_dim 128.
//
// One CTA = 128 query rows of one (sequence, KV head): row r = (query token q0 + r / GROUP, head r % GROUP),
// i.e. TMEM lane r. Per 64-token K/V tile j of the paged pool:
//
//   warp 9 (producer) : two bulk copies (K tile, V tile: 16 KB each, already in the canonical
//                       SWIZZLE_128B form — kv_swizzled_elem) into a 4-stage ring
//   warp 8 (MMA)      : S_j  = Q K_j^T        8 x tcgen05.mma 128x64x16, A = Q in TENSOR MEMORY,
//                                              B = K_j (shared, K-major)            -> TMEM S[j & 1]
//                       OT_j = P_j V_j         4 x tcgen05.mma 128x128x16, A = P_j in TENSOR MEMORY,
//                                              B = V_j (shared, MN-major descriptor) -> TMEM OT[j & 1]
//                       issue order S_0, S_1, OT_0, S_2, OT_1, ...  (the softmax of tile j+1 never waits
//                       for the P V product of tile j)
//   warps 0-7 (softmax): TWO threads per query row (warp w and w + 4 share the TMEM lanes of quarter w & 3):
//                       each takes 32 of the 64 S columns and 64 of the 128 output dims; tcgen05.ld S_j ->
//                       causal mask, row max (the two halves meet through shared memory), running sum in fp32,
//                       P_j = exp2((S_j - m_j) * scale) as fp16 pairs into tensor memory (tcgen05.st).
//                       Two warps per scheduler hide each other's TMEM / MUFU latencies; the first version (one
//                       thread per row) ran at 3 350 cycles per tile.
//
// O = sum_j P_j V_j ACCUMULATES IN TENSOR MEMORY over all tiles (one 128-column accumulator; the MMAs of tile j
// add to what tiles 0..j-1 left). The running maximum m is raised lazily — only when a tile's maximum exceeds
// it by more than 2^8 — so most tiles do not touch O at all; when a warp does raise it (the first tiles of a
// row, rarely afterwards) it waits for the P V product of the previous tile, multiplies its rows of O by
// exp2(m_old - m_new) in place (tcgen05.ld / tcgen05.st) and only then publishes P_j, which is what the MMA warp
// waits for before it issues O += P_j V_j. Round 1 computed every tile's product into a fresh accumulator and
// folded it into registers (two tcgen05.ld of 32 columns + 64 adds per thread per tile: ~300 of the ~770 ns a
// tile's softmax took, beside ~750 ns of MMA issue; ~980 ns per 128 x 64 tile, tensor pipe 26 % busy).
// TMEM: 2 x 64 (S) + 128 (O) + 64 (Q) + 2 x 32 (P) = 384 of 512 columns; shared memory only holds the K / V ring.
//
// Replaces the mma.sync prefill kernel of kb_attn.cu for head_dim 128 (that one stays for head_dim 64 and
// as the A/B reference: KB_ATTN_PREFILL_LEGACY=1). Reference counterpart: none (src/llm/portkey.py:379).
#include "kb_common.cuh"

#include <stdlib.h>

namespace kb {

namespace {

constexpr int TC_ROWS = 128;      // query rows per CTA = TMEM lanes
constexpr int TC_NT = 64;         // tokens per K/V tile = one KV block
constexpr int TC_HD = 128;
constexpr int TC_STAGES = 4;
constexpr int TC_TILE_BYTES = TC_NT * TC_HD * 2;          // 16 KB (K or V)
constexpr int TC_THREADS = 320;     // 8 softmax warps + MMA warp + producer warp
constexpr int TC_SOFTMAX_THREADS = 256;
constexpr int TC_SMEM = 1024 + TC_STAGES * 2 * TC_TILE_BYTES + 256 + 2 * 2 * TC_ROWS * 4;
constexpr float TC_LOG2E = 1.4426950408889634f;

// K-major, SWIZZLE_128B: rows of 128 bytes, 8-row groups 1024 bytes apart.
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// MN-major, SWIZZLE_128B (cute: Swizzle<3,4,3> o ((8,n),(8,k)) : ((1,LBO),(8,SBO)) in 16-byte units):
// a K row holds 64 MN elements (128 bytes); 8-row K groups are `sbo` bytes apart, 64-element MN chunks
// `lbo` bytes apart. For a V tile: rows = tokens, MN = head dims, lbo = plane stride, sbo = 1024.
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)(lbo >> 4) << 16;
    d |= (uint64_t)(sbo >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// fp16 x fp16 -> fp32 (A and B format fields 0 = fp16), M = 128; b_mn: B operand is MN-major.
__device__ __forceinline__ uint32_t idesc_f16(int n, int b_mn) {
    return (1u << 4) | ((uint32_t)b_mn << 16) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(TC_ROWS >> 4) << 24);
}

// The MMA and producer warps wait with one thread each; a tight try_wait loop in those warps competes for
// issue slots with the two softmax warps on the same scheduler (20 % of the samples of the first version were
// BRA / SYNCS / NOP), so they sleep between polls.
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) __nanosleep(64);
}

// D[tmem] (+)= A[tmem] * B[smem desc]: the A operand (Q, P) lives in tensor memory — lane = row, one 32-bit
// column per two consecutive K elements — so an MMA only fetches B from shared memory.
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// thread i of the warp writes N consecutive 32-bit columns of TMEM lane (base_lane + i)
__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ float ex2_approx(float x) {   // one MUFU.EX2; ex2(-inf) = 0
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// Per-tile timestamps of CTA (0,0,0) (profiling aid: `make EXTRA=-DKB_ATTN_TRACE`, tools/attn_trace.py).
#ifdef KB_ATTN_TRACE
__device__ unsigned long long g_attn_trace[64 * 8];
#define KB_ATRACE(tile, slot)                                                                          \
    do {                                                                                               \
        if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (tile) < 64) {                    \
            unsigned long long t__;                                                                    \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t__));                                    \
            g_attn_trace[(tile) * 8 + (slot)] = t__;                                                   \
        }                                                                                              \
    } while (0)
#else
#define KB_ATRACE(tile, slot) do { } while (0)
#endif

template <int GROUP>
__global__ void __launch_bounds__(TC_THREADS, 1)
attn_prefill_tc_kernel(const __half* __restrict__ q, int q_stride, const __half* __restrict__ kpool,
                       const __half* __restrict__ vpool, const int* __restrict__ cu_seqlens,
                       const int* __restrict__ ctx_lens, const int* __restrict__ block_tables,
                       uint16_t* __restrict__ out, int max_blocks, int n_heads, int n_kv_heads, int out_f16) {
    constexpr int TQ = TC_ROWS / GROUP;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* kv_s = smem;                                  // ring: stage x (K tile, V tile)
    uint64_t* kv_full = reinterpret_cast<uint64_t*>(kv_s + TC_STAGES * 2 * TC_TILE_BYTES);
    uint64_t* kv_empty = kv_full + TC_STAGES;
    uint64_t* s_full = kv_empty + TC_STAGES;               // [2]
    uint64_t* s_empty = s_full + 2;                        // [2]
    uint64_t* p_full = s_empty + 2;                        // [2]
    uint64_t* pv_done = p_full + 2;                        // [1] phase j completes when O += P_j V_j has retired
    uint64_t* q_ready = pv_done + 1;                       // [1]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(q_ready + 1);
    float* mx_x = reinterpret_cast<float*>(kv_s + TC_STAGES * 2 * TC_TILE_BYTES + 256);   // [tile parity][half][row]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 8 && lane == 0) {
        for (int s = 0; s < TC_STAGES; ++s) {
            mbar_init(&kv_full[s], 1);
            mbar_init(&kv_empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&s_full[a], 1);
            mbar_init(&s_empty[a], 8);
            mbar_init(&p_full[a], 8);
        }
        mbar_init(pv_done, 1);
        mbar_init(q_ready, 8);
        fence_mbar_init();
    } else if (warp == 9) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_launch();

    const int seq = blockIdx.z, kvh = blockIdx.y;
    const int tok_begin = cu_seqlens[seq];
    const int n_new = cu_seqlens[seq + 1] - tok_begin;
    const int q0 = ((int)gridDim.x - 1 - (int)blockIdx.x) * TQ;      // heavy (late) query tiles first
    const bool active = q0 < n_new;
    const int ctx = ctx_lens[seq];
    const int q_end = min(n_new, q0 + TQ);
    const int kv_end = ctx + q_end;
    const int n_tiles = active ? (kv_end + TC_NT - 1) / TC_NT : 0;
    const uint32_t s_tmem = tmem_base;                     // S[a]  at columns a * 64
    const uint32_t o_tmem = tmem_base + 128;               // O (accumulated over all tiles) at columns 128..255
    const uint32_t q_tmem = tmem_base + 384;               // Q as the A operand of S: 128 fp16 = 64 columns
    const uint32_t p_tmem = tmem_base + 448;               // P[a] as the A operand of OT: 64 fp16 = 32 columns each

    if (warp == 9) {
        // ===== producer =====
        if (lane == 0) {
            const int* bt = block_tables + (size_t)seq * max_blocks;
            for (int j = 0; j < n_tiles; ++j) {
                const int st = j % TC_STAGES;
                mbar_wait_relaxed(&kv_empty[st], (uint32_t)(((j / TC_STAGES) & 1) ^ 1));
                const size_t off = ((size_t)bt[j] * n_kv_heads + kvh) * KB_BLOCK_TOKENS * TC_HD;
                uint8_t* ks = kv_s + st * 2 * TC_TILE_BYTES;
                mbar_arrive_expect_tx(&kv_full[st], 2 * TC_TILE_BYTES);
                bulk_g2s(ks, kpool + off, TC_TILE_BYTES, &kv_full[st]);
                bulk_g2s(ks + TC_TILE_BYTES, vpool + off, TC_TILE_BYTES, &kv_full[st]);
            }
        }
    } else if (warp == 8) {
        // ===== MMA issuer =====
        if (lane == 0 && n_tiles > 0) {
            const uint32_t id_s = idesc_f16(TC_NT, 0);        // S  = Q K^T : N = 64 tokens, B K-major
            const uint32_t id_o = idesc_f16(TC_HD, 1);        // OT = P V   : N = 128 dims,  B MN-major
            mbar_wait(q_ready, 0);
            tc_fence_after();
            auto issue_s = [&](int j) {
                const int st = j % TC_STAGES, a = j & 1;
                mbar_wait_relaxed(&kv_full[st], (uint32_t)((j / TC_STAGES) & 1));
                mbar_wait_relaxed(&s_empty[a], (uint32_t)(((j >> 1) & 1) ^ 1));
                tc_fence_after();
                KB_ATRACE(j > 0 ? j - 1 : 0, 7);      // waits of issue_s(j) over (recorded in row j - 1)
                // one descriptor per tile; the 8 k-steps differ by constant offsets (16-byte units)
                const uint64_t kdesc = desc_kmajor(smem_u32(kv_s + st * 2 * TC_TILE_BYTES));
                const uint32_t d = s_tmem + (uint32_t)(a * 64);
#pragma unroll
                for (int p = 0; p < 2; ++p)
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_f16_ts(d, q_tmem + (uint32_t)(p * 32 + k * 8),
                                     kdesc + (uint64_t)((p * (TC_NT * 128) + k * 32) >> 4), id_s, (uint32_t)((p | k) != 0));
                umma_commit(&s_full[a]);
            };
            auto issue_o = [&](int j) {
                const int st = j % TC_STAGES, a = j & 1;
                mbar_wait_relaxed(&p_full[a], (uint32_t)((j >> 1) & 1));   // P_j written (and O rescaled if m moved)
                tc_fence_after();
                const uint64_t vdesc = desc_mnmajor(smem_u32(kv_s + st * 2 * TC_TILE_BYTES + TC_TILE_BYTES), TC_NT * 128, 1024);
                const uint32_t pa = p_tmem + (uint32_t)(a * 32);
#pragma unroll
                for (int k = 0; k < 4; ++k)       // 16 tokens per step: two 8-row groups of the V tile
                    umma_f16_ts(o_tmem, pa + (uint32_t)(k * 8), vdesc + (uint64_t)((k * 2048) >> 4), id_o,
                                (uint32_t)((j | k) != 0));               // tile 0 starts O, every later MMA adds to it
                umma_commit(pv_done);
                umma_commit(&kv_empty[st]);       // K_j and V_j are both consumed
            };
            issue_s(0);
            for (int j = 0; j < n_tiles; ++j) {
                KB_ATRACE(j, 0);                      // MMA thread: iteration j begins
                if (j + 1 < n_tiles) issue_s(j + 1);
                KB_ATRACE(j, 1);                      // S_{j+1} issued
                issue_o(j);
                KB_ATRACE(j, 2);                      // OT_j issued
            }
        }
    } else {
        // ===== softmax warps: two threads per query row (TMEM lane) =====
        const int quarter = warp & 3, half = warp >> 2;
        const int r = quarter * 32 + lane;                 // row = TMEM lane
        const int tl = q0 + r / GROUP;
        const int head = kvh * GROUP + r % GROUP;
        const bool valid = active && tl < n_new;
        const int lim = valid ? ctx + tl + 1 : kv_end;     // exclusive key limit of this row
        // this thread's 64 dims of the Q row -> tensor memory (A operand of S = Q K^T)
        if (active) {
            const uint4* src = reinterpret_cast<const uint4*>(q + (size_t)(tok_begin + min(tl, n_new - 1)) * q_stride +
                                                              (size_t)head * TC_HD) + half * 8;
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                uint32_t w[16];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint4 v = valid ? src[4 * c + i] : make_uint4(0u, 0u, 0u, 0u);
                    w[4 * i] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w;
                }
                tmem_st_32x16(q_tmem + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(half * 32 + c * 16), w);
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(q_ready);
        }
        float m = -INFINITY, l = 0.f;                      // m: shared by both halves; l: this half's columns
        const float sc = TC_LOG2E * rsqrtf((float)TC_HD);
        const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
        const uint32_t o_mine = o_tmem + lane_off + (uint32_t)(half * 64);   // this thread's 64 dims of its row of O

        for (int j = 0; j < n_tiles; ++j) {
            const int a = j & 1;
            // ---- softmax of tile j: this thread's 32 columns ----
            if (threadIdx.x == 0) KB_ATRACE(j, 3);     // softmax: iteration j begins (waits for S_j)
            mbar_wait(&s_full[a], (uint32_t)((j >> 1) & 1));
            tc_fence_after();
            if (threadIdx.x == 0) KB_ATRACE(j, 4);     // S_j ready
            uint32_t u[32];
            float s[32];
            tmem_ld_32x32(s_tmem + lane_off + (uint32_t)(a * 64 + half * 32), u);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_empty[a]);       // the MMA warp may overwrite S[a] with tile j + 2

            const int tok0 = j * TC_NT + half * 32;
            float mx = -INFINITY;
            if (tok0 + 32 > lim) {                         // causal / length mask: only the last tiles of a row
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float v = (tok0 + i >= lim) ? -INFINITY : __uint_as_float(u[i]);
                    s[i] = v;
                    mx = fmaxf(mx, v);
                }
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    s[i] = __uint_as_float(u[i]);
                    mx = fmaxf(mx, s[i]);
                }
            }
            float* xb = mx_x + a * (2 * TC_ROWS);
            xb[half * TC_ROWS + r] = mx;
            asm volatile("bar.sync 1, %0;" ::"n"(TC_SOFTMAX_THREADS) : "memory");
            // Running max in scaled (log2) units, raised LAZILY: it only moves when the tile's max exceeds it by
            // more than 8, so P stays below 2^8 (fp16: 10 mantissa bits from 2^-14 up, sums in fp32) and most
            // tiles leave the accumulated O alone (corr == 1).
            const float tile_mx = fmaxf(mx, xb[(half ^ 1) * TC_ROWS + r]) * sc;
            const float m_new = (tile_mx > m + 8.f) ? tile_mx : m;
            const float corr = ex2_approx(m - m_new);       // m = -inf on the first tile -> 0
            m = m_new;
            float sum = 0.f;
            uint32_t pw[16];                               // this thread's 32 probabilities, fp16 pairs
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float p0 = ex2_approx(fmaf(s[2 * i], sc, -m_new));
                const float p1 = ex2_approx(fmaf(s[2 * i + 1], sc, -m_new));
                sum += p0 + p1;
                pw[i] = pack_f16(p0, p1);
            }
            l = l * corr + sum;
            // P_j -> tensor memory (A operand of O += P_j V_j); the product of tile j - 2, the last reader of this
            // buffer, has retired: tile j - 1's P could only be published after S_{j-1}, issued behind it
            tmem_st_32x16(p_tmem + lane_off + (uint32_t)(a * 32 + half * 16), pw);
            // ---- the running max moved: rescale this warp's rows of O in tensor memory (rare) ----
            if (j > 0 && __any_sync(0xffffffffu, corr != 1.f)) {
                mbar_wait(pv_done, (uint32_t)((j - 1) & 1));       // O holds tiles 0 .. j-1, nothing is in flight
                tc_fence_after();
                if (threadIdx.x == 0) KB_ATRACE(j, 6);             // rescale begins
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    tmem_ld_32x32(o_mine + (uint32_t)(c * 32), u);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) u[i] = __float_as_uint(__uint_as_float(u[i]) * corr);
                    tmem_st_32x16(o_mine + (uint32_t)(c * 32), reinterpret_cast<const uint32_t(&)[16]>(u[0]));
                    tmem_st_32x16(o_mine + (uint32_t)(c * 32 + 16), reinterpret_cast<const uint32_t(&)[16]>(u[16]));
                }
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&p_full[a]);        // the MMA warp may issue O += P_j V_j
            if (threadIdx.x == 0) KB_ATRACE(j, 5);     // P_j written
        }
        if (n_tiles > 0) {
            mbar_wait(pv_done, (uint32_t)((n_tiles - 1) & 1));
            tc_fence_after();
            // the row sum is the sum of both halves' partial sums (same running max in both)
            float* xb = mx_x + (n_tiles & 1) * (2 * TC_ROWS);
            xb[half * TC_ROWS + r] = l;
            asm volatile("bar.sync 1, %0;" ::"n"(TC_SOFTMAX_THREADS) : "memory");
            l += xb[(half ^ 1) * TC_ROWS + r];
            const float inv = 1.f / l;
            uint16_t* op = out + ((size_t)(tok_begin + tl) * n_heads + head) * TC_HD + half * 64;
            uint32_t u[32];
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                tmem_ld_32x32(o_mine + (uint32_t)(c * 32), u);
                tmem_ld_wait();
                if (valid) {
#pragma unroll
                    for (int q4 = 0; q4 < 4; ++q4) {
                        uint4 w;
                        w.x = pack_act(__uint_as_float(u[8 * q4 + 0]) * inv, __uint_as_float(u[8 * q4 + 1]) * inv, out_f16);
                        w.y = pack_act(__uint_as_float(u[8 * q4 + 2]) * inv, __uint_as_float(u[8 * q4 + 3]) * inv, out_f16);
                        w.z = pack_act(__uint_as_float(u[8 * q4 + 4]) * inv, __uint_as_float(u[8 * q4 + 5]) * inv, out_f16);
                        w.w = pack_act(__uint_as_float(u[8 * q4 + 6]) * inv, __uint_as_float(u[8 * q4 + 7]) * inv, out_f16);
                        *reinterpret_cast<uint4*>(op + 32 * c + 8 * q4) = w;
                    }
                }
            }
            tc_fence_before();
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 9) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ---------------------------------------------------------------------------------------------
// 128-key K/V tiles, two MMA-issuing threads — the default since round 2.
//
// What bounds attn_prefill_tc_kernel is not the softmax arithmetic (removing every MUFU or F2FP left its time
// unchanged) but its ONE MMA-issuing thread: with the softmax warps reduced to their barrier protocol the kernel
// still took 78 % of its time. tools/mma_rate.cu: an MMA costs ~66 cycles of issue for any N <= 128 when one thread
// issues (so the eight 128 x 64 x 16 MMAs of S = Q K^T cost as much as eight 128 x 128 x 16 ones), and two threads
// issuing N = 128 MMAs reach the pipe's floor of 64 cycles each. Hence:
//   * a K/V tile is 128 keys (two KV blocks): S = Q K^T is 8 MMAs of 128 x 128 x 16, O += P V is 8 MMAs of
//     128 x 128 x 16 over 128 tokens — 16 instructions per 128 keys instead of 24, every one of full width, and the
//     barrier protocol runs once per 128 keys;
//   * warp 8 issues the S MMAs, warp 9 the P V MMAs, warps 10 / 11 feed a K ring (4 slots) and a V ring (2 slots) of
//     32 KB each (4 bulk copies of 8 KB per tile and ring: each 64-dim plane of the two blocks side by side, which
//     is again the canonical SWIZZLE_128B form);
//   * TMEM: S[2] (2 x 128 columns; each thread writes its 32 columns of P_j over the S values it alone has read) +
//     O (128) + Q (64) = 448; shared memory: 6 x 32 KB.
// S_{j+2} overwrites the buffer P_j lives in, and the two MMA threads are not ordered with each other: the S thread
// waits for O += P_j V_j to retire (pv_done[j & 1]) before it issues S_{j+2}.
// An odd number of KV blocks leaves the last tile half filled: its keys are masked (their S values are never
// used, P = 0) and the V rows behind them are zeroed once at kernel start so that 0 x stale-memory stays 0.
// ---------------------------------------------------------------------------------------------
constexpr int TC2_Q_BYTES = TC_ROWS * TC_HD * 2;           // 32 KB per query tile
// D[tmem] (+)= A[smem desc] * B[smem desc]
__device__ __forceinline__ void umma_f16_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}

constexpr int T128_NT = 128;
constexpr int T128_STAGES = 3;                              // (independent-halves variant below: one joint K/V ring)
constexpr int T128_HALF_BYTES = T128_NT * TC_HD * 2;        // K (or V) of one 128-key tile: 32 KB
constexpr int T128_THREADS = 352;                           // 8 softmax warps + S-MMA warp + PV-MMA warp + producer
// The default kernel keeps K and V in SEPARATE rings (tools/attn_trace.py: with one joint ring of three 64 KB stages the
// S thread waited 0.2 - 1.3 us per tile for K to land — a stage was only handed back when O += P V of its tile had
// retired, one tile period before the K of the tile that reuses it is needed, and a bulk copy from L2 / HBM takes about
// that long). A K slot is free as soon as S = Q K^T of its tile has retired, a V slot when O += P V has.
constexpr int T128_K_STAGES = 4;
constexpr int T128_V_STAGES = 2;
constexpr int T128D_THREADS = 384;                          // ... + a second producer warp (one per ring)
constexpr int T128_SMEM = 1024 + (T128_K_STAGES + T128_V_STAGES) * T128_HALF_BYTES + 256 + 2 * 2 * TC_ROWS * 4;

template <int GROUP>
__global__ void __launch_bounds__(T128D_THREADS, 1)
attn_prefill_tc128_kernel(const __half* __restrict__ q, int q_stride, const __half* __restrict__ kpool,
                          const __half* __restrict__ vpool, const int* __restrict__ cu_seqlens,
                          const int* __restrict__ ctx_lens, const int* __restrict__ block_tables,
                          uint16_t* __restrict__ out, int max_blocks, int n_heads, int n_kv_heads, int out_f16) {
    constexpr int TQ = TC_ROWS / GROUP;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    constexpr int KS = T128_K_STAGES, VS = T128_V_STAGES;
    uint8_t* k_s = smem;                                   // K ring: KS slots of 32 KB
    uint8_t* v_s = k_s + KS * T128_HALF_BYTES;             // V ring: VS slots of 32 KB
    uint64_t* k_full = reinterpret_cast<uint64_t*>(v_s + VS * T128_HALF_BYTES);
    uint64_t* k_empty = k_full + KS;
    uint64_t* v_full = k_empty + KS;
    uint64_t* v_empty = v_full + VS;
    uint64_t* s_full = v_empty + VS;                       // [2]
    uint64_t* p_full = s_full + 2;                         // [2]
    uint64_t* pv_done = p_full + 2;                        // [2]
    uint64_t* q_ready = pv_done + 2;                       // [1]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(q_ready + 1);
    float* mx_x = reinterpret_cast<float*>(v_s + VS * T128_HALF_BYTES + 256);   // [parity][half][row]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 8 && lane == 0) {
        for (int s = 0; s < KS; ++s) {
            mbar_init(&k_full[s], 1);
            mbar_init(&k_empty[s], 1);
        }
        for (int s = 0; s < VS; ++s) {
            mbar_init(&v_full[s], 1);
            mbar_init(&v_empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&s_full[a], 1);
            mbar_init(&p_full[a], 8);
            mbar_init(&pv_done[a], 1);
        }
        mbar_init(q_ready, 8);
        fence_mbar_init();
    } else if (warp == 9) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    } else if (warp < 8) {
        // V rows of the second block of every slot: finite before anything multiplies them by P = 0
        for (int st = 0; st < VS; ++st)
            for (int pl = 0; pl < 2; ++pl) {
                uint4* z = reinterpret_cast<uint4*>(v_s + st * T128_HALF_BYTES + pl * (T128_NT * 128) + 64 * 128);
                for (int i = threadIdx.x; i < 64 * 128 / 16; i += TC_SOFTMAX_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
            }
        fence_proxy_async();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_launch();

    const int seq = blockIdx.z, kvh = blockIdx.y;
    const int tok_begin = cu_seqlens[seq];
    const int n_new = cu_seqlens[seq + 1] - tok_begin;
    const int q0 = ((int)gridDim.x - 1 - (int)blockIdx.x) * TQ;      // heavy (late) query tiles first
    const bool active = q0 < n_new;
    const int ctx = ctx_lens[seq];
    const int q_end = min(n_new, q0 + TQ);
    const int kv_end = ctx + q_end;
    const int n_blk = active ? (kv_end + TC_NT - 1) / TC_NT : 0;     // 64-token KV blocks this CTA reads
    const int n_tiles = (n_blk + 1) >> 1;                            // 128-key tiles
    const uint32_t s_tmem = tmem_base;                     // S[a] at columns a * 128 (P_a over its first 64)
    const uint32_t o_tmem = tmem_base + 256;               // O, 128 columns
    const uint32_t q_tmem = tmem_base + 384;               // Q as the A operand of S: 128 fp16 = 64 columns

    if (warp >= 10) {
        // ===== producers: warp 10 feeds the K ring, warp 11 the V ring =====
        if (lane == 0) {
            const bool is_v = warp == 11;
            const int ns = is_v ? VS : KS;
            uint64_t* full = is_v ? v_full : k_full;
            uint64_t* empty = is_v ? v_empty : k_empty;
            uint8_t* ring = is_v ? v_s : k_s;
            const __half* pool = is_v ? vpool : kpool;
            const int* bt = block_tables + (size_t)seq * max_blocks;
            for (int j = 0; j < n_tiles; ++j) {
                const int st = j % ns;
                mbar_wait_relaxed(&empty[st], (uint32_t)(((j / ns) & 1) ^ 1));
                const int nb = min(2, n_blk - 2 * j);
                uint8_t* dst = ring + st * T128_HALF_BYTES;
                mbar_arrive_expect_tx(&full[st], (uint32_t)(nb * TC_TILE_BYTES));
                for (int b = 0; b < nb; ++b) {
                    const size_t off = ((size_t)bt[2 * j + b] * n_kv_heads + kvh) * KB_BLOCK_TOKENS * TC_HD;
#pragma unroll
                    for (int pl = 0; pl < 2; ++pl)         // plane pl of block b -> rows [64 b, 64 b + 64) of plane pl
                        bulk_g2s(dst + pl * (T128_NT * 128) + b * (64 * 128), pool + off + pl * (64 * 64), 64 * 128,
                                 &full[st]);
                }
            }
        }
    } else if (warp == 8) {
        // ===== S = Q K^T issuer =====
        if (lane == 0 && n_tiles > 0) {
            const uint32_t id_s = idesc_f16(T128_NT, 0);       // N = 128 keys, B K-major
            mbar_wait(q_ready, 0);
            tc_fence_after();
            for (int j = 0; j < n_tiles; ++j) {
                const int st = j % KS, a = j & 1;
                mbar_wait_relaxed(&k_full[st], (uint32_t)((j / KS) & 1));
                KB_ATRACE(j, 0);                               // S thread: K of tile j landed
                if (j >= 2) mbar_wait_relaxed(&pv_done[a], (uint32_t)(((j - 2) >> 1) & 1));   // P_{j-2} has been consumed
                tc_fence_after();
                KB_ATRACE(j, 1);                               // S thread: O += P V of tile j - 2 seen retired
                const uint64_t kdesc = desc_kmajor(smem_u32(k_s + st * T128_HALF_BYTES));
                const uint32_t d = s_tmem + (uint32_t)(a * 128);
#pragma unroll
                for (int pl = 0; pl < 2; ++pl)
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_f16_ts(d, q_tmem + (uint32_t)(pl * 32 + k * 8),
                                    kdesc + (uint64_t)((pl * (T128_NT * 128) + k * 32) >> 4), id_s, (uint32_t)((pl | k) != 0));
                umma_commit(&s_full[a]);
                umma_commit(&k_empty[st]);                     // K_j is consumed as soon as S_j has retired
                KB_ATRACE(j, 2);                               // S thread: S_j issued and committed
            }
        }
    } else if (warp == 9) {
        // ===== O += P V issuer =====
        if (lane == 0 && n_tiles > 0) {
            const uint32_t id_o = idesc_f16(TC_HD, 1);         // N = 128 dims, B MN-major
            for (int j = 0; j < n_tiles; ++j) {
                const int st = j % VS, a = j & 1;
                mbar_wait_relaxed(&v_full[st], (uint32_t)((j / VS) & 1));
                mbar_wait_relaxed(&p_full[a], (uint32_t)((j >> 1) & 1));   // P_j written (and O rescaled if m moved)
                tc_fence_after();
                KB_ATRACE(j, 5);                               // PV thread: V_j landed and P_j seen
                const uint64_t vdesc = desc_mnmajor(smem_u32(v_s + st * T128_HALF_BYTES), T128_NT * 128, 1024);
                const uint32_t pa = s_tmem + (uint32_t)(a * 128);
#pragma unroll
                for (int k = 0; k < 8; ++k)       // 16 tokens per step; P of keys [64 h, 64 h + 64) sits at columns 64 h ..
                    umma_f16_ts(o_tmem, pa + (uint32_t)((k >> 2) * 64 + (k & 3) * 8), vdesc + (uint64_t)((k * 2048) >> 4),
                                id_o, (uint32_t)((j | k) != 0));
                umma_commit(&pv_done[a]);
                umma_commit(&v_empty[st]);        // V_j is consumed
                KB_ATRACE(j, 6);                               // PV thread: O += P_j V_j issued and committed
            }
        }
    } else {
        // ===== softmax warps: two threads per query row (TMEM lane), 64 of the 128 S columns each =====
        const int quarter = warp & 3, half = warp >> 2;
        const int r = quarter * 32 + lane;
        const int tl = q0 + r / GROUP;
        const int head = kvh * GROUP + r % GROUP;
        const bool valid = active && tl < n_new;
        const int lim = valid ? ctx + tl + 1 : kv_end;
        const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
        if (active) {
            const uint4* src = reinterpret_cast<const uint4*>(q + (size_t)(tok_begin + min(tl, n_new - 1)) * q_stride +
                                                              (size_t)head * TC_HD) + half * 8;
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                uint32_t w[16];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint4 v = valid ? src[4 * c + i] : make_uint4(0u, 0u, 0u, 0u);
                    w[4 * i] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w;
                }
                tmem_st_32x16(q_tmem + lane_off + (uint32_t)(half * 32 + c * 16), w);
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(q_ready);
        }
        float m = -INFINITY, l = 0.f;
        const float sc = TC_LOG2E * rsqrtf((float)TC_HD);
        const uint32_t o_mine = o_tmem + lane_off + (uint32_t)(half * 64);

        for (int j = 0; j < n_tiles; ++j) {
            const int a = j & 1;
            mbar_wait(&s_full[a], (uint32_t)((j >> 1) & 1));
            tc_fence_after();
            if (threadIdx.x == 0) KB_ATRACE(j, 3);         // softmax: S_j seen complete
            const uint32_t s_mine = s_tmem + lane_off + (uint32_t)(a * 128 + half * 64);
            const int tok0 = j * T128_NT + half * 64;
            const bool masked = tok0 + 64 > lim;           // causal / length mask: only the last tiles of a row
            // OPTIMISTIC pass: exponentiate against the running max of the tiles before (m), 32 columns at a time,
            // without waiting for this tile's row max. After the first tiles of a row the max rarely moves by more than
            // the lazy threshold, so this is usually the final result — and with no barrier between the tensor-memory
            // loads and the MUFU work the warps drift apart, so one warp's tcgen05.ld (64 B/clk per SM: 64 KB of S per
            // tile) runs under another's exponentials instead of all warps doing first the one, then the other.
            uint32_t pk[32];                               // this thread's 64 probabilities, fp16 pairs
            float mx = -INFINITY, sum = 0.f;
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                uint32_t u[32];
                tmem_ld_32x32(s_mine + (uint32_t)(c * 32), u);
                tmem_ld_wait();
                if (masked) {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (tok0 + c * 32 + i >= lim) u[i] = 0xff800000u;     // -inf
                }
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const float s0 = __uint_as_float(u[2 * i]), s1 = __uint_as_float(u[2 * i + 1]);
                    mx = fmaxf(mx, fmaxf(s0, s1));
                    const float p0 = ex2_approx(fmaf(s0, sc, -m));           // m = -inf on the first tile: redone below
                    const float p1 = ex2_approx(fmaf(s1, sc, -m));
                    sum += p0 + p1;
                    pk[c * 16 + i] = pack_f16(p0, p1);
                }
            }
            float* xb = mx_x + a * (2 * TC_ROWS);
            xb[half * TC_ROWS + r] = mx;
            asm volatile("bar.sync 1, %0;" ::"n"(TC_SOFTMAX_THREADS) : "memory");   // both halves hold S_j's row max
            const float tile_mx = fmaxf(mx, xb[(half ^ 1) * TC_ROWS + r]) * sc;
            const float m_new = (tile_mx > m + 8.f) ? tile_mx : m;      // lazily raised running max (log2 units)
            const float corr = ex2_approx(m - m_new);                   // m = -inf on the first tile -> 0
            if (__any_sync(0xffffffffu, m_new != m)) {
                // the max moved for some row of this warp (both threads of a row decide alike): second pass against
                // the new max, from S itself — P has not been written over it yet
                sum = 0.f;
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    uint32_t u[32];
                    tmem_ld_32x32(s_mine + (uint32_t)(c * 32), u);
                    tmem_ld_wait();
                    if (masked) {
#pragma unroll
                        for (int i = 0; i < 32; ++i)
                            if (tok0 + c * 32 + i >= lim) u[i] = 0xff800000u;
                    }
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const float p0 = ex2_approx(fmaf(__uint_as_float(u[2 * i]), sc, -m_new));
                        const float p1 = ex2_approx(fmaf(__uint_as_float(u[2 * i + 1]), sc, -m_new));
                        sum += p0 + p1;
                        pk[c * 16 + i] = pack_f16(p0, p1);
                    }
                }
            }
            m = m_new;
            l = l * corr + sum;
            // P_j: this thread's 64 probabilities = 32 columns, written over the first half of ITS OWN 64 columns of
            // S[a] — columns no other thread reads, so a second pass of the row's other thread cannot be overtaken
            tmem_st_32x16(s_mine, reinterpret_cast<const uint32_t(&)[16]>(pk[0]));
            tmem_st_32x16(s_mine + 16u, reinterpret_cast<const uint32_t(&)[16]>(pk[16]));
            // ---- the running max moved: rescale this warp's rows of O in tensor memory (rare) ----
            if (j > 0 && __any_sync(0xffffffffu, corr != 1.f)) {
                mbar_wait(&pv_done[(j - 1) & 1], (uint32_t)(((j - 1) >> 1) & 1));   // O holds tiles 0 .. j-1
                tc_fence_after();
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    uint32_t u[32];
                    tmem_ld_32x32(o_mine + (uint32_t)(c * 32), u);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) u[i] = __float_as_uint(__uint_as_float(u[i]) * corr);
                    tmem_st_32x16(o_mine + (uint32_t)(c * 32), reinterpret_cast<const uint32_t(&)[16]>(u[0]));
                    tmem_st_32x16(o_mine + (uint32_t)(c * 32 + 16), reinterpret_cast<const uint32_t(&)[16]>(u[16]));
                }
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&p_full[a]);        // the PV warp may issue O += P_j V_j
            if (threadIdx.x == 0) KB_ATRACE(j, 4);         // softmax: P_j published (warp 0)
        }
        if (n_tiles > 0) {
            mbar_wait(&pv_done[(n_tiles - 1) & 1], (uint32_t)(((n_tiles - 1) >> 1) & 1));
            tc_fence_after();
            float* xb = mx_x + (n_tiles & 1) * (2 * TC_ROWS);
            xb[half * TC_ROWS + r] = l;
            asm volatile("bar.sync 1, %0;" ::"n"(TC_SOFTMAX_THREADS) : "memory");
            l += xb[(half ^ 1) * TC_ROWS + r];
            const float inv = 1.f / l;
            uint16_t* op = out + ((size_t)(tok_begin + tl) * n_heads + head) * TC_HD + half * 64;
            uint32_t u[32];
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                tmem_ld_32x32(o_mine + (uint32_t)(c * 32), u);
                tmem_ld_wait();
                if (valid) {
#pragma unroll
                    for (int q4 = 0; q4 < 4; ++q4) {
                        uint4 w;
                        w.x = pack_act(__uint_as_float(u[8 * q4 + 0]) * inv, __uint_as_float(u[8 * q4 + 1]) * inv, out_f16);
                        w.y = pack_act(__uint_as_float(u[8 * q4 + 2]) * inv, __uint_as_float(u[8 * q4 + 3]) * inv, out_f16);
                        w.z = pack_act(__uint_as_float(u[8 * q4 + 4]) * inv, __uint_as_float(u[8 * q4 + 5]) * inv, out_f16);
                        w.w = pack_act(__uint_as_float(u[8 * q4 + 6]) * inv, __uint_as_float(u[8 * q4 + 7]) * inv, out_f16);
                        *reinterpret_cast<uint4*>(op + 32 * c + 8 * q4) = w;
                    }
                }
            }
            tc_fence_before();
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 9) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ---------------------------------------------------------------------------------------------
// 128-key tiles with INDEPENDENT key halves (KB_ATTN_T128H, see attn_prefill_tc for which one is the default).
//
// In attn_prefill_tc128_kernel the two threads of a row still meet once per tile (row max through shared memory
// + bar.sync), so all 8 softmax warps read tensor memory at the same time (64 KB of S at 64 B/clk = 1 024 cycles)
// and then all exponentiate at the same time (16 384 exp2 at 16 / clk = 1 024 cycles): two throughput-bound
// phases in series. Here the two halves never meet inside the loop: warps 0-3 own keys [0, 64) of every tile,
// warps 4-7 keys [64, 128), each with its own running max, row sum and O accumulator in tensor memory
// (O_h = sum_j P_{j,h} V_{j,h}, a split of the keys as in split-KV decode) and its own barriers, so they drift
// apart and one half's tensor-memory reads run under the other half's MUFU phase. The halves are merged once, at
// the end: out = (O_0 2^(m_0 - M) + O_1 2^(m_1 - M)) / (l_0 2^(m_0 - M) + l_1 2^(m_1 - M)).
//   TMEM: S[2] (2 x 128) + O_0 (128) + O_1 (128) = 512 columns; P_{j,h} over the first 32 columns of ITS half of
//   S[j & 1] (no other thread ever touches them). Q moves to shared memory (32 KB, K-major SWIZZLE_128B, SS-form
//   MMA) and the K/V ring has 2 stages of 64 KB.
// ---------------------------------------------------------------------------------------------
constexpr int T128H_STAGES = 2;
constexpr int T128H_SMEM = 1024 + TC2_Q_BYTES + T128H_STAGES * 2 * T128_HALF_BYTES + 256 + 2 * TC_ROWS * 2 * 4;

template <int GROUP>
__global__ void __launch_bounds__(T128_THREADS, 1)
attn_prefill_tc128h_kernel(const __half* __restrict__ q, int q_stride, const __half* __restrict__ kpool,
                           const __half* __restrict__ vpool, const int* __restrict__ cu_seqlens,
                           const int* __restrict__ ctx_lens, const int* __restrict__ block_tables,
                           uint16_t* __restrict__ out, int max_blocks, int n_heads, int n_kv_heads, int out_f16) {
    constexpr int TQ = TC_ROWS / GROUP;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* q_s = smem;
    uint8_t* kv_s = q_s + TC2_Q_BYTES;
    uint64_t* kv_full = reinterpret_cast<uint64_t*>(kv_s + T128H_STAGES * 2 * T128_HALF_BYTES);
    uint64_t* kv_empty = kv_full + T128H_STAGES;
    uint64_t* s_full = kv_empty + T128H_STAGES;            // [2]
    uint64_t* p_full = s_full + 2;                         // [half][2]
    uint64_t* pv_done = p_full + 4;                        // [half][2]
    uint64_t* q_ready = pv_done + 4;                       // [1]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(q_ready + 1);
    float* ml_x = reinterpret_cast<float*>(kv_s + T128H_STAGES * 2 * T128_HALF_BYTES + 256);   // [half][row][m, l]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 8 && lane == 0) {
        for (int s = 0; s < T128H_STAGES; ++s) {
            mbar_init(&kv_full[s], 1);
            mbar_init(&kv_empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) mbar_init(&s_full[a], 1);
        for (int a = 0; a < 4; ++a) {
            mbar_init(&p_full[a], 4);
            mbar_init(&pv_done[a], 1);
        }
        mbar_init(q_ready, 8);
        fence_mbar_init();
    } else if (warp == 9) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    } else if (warp < 8) {
        // V rows of the second block of every stage: finite before anything multiplies them by P = 0
        for (int st = 0; st < T128H_STAGES; ++st)
            for (int pl = 0; pl < 2; ++pl) {
                uint4* z = reinterpret_cast<uint4*>(kv_s + st * 2 * T128_HALF_BYTES + T128_HALF_BYTES + pl * (T128_NT * 128) +
                                                    64 * 128);
                for (int i = threadIdx.x; i < 64 * 128 / 16; i += TC_SOFTMAX_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
            }
        fence_proxy_async();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_launch();

    const int seq = blockIdx.z, kvh = blockIdx.y;
    const int tok_begin = cu_seqlens[seq];
    const int n_new = cu_seqlens[seq + 1] - tok_begin;
    const int q0 = ((int)gridDim.x - 1 - (int)blockIdx.x) * TQ;      // heavy (late) query tiles first
    const bool active = q0 < n_new;
    const int ctx = ctx_lens[seq];
    const int q_end = min(n_new, q0 + TQ);
    const int kv_end = ctx + q_end;
    const int n_blk = active ? (kv_end + TC_NT - 1) / TC_NT : 0;     // 64-token KV blocks this CTA reads
    const int n_tiles = (n_blk + 1) >> 1;                            // 128-key tiles
    const uint32_t s_tmem = tmem_base;                     // S[a] at columns a * 128; half h owns columns [64 h, 64 h + 64)
    const uint32_t o_tmem = tmem_base + 256;               // O_h at columns 256 + 128 h

    if (warp == 10) {
        // ===== producer =====
        if (lane == 0) {
            const int* bt = block_tables + (size_t)seq * max_blocks;
            for (int j = 0; j < n_tiles; ++j) {
                const int st = j % T128H_STAGES;
                mbar_wait_relaxed(&kv_empty[st], (uint32_t)(((j / T128H_STAGES) & 1) ^ 1));
                const int nb = min(2, n_blk - 2 * j);
                uint8_t* ks = kv_s + st * 2 * T128_HALF_BYTES;
                mbar_arrive_expect_tx(&kv_full[st], (uint32_t)(nb * 2 * TC_TILE_BYTES));
                for (int b = 0; b < nb; ++b) {
                    const size_t off = ((size_t)bt[2 * j + b] * n_kv_heads + kvh) * KB_BLOCK_TOKENS * TC_HD;
#pragma unroll
                    for (int pl = 0; pl < 2; ++pl) {
                        bulk_g2s(ks + pl * (T128_NT * 128) + b * (64 * 128), kpool + off + pl * (64 * 64), 64 * 128,
                                 &kv_full[st]);
                        bulk_g2s(ks + T128_HALF_BYTES + pl * (T128_NT * 128) + b * (64 * 128), vpool + off + pl * (64 * 64),
                                 64 * 128, &kv_full[st]);
                    }
                }
            }
        }
    } else if (warp == 8) {
        // ===== S = Q K^T issuer =====
        if (lane == 0 && n_tiles > 0) {
            const uint32_t id_s = idesc_f16(T128_NT, 0);       // N = 128 keys, B K-major
            mbar_wait(q_ready, 0);
            tc_fence_after();
            const uint64_t qdesc = desc_kmajor(smem_u32(q_s));
            for (int j = 0; j < n_tiles; ++j) {
                const int st = j % T128H_STAGES, a = j & 1;
                mbar_wait_relaxed(&kv_full[st], (uint32_t)((j / T128H_STAGES) & 1));
                if (j >= 2) {                                  // both halves' P_{j-2} have been consumed
                    mbar_wait_relaxed(&pv_done[a], (uint32_t)(((j - 2) >> 1) & 1));
                    mbar_wait_relaxed(&pv_done[2 + a], (uint32_t)(((j - 2) >> 1) & 1));
                }
                tc_fence_after();
                const uint64_t kdesc = desc_kmajor(smem_u32(kv_s + st * 2 * T128_HALF_BYTES));
                const uint32_t d = s_tmem + (uint32_t)(a * 128);
#pragma unroll
                for (int pl = 0; pl < 2; ++pl)
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_f16_ss(d, qdesc + (uint64_t)((pl * (TC_ROWS * 128) + k * 32) >> 4),
                                    kdesc + (uint64_t)((pl * (T128_NT * 128) + k * 32) >> 4), id_s, (uint32_t)((pl | k) != 0));
                umma_commit(&s_full[a]);
            }
        }
    } else if (warp == 9) {
        // ===== O_h += P_h V_h issuer: whichever half publishes its P first goes first =====
        if (lane == 0 && n_tiles > 0) {
            const uint32_t id_o = idesc_f16(TC_HD, 1);         // N = 128 dims, B MN-major
            for (int j = 0; j < n_tiles; ++j) {
                const int st = j % T128H_STAGES, a = j & 1;
                const uint32_t par = (uint32_t)((j >> 1) & 1);
                const uint64_t vdesc = desc_mnmajor(smem_u32(kv_s + st * 2 * T128_HALF_BYTES + T128_HALF_BYTES),
                                                    T128_NT * 128, 1024);
                uint32_t todo = 3u;
                while (todo) {
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        if ((todo >> h) & 1u) {
                            if (!mbar_try_wait(&p_full[h * 2 + a], par)) continue;
                            tc_fence_after();
                            const uint32_t pa = s_tmem + (uint32_t)(a * 128 + h * 64);
                            const uint32_t d = o_tmem + (uint32_t)(h * 128);
#pragma unroll
                            for (int k = 0; k < 4; ++k)       // this half's 64 tokens, 16 per step
                                umma_f16_ts(d, pa + (uint32_t)(k * 8),
                                            vdesc + (uint64_t)((h * (64 * 128) + k * 2048) >> 4), id_o, (uint32_t)((j | k) != 0));
                            umma_commit(&pv_done[h * 2 + a]);
                            todo &= ~(1u << h);
                        }
                    }
                    if (todo == 3u) __nanosleep(32);
                }
                umma_commit(&kv_empty[st]);       // K_j and V_j are consumed by both halves
            }
        }
    } else {
        // ===== softmax warps: half h = warp / 4 owns keys [64 h, 64 h + 64) of every tile, one thread per row =====
        const int quarter = warp & 3, half = warp >> 2;
        const int r = quarter * 32 + lane;
        const int tl = q0 + r / GROUP;
        const int head = kvh * GROUP + r % GROUP;
        const bool valid = active && tl < n_new;
        const int lim = valid ? ctx + tl + 1 : kv_end;
        const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
        if (active) {
            const uint4* src = reinterpret_cast<const uint4*>(q + (size_t)(tok_begin + min(tl, n_new - 1)) * q_stride +
                                                              (size_t)head * TC_HD) + half * 8;
            uint8_t* dst = q_s + half * (TC_ROWS * 128) + r * 128;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const uint4 v = valid ? src[c] : make_uint4(0u, 0u, 0u, 0u);
                *reinterpret_cast<uint4*>(dst + ((c ^ (r & 7)) << 4)) = v;
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(q_ready);
        }
        float m = -INFINITY, l = 0.f;                      // of THIS half's keys only
        const float sc = TC_LOG2E * rsqrtf((float)TC_HD);
        const uint32_t o_h = o_tmem + lane_off + (uint32_t)(half * 128);

        for (int j = 0; j < n_tiles; ++j) {
            const int a = j & 1;
            mbar_wait(&s_full[a], (uint32_t)((j >> 1) & 1));
            tc_fence_after();
            uint32_t u0[32], u1[32];
            const uint32_t s_mine = s_tmem + lane_off + (uint32_t)(a * 128 + half * 64);
            tmem_ld_32x32(s_mine, u0);
            tmem_ld_32x32(s_mine + 32u, u1);
            tmem_ld_wait();
            const int tok0 = j * T128_NT + half * 64;
            if (tok0 + 64 > lim) {                         // causal / length mask: only the last tiles of a row
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    if (tok0 + i >= lim) u0[i] = 0xff800000u;             // -inf
                    if (tok0 + 32 + i >= lim) u1[i] = 0xff800000u;
                }
            }
            float mx = -INFINITY;
#pragma unroll
            for (int i = 0; i < 32; ++i) mx = fmaxf(mx, fmaxf(__uint_as_float(u0[i]), __uint_as_float(u1[i])));
            // lazily raised running max of this half (log2 units). A half whose keys are all masked so far keeps
            // m = -inf: use 0 as the reference then (every P is exp2(-inf) = 0) instead of -inf - -inf.
            const float tile_mx = mx * sc;
            const float m_new = (tile_mx > m + 8.f) ? tile_mx : m;
            const float corr = (m_new == -INFINITY) ? 1.f : ex2_approx(m - m_new);
            const float ref = (m_new == -INFINITY) ? 0.f : m_new;
            m = m_new;
            float sum = 0.f;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float p0 = ex2_approx(fmaf(__uint_as_float(u0[2 * i]), sc, -ref));
                const float p1 = ex2_approx(fmaf(__uint_as_float(u0[2 * i + 1]), sc, -ref));
                const float p2 = ex2_approx(fmaf(__uint_as_float(u1[2 * i]), sc, -ref));
                const float p3 = ex2_approx(fmaf(__uint_as_float(u1[2 * i + 1]), sc, -ref));
                sum += (p0 + p1) + (p2 + p3);
                u0[i] = pack_f16(p0, p1);
                u1[i] = pack_f16(p2, p3);
            }
            l = l * corr + sum;
            tmem_st_32x16(s_mine, reinterpret_cast<const uint32_t(&)[16]>(u0[0]));          // P_{j,h}: 32 columns over
            tmem_st_32x16(s_mine + 16u, reinterpret_cast<const uint32_t(&)[16]>(u1[0]));    // the S values just read
            if (j > 0 && __any_sync(0xffffffffu, corr != 1.f)) {       // this half's max moved: rescale ITS O (rare)
                mbar_wait(&pv_done[half * 2 + ((j - 1) & 1)], (uint32_t)(((j - 1) >> 1) & 1));
                tc_fence_after();
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    tmem_ld_32x32(o_h + (uint32_t)(c * 32), u0);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) u0[i] = __float_as_uint(__uint_as_float(u0[i]) * corr);
                    tmem_st_32x16(o_h + (uint32_t)(c * 32), reinterpret_cast<const uint32_t(&)[16]>(u0[0]));
                    tmem_st_32x16(o_h + (uint32_t)(c * 32 + 16), reinterpret_cast<const uint32_t(&)[16]>(u0[16]));
                }
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&p_full[half * 2 + a]);
        }
        if (n_tiles > 0) {
            // merge the halves: each thread publishes (m, l) of its half, then takes 64 of the row's 128 dims
            ml_x[(half * TC_ROWS + r) * 2] = m;
            ml_x[(half * TC_ROWS + r) * 2 + 1] = l;
            mbar_wait(&pv_done[(n_tiles - 1) & 1], (uint32_t)(((n_tiles - 1) >> 1) & 1));          // O_0 complete
            mbar_wait(&pv_done[2 + ((n_tiles - 1) & 1)], (uint32_t)(((n_tiles - 1) >> 1) & 1));    // O_1 complete
            tc_fence_after();
            asm volatile("bar.sync 1, %0;" ::"n"(TC_SOFTMAX_THREADS) : "memory");
            const float m0 = ml_x[r * 2], l0 = ml_x[r * 2 + 1];
            const float m1 = ml_x[(TC_ROWS + r) * 2], l1 = ml_x[(TC_ROWS + r) * 2 + 1];
            const float M = fmaxf(m0, m1);                 // finite: key 0 of the row is never masked
            const float w0 = (m0 == -INFINITY) ? 0.f : ex2_approx(m0 - M);
            const float w1 = (m1 == -INFINITY) ? 0.f : ex2_approx(m1 - M);
            const float inv = 1.f / (l0 * w0 + l1 * w1);
            const float f0 = w0 * inv, f1 = w1 * inv;
            uint16_t* op = out + ((size_t)(tok_begin + tl) * n_heads + head) * TC_HD + half * 64;
            const uint32_t o0 = o_tmem + lane_off + (uint32_t)(half * 64);          // O_0, this thread's 64 dims
            uint32_t u[32], v[32];
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                tmem_ld_32x32(o0 + (uint32_t)(c * 32), u);
                tmem_ld_32x32(o0 + 128u + (uint32_t)(c * 32), v);                   // O_1, same dims
                tmem_ld_wait();
                if (valid) {
#pragma unroll
                    for (int q4 = 0; q4 < 4; ++q4) {
                        float y[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e)
                            y[e] = (w0 != 0.f ? __uint_as_float(u[8 * q4 + e]) * f0 : 0.f) +
                                   (w1 != 0.f ? __uint_as_float(v[8 * q4 + e]) * f1 : 0.f);
                        uint4 w;
                        w.x = pack_act(y[0], y[1], out_f16);
                        w.y = pack_act(y[2], y[3], out_f16);
                        w.z = pack_act(y[4], y[5], out_f16);
                        w.w = pack_act(y[6], y[7], out_f16);
                        *reinterpret_cast<uint4*>(op + 32 * c + 8 * q4) = w;
                    }
                }
            }
            tc_fence_before();
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 9) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

template <int GROUP>
int launch_tc128h(const void* q, int q_stride, const void* kp, const void* vp, const int* cu, const int* ctx,
                  const int* bt, void* out, int n_seqs, int max_new, int max_blocks, int n_heads, int n_kv_heads,
                  int out_f16, cudaStream_t stream) {
    auto kern = attn_prefill_tc128h_kernel<GROUP>;
    static bool configured[64] = {};
    int dev = 0;
    KB_CUDA(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        KB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, T128H_SMEM));
        configured[dev & 63] = true;
    }
    constexpr int TQ = TC_ROWS / GROUP;
    dim3 grid((max_new + TQ - 1) / TQ, n_kv_heads, n_seqs);
    KB_CUDA(launch_kernel(kern, grid, dim3(T128_THREADS), T128H_SMEM, stream, (const __half*)q, q_stride,
                          (const __half*)kp, (const __half*)vp, cu, ctx, bt, (uint16_t*)out, max_blocks,
                          n_heads, n_kv_heads, out_f16));
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

template <int GROUP>
int launch_tc128(const void* q, int q_stride, const void* kp, const void* vp, const int* cu, const int* ctx,
                 const int* bt, void* out, int n_seqs, int max_new, int max_blocks, int n_heads, int n_kv_heads,
                 int out_f16, cudaStream_t stream) {
    auto kern = attn_prefill_tc128_kernel<GROUP>;
    static bool configured[64] = {};
    int dev = 0;
    KB_CUDA(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        KB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, T128_SMEM));
        configured[dev & 63] = true;
    }
    constexpr int TQ = TC_ROWS / GROUP;
    dim3 grid((max_new + TQ - 1) / TQ, n_kv_heads, n_seqs);
    KB_CUDA(launch_kernel(kern, grid, dim3(T128D_THREADS), T128_SMEM, stream, (const __half*)q, q_stride,
                          (const __half*)kp, (const __half*)vp, cu, ctx, bt, (uint16_t*)out, max_blocks,
                          n_heads, n_kv_heads, out_f16));
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

// ---------------------------------------------------------------------------------------------
// Two query tiles per CTA ("ping-pong"); opt-in with KB_ATTN_TC2=1 (see attn_prefill_tc for the measurement).
//
// One query tile keeps the tensor pipe ~27 % busy: per 64-token K/V tile the chain S = Q K^T -> tcgen05.ld of S
// (TMEM reads at 64 B/clk: 512 cycles for 128 x 64 fp32) -> 8 192 exp2 (MUFU, 16 / clk: 512 cycles) -> P ->
// O += P V is serial in one tile, ~1 900 cycles against 512 cycles of MMA. Here a CTA owns TWO consecutive query
// tiles A and B (256 rows) of the same (sequence, KV head): each has its own S double buffer and O accumulator in
// tensor memory and its own 8 softmax warps, and both share ONE K/V stream — every K/V tile is staged once for
// 256 query rows instead of 128. The MMA warp issues  S_A(j) S_B(j) | O_A += P V (j-1)  O_B += P V (j-1) | S_A(j+1) ...
// so while tile A's warps load, exponentiate and store, the tensor pipe works for tile B and vice versa.
//
//   TMEM (512 columns): per tile  S[2] (2 x 64)  +  O (128)  = 256;  P_j (fp16 pairs, 32 columns) is written over
//   the first half of S[j & 1] once both halves of a row have S_j in registers — S_{j+2}, the next writer of that
//   buffer, is issued after O += P_j V_j has read it (MMAs execute in issue order), so no "S buffer free" barrier.
//   Shared memory: Q_A, Q_B as K-major SWIZZLE_128B A operands (2 x 32 KB; the softmax threads write them, the
//   first round kept Q in tensor memory, which has no room for it any more) + a 4-stage K/V ring (128 KB).
//   Threads: 16 softmax warps (A: 0-7, B: 8-15; warp w and w + 4 of a tile share the rows of TMEM quarter w & 3
//   and take 32 of the 64 S columns / 64 of the 128 output dims each) + MMA warp + producer warp = 576.
// ---------------------------------------------------------------------------------------------
constexpr int TC2_THREADS = 576;
constexpr int TC2_SMEM = 1024 + 2 * TC2_Q_BYTES + TC_STAGES * 2 * TC_TILE_BYTES + 256 + 2 * 2 * 2 * TC_ROWS * 4;

template <int GROUP>
__global__ void __launch_bounds__(TC2_THREADS, 1)
attn_prefill_tc2_kernel(const __half* __restrict__ q, int q_stride, const __half* __restrict__ kpool,
                        const __half* __restrict__ vpool, const int* __restrict__ cu_seqlens,
                        const int* __restrict__ ctx_lens, const int* __restrict__ block_tables,
                        uint16_t* __restrict__ out, int max_blocks, int n_heads, int n_kv_heads, int out_f16) {
    constexpr int TQ = TC_ROWS / GROUP;                    // query tokens per tile
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* q_s = smem;                                   // Q_A | Q_B
    uint8_t* kv_s = q_s + 2 * TC2_Q_BYTES;                 // ring: stage x (K tile, V tile)
    uint64_t* kv_full = reinterpret_cast<uint64_t*>(kv_s + TC_STAGES * 2 * TC_TILE_BYTES);
    uint64_t* kv_empty = kv_full + TC_STAGES;
    uint64_t* s_full = kv_empty + TC_STAGES;               // [tile][2]
    uint64_t* p_full = s_full + 4;                         // [tile][2]
    uint64_t* pv_done = p_full + 4;                        // [tile]
    uint64_t* q_ready = pv_done + 2;                       // [1]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(q_ready + 1);
    float* mx_x = reinterpret_cast<float*>(kv_s + TC_STAGES * 2 * TC_TILE_BYTES + 256);   // [tile][parity][half][row]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int seq = blockIdx.z, kvh = blockIdx.y;
    if (warp == 16 && lane == 0) {
        for (int s = 0; s < TC_STAGES; ++s) {
            mbar_init(&kv_full[s], 1);
            mbar_init(&kv_empty[s], 1);
        }
        for (int a = 0; a < 4; ++a) {
            mbar_init(&s_full[a], 1);
            mbar_init(&p_full[a], 8);
        }
        mbar_init(&pv_done[0], 1);
        mbar_init(&pv_done[1], 1);
        mbar_init(q_ready, 16);
        fence_mbar_init();
    } else if (warp == 17) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_launch();

    const int tok_begin = cu_seqlens[seq];
    const int n_new = cu_seqlens[seq + 1] - tok_begin;
    const int pair = (int)gridDim.x - 1 - (int)blockIdx.x;          // heavy (late) query tiles first
    const int q0_pair = pair * 2 * TQ;
    const bool active = q0_pair < n_new;
    const bool b_active = q0_pair + TQ < n_new;                     // the second tile has at least one real row
    const int ctx = ctx_lens[seq];
    const int q_end = min(n_new, q0_pair + 2 * TQ);
    const int kv_end = ctx + q_end;                                 // keys visible to the last query of the pair
    const int n_tiles = active ? (kv_end + TC_NT - 1) / TC_NT : 0;

    if (warp == 17) {
        // ===== producer =====
        if (lane == 0) {
            const int* bt = block_tables + (size_t)seq * max_blocks;
            for (int j = 0; j < n_tiles; ++j) {
                const int st = j % TC_STAGES;
                mbar_wait_relaxed(&kv_empty[st], (uint32_t)(((j / TC_STAGES) & 1) ^ 1));
                const size_t off = ((size_t)bt[j] * n_kv_heads + kvh) * KB_BLOCK_TOKENS * TC_HD;
                uint8_t* ks = kv_s + st * 2 * TC_TILE_BYTES;
                mbar_arrive_expect_tx(&kv_full[st], 2 * TC_TILE_BYTES);
                bulk_g2s(ks, kpool + off, TC_TILE_BYTES, &kv_full[st]);
                bulk_g2s(ks + TC_TILE_BYTES, vpool + off, TC_TILE_BYTES, &kv_full[st]);
            }
        }
    } else if (warp == 16) {
        // ===== MMA issuer =====
        if (lane == 0 && n_tiles > 0) {
            const uint32_t id_s = idesc_f16(TC_NT, 0);         // S = Q K^T : N = 64 tokens, B K-major
            const uint32_t id_o = idesc_f16(TC_HD, 1);         // O += P V  : N = 128 dims,  B MN-major
            const int n_x = b_active ? 2 : 1;
            mbar_wait(q_ready, 0);
            tc_fence_after();
            auto issue_s = [&](int j) {                        // S_A(j), S_B(j)
                const int st = j % TC_STAGES, a = j & 1;
                mbar_wait_relaxed(&kv_full[st], (uint32_t)((j / TC_STAGES) & 1));
                tc_fence_after();
                const uint64_t kdesc = desc_kmajor(smem_u32(kv_s + st * 2 * TC_TILE_BYTES));
                for (int x = 0; x < n_x; ++x) {
                    const uint64_t qdesc = desc_kmajor(smem_u32(q_s + x * TC2_Q_BYTES));
                    const uint32_t d = tmem_base + (uint32_t)(x * 256 + a * 64);
#pragma unroll
                    for (int p = 0; p < 2; ++p)
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_f16_ss(d, qdesc + (uint64_t)((p * (TC_ROWS * 128) + k * 32) >> 4),
                                        kdesc + (uint64_t)((p * (TC_NT * 128) + k * 32) >> 4), id_s, (uint32_t)((p | k) != 0));
                    umma_commit(&s_full[x * 2 + a]);
                }
            };
            auto issue_o = [&](int j) {                        // O_A += P_A(j) V_j, O_B += P_B(j) V_j
                const int st = j % TC_STAGES, a = j & 1;
                const uint64_t vdesc = desc_mnmajor(smem_u32(kv_s + st * 2 * TC_TILE_BYTES + TC_TILE_BYTES), TC_NT * 128, 1024);
                for (int x = 0; x < n_x; ++x) {
                    mbar_wait_relaxed(&p_full[x * 2 + a], (uint32_t)((j >> 1) & 1));   // P written (O rescaled if m moved)
                    tc_fence_after();
                    const uint32_t pa = tmem_base + (uint32_t)(x * 256 + a * 64);      // P over the first half of S[a]
                    const uint32_t d = tmem_base + (uint32_t)(x * 256 + 128);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_f16_ts(d, pa + (uint32_t)(k * 8), vdesc + (uint64_t)((k * 2048) >> 4), id_o,
                                    (uint32_t)((j | k) != 0));
                    umma_commit(&pv_done[x]);
                }
                umma_commit(&kv_empty[st]);                    // K_j and V_j are consumed by both tiles
            };
            issue_s(0);
            for (int j = 0; j < n_tiles; ++j) {
                if (j + 1 < n_tiles) issue_s(j + 1);
                issue_o(j);
            }
        }
    } else {
        // ===== softmax warps: tile x = warp / 8, two threads per query row (TMEM lane) =====
        const int x = warp >> 3, w8 = warp & 7;
        const int quarter = w8 & 3, half = w8 >> 2;
        const int r = quarter * 32 + lane;                 // row of the tile = TMEM lane
        const int q0 = q0_pair + x * TQ;
        const int tl = q0 + r / GROUP;
        const int head = kvh * GROUP + r % GROUP;
        const bool tile_on = active && (x == 0 || b_active);
        const bool valid = tile_on && tl < n_new;
        const int lim = valid ? ctx + tl + 1 : kv_end;     // exclusive key limit of this row
        // this thread's 64 dims of its Q row -> shared memory, the K-major SWIZZLE_128B A operand of S = Q K^T:
        // [dim plane = half][row][64 dims], 16-byte chunks XOR-ed with (row & 7)
        if (active) {
            if (tile_on) {
                const uint4* src = reinterpret_cast<const uint4*>(q + (size_t)(tok_begin + min(tl, n_new - 1)) * q_stride +
                                                                  (size_t)head * TC_HD) + half * 8;
                uint8_t* dst = q_s + x * TC2_Q_BYTES + half * (TC_ROWS * 128) + r * 128;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const uint4 v = valid ? src[c] : make_uint4(0u, 0u, 0u, 0u);
                    *reinterpret_cast<uint4*>(dst + ((c ^ (r & 7)) << 4)) = v;
                }
                fence_proxy_async();                       // generic-proxy writes -> visible to the tensor core's reads
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(q_ready);
        }
        float m = -INFINITY, l = 0.f;                      // m: shared by both halves; l: this half's columns
        const float sc = TC_LOG2E * rsqrtf((float)TC_HD);
        const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
        const uint32_t t_x = tmem_base + lane_off + (uint32_t)(x * 256);
        const uint32_t o_mine = t_x + 128u + (uint32_t)(half * 64);          // this thread's 64 dims of its row of O
        float* mx_t = mx_x + x * (2 * 2 * TC_ROWS);
        const int n_my = tile_on ? n_tiles : 0;

        for (int j = 0; j < n_my; ++j) {
            const int a = j & 1;
            mbar_wait(&s_full[x * 2 + a], (uint32_t)((j >> 1) & 1));
            tc_fence_after();
            uint32_t u[32];
            float s[32];
            tmem_ld_32x32(t_x + (uint32_t)(a * 64 + half * 32), u);
            tmem_ld_wait();

            const int tok0 = j * TC_NT + half * 32;
            float mx = -INFINITY;
            if (tok0 + 32 > lim) {                         // causal / length mask: only the last tiles of a row
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float v = (tok0 + i >= lim) ? -INFINITY : __uint_as_float(u[i]);
                    s[i] = v;
                    mx = fmaxf(mx, v);
                }
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    s[i] = __uint_as_float(u[i]);
                    mx = fmaxf(mx, s[i]);
                }
            }
            float* xb = mx_t + a * (2 * TC_ROWS);
            xb[half * TC_ROWS + r] = mx;
            // both halves of every row of this tile have S_j in registers from here on (P may overwrite it)
            if (x == 0) asm volatile("bar.sync 1, 256;" ::: "memory");
            else asm volatile("bar.sync 2, 256;" ::: "memory");
            const float tile_mx = fmaxf(mx, xb[(half ^ 1) * TC_ROWS + r]) * sc;
            // lazily raised running max (log2 units): only when the tile's max exceeds it by more than 8. A row whose
            // keys of this tile are all masked (tile A at the pair's last K/V tile) keeps m and gets P = 0.
            const float m_new = (tile_mx > m + 8.f) ? tile_mx : m;
            const float corr = ex2_approx(m - m_new);       // m = -inf on the first tile -> 0
            m = m_new;
            float sum = 0.f;
            uint32_t pw[16];                               // this thread's 32 probabilities, fp16 pairs
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float p0 = ex2_approx(fmaf(s[2 * i], sc, -m_new));
                const float p1 = ex2_approx(fmaf(s[2 * i + 1], sc, -m_new));
                sum += p0 + p1;
                pw[i] = pack_f16(p0, p1);
            }
            l = l * corr + sum;
            tmem_st_32x16(t_x + (uint32_t)(a * 64 + half * 16), pw);       // P_j over the first 32 columns of S[a]
            // ---- the running max moved: rescale this warp's rows of O in tensor memory (rare) ----
            if (j > 0 && __any_sync(0xffffffffu, corr != 1.f)) {
                mbar_wait(&pv_done[x], (uint32_t)((j - 1) & 1));   // O holds tiles 0 .. j-1, nothing is in flight
                tc_fence_after();
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    tmem_ld_32x32(o_mine + (uint32_t)(c * 32), u);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) u[i] = __float_as_uint(__uint_as_float(u[i]) * corr);
                    tmem_st_32x16(o_mine + (uint32_t)(c * 32), reinterpret_cast<const uint32_t(&)[16]>(u[0]));
                    tmem_st_32x16(o_mine + (uint32_t)(c * 32 + 16), reinterpret_cast<const uint32_t(&)[16]>(u[16]));
                }
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&p_full[x * 2 + a]);    // the MMA warp may issue O += P_j V_j
        }
        if (n_my > 0) {
            mbar_wait(&pv_done[x], (uint32_t)((n_my - 1) & 1));
            tc_fence_after();
            // the row sum is the sum of both halves' partial sums (same running max in both)
            float* xb = mx_t + (n_my & 1) * (2 * TC_ROWS);
            xb[half * TC_ROWS + r] = l;
            if (x == 0) asm volatile("bar.sync 1, 256;" ::: "memory");
            else asm volatile("bar.sync 2, 256;" ::: "memory");
            l += xb[(half ^ 1) * TC_ROWS + r];
            const float inv = 1.f / l;
            uint16_t* op = out + ((size_t)(tok_begin + tl) * n_heads + head) * TC_HD + half * 64;
            uint32_t u[32];
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                tmem_ld_32x32(o_mine + (uint32_t)(c * 32), u);
                tmem_ld_wait();
                if (valid) {
#pragma unroll
                    for (int q4 = 0; q4 < 4; ++q4) {
                        uint4 w;
                        w.x = pack_act(__uint_as_float(u[8 * q4 + 0]) * inv, __uint_as_float(u[8 * q4 + 1]) * inv, out_f16);
                        w.y = pack_act(__uint_as_float(u[8 * q4 + 2]) * inv, __uint_as_float(u[8 * q4 + 3]) * inv, out_f16);
                        w.z = pack_act(__uint_as_float(u[8 * q4 + 4]) * inv, __uint_as_float(u[8 * q4 + 5]) * inv, out_f16);
                        w.w = pack_act(__uint_as_float(u[8 * q4 + 6]) * inv, __uint_as_float(u[8 * q4 + 7]) * inv, out_f16);
                        *reinterpret_cast<uint4*>(op + 32 * c + 8 * q4) = w;
                    }
                }
            }
            tc_fence_before();
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 17) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ---------------------------------------------------------------------------------------------
// Small-footprint variant: ONE query tile per CTA like attn_prefill_tc_kernel, but sized so that TWO CTAs fit an
// SM (256 TMEM columns: S[2] + O, P written over S, Q in shared memory; 32 KB of Q + a 2-stage K/V ring = 96 KB;
// 320 threads at <= 102 registers). A tile's chain  S -> tcgen05.ld -> exp2 -> P -> O += P V  is serial and takes
// ~1 900 cycles of which the tensor pipe works ~740 (tools/mma_rate.cu: a 128 x 64 x 16 MMA costs as much issue
// time as a 128 x 128 x 16 one, 62 cycles); two independent CTAs on the SM drift apart in phase, so one's
// tcgen05.ld / MUFU phase overlaps the other's MMAs — the overlap the two-tiles-per-CTA kernel did not get because
// its two tiles ran in lockstep.
// ---------------------------------------------------------------------------------------------
constexpr int TCS_STAGES = 2;
constexpr int TCS_SMEM = 1024 + TC2_Q_BYTES + TCS_STAGES * 2 * TC_TILE_BYTES + 256 + 2 * 2 * TC_ROWS * 4;

template <int GROUP>
__global__ void __launch_bounds__(TC_THREADS, 2)
attn_prefill_tcs_kernel(const __half* __restrict__ q, int q_stride, const __half* __restrict__ kpool,
                        const __half* __restrict__ vpool, const int* __restrict__ cu_seqlens,
                        const int* __restrict__ ctx_lens, const int* __restrict__ block_tables,
                        uint16_t* __restrict__ out, int max_blocks, int n_heads, int n_kv_heads, int out_f16) {
    constexpr int TQ = TC_ROWS / GROUP;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* q_s = smem;
    uint8_t* kv_s = q_s + TC2_Q_BYTES;
    uint64_t* kv_full = reinterpret_cast<uint64_t*>(kv_s + TCS_STAGES * 2 * TC_TILE_BYTES);
    uint64_t* kv_empty = kv_full + TCS_STAGES;
    uint64_t* s_full = kv_empty + TCS_STAGES;              // [2]
    uint64_t* p_full = s_full + 2;                         // [2]
    uint64_t* pv_done = p_full + 2;                        // [1]
    uint64_t* q_ready = pv_done + 1;                       // [1]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(q_ready + 1);
    float* mx_x = reinterpret_cast<float*>(kv_s + TCS_STAGES * 2 * TC_TILE_BYTES + 256);   // [parity][half][row]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 8 && lane == 0) {
        for (int s = 0; s < TCS_STAGES; ++s) {
            mbar_init(&kv_full[s], 1);
            mbar_init(&kv_empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&s_full[a], 1);
            mbar_init(&p_full[a], 8);
        }
        mbar_init(pv_done, 1);
        mbar_init(q_ready, 8);
        fence_mbar_init();
    } else if (warp == 9) {
        tmem_alloc(tmem_slot, 256);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_launch();

    const int seq = blockIdx.z, kvh = blockIdx.y;
    const int tok_begin = cu_seqlens[seq];
    const int n_new = cu_seqlens[seq + 1] - tok_begin;
    const int q0 = ((int)gridDim.x - 1 - (int)blockIdx.x) * TQ;      // heavy (late) query tiles first
    const bool active = q0 < n_new;
    const int ctx = ctx_lens[seq];
    const int q_end = min(n_new, q0 + TQ);
    const int kv_end = ctx + q_end;
    const int n_tiles = active ? (kv_end + TC_NT - 1) / TC_NT : 0;

    if (warp == 9) {
        if (lane == 0) {
            const int* bt = block_tables + (size_t)seq * max_blocks;
            for (int j = 0; j < n_tiles; ++j) {
                const int st = j % TCS_STAGES;
                mbar_wait_relaxed(&kv_empty[st], (uint32_t)(((j / TCS_STAGES) & 1) ^ 1));
                const size_t off = ((size_t)bt[j] * n_kv_heads + kvh) * KB_BLOCK_TOKENS * TC_HD;
                uint8_t* ks = kv_s + st * 2 * TC_TILE_BYTES;
                mbar_arrive_expect_tx(&kv_full[st], 2 * TC_TILE_BYTES);
                bulk_g2s(ks, kpool + off, TC_TILE_BYTES, &kv_full[st]);
                bulk_g2s(ks + TC_TILE_BYTES, vpool + off, TC_TILE_BYTES, &kv_full[st]);
            }
        }
    } else if (warp == 8) {
        if (lane == 0 && n_tiles > 0) {
            const uint32_t id_s = idesc_f16(TC_NT, 0);
            const uint32_t id_o = idesc_f16(TC_HD, 1);
            mbar_wait(q_ready, 0);
            tc_fence_after();
            const uint64_t qdesc = desc_kmajor(smem_u32(q_s));
            auto issue_s = [&](int j) {
                const int st = j % TCS_STAGES, a = j & 1;
                mbar_wait_relaxed(&kv_full[st], (uint32_t)((j / TCS_STAGES) & 1));
                tc_fence_after();
                const uint64_t kdesc = desc_kmajor(smem_u32(kv_s + st * 2 * TC_TILE_BYTES));
                const uint32_t d = tmem_base + (uint32_t)(a * 64);
#pragma unroll
                for (int p = 0; p < 2; ++p)
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_f16_ss(d, qdesc + (uint64_t)((p * (TC_ROWS * 128) + k * 32) >> 4),
                                    kdesc + (uint64_t)((p * (TC_NT * 128) + k * 32) >> 4), id_s, (uint32_t)((p | k) != 0));
                umma_commit(&s_full[a]);
            };
            auto issue_o = [&](int j) {
                const int st = j % TCS_STAGES, a = j & 1;
                mbar_wait_relaxed(&p_full[a], (uint32_t)((j >> 1) & 1));
                tc_fence_after();
                const uint64_t vdesc = desc_mnmajor(smem_u32(kv_s + st * 2 * TC_TILE_BYTES + TC_TILE_BYTES), TC_NT * 128, 1024);
                const uint32_t pa = tmem_base + (uint32_t)(a * 64);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_f16_ts(tmem_base + 128u, pa + (uint32_t)(k * 8), vdesc + (uint64_t)((k * 2048) >> 4), id_o,
                                (uint32_t)((j | k) != 0));
                umma_commit(pv_done);
                umma_commit(&kv_empty[st]);
            };
            issue_s(0);
            for (int j = 0; j < n_tiles; ++j) {
                if (j + 1 < n_tiles) issue_s(j + 1);
                issue_o(j);
            }
        }
    } else {
        const int quarter = warp & 3, half = warp >> 2;
        const int r = quarter * 32 + lane;
        const int tl = q0 + r / GROUP;
        const int head = kvh * GROUP + r % GROUP;
        const bool valid = active && tl < n_new;
        const int lim = valid ? ctx + tl + 1 : kv_end;
        if (active) {
            const uint4* src = reinterpret_cast<const uint4*>(q + (size_t)(tok_begin + min(tl, n_new - 1)) * q_stride +
                                                              (size_t)head * TC_HD) + half * 8;
            uint8_t* dst = q_s + half * (TC_ROWS * 128) + r * 128;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const uint4 v = valid ? src[c] : make_uint4(0u, 0u, 0u, 0u);
                *reinterpret_cast<uint4*>(dst + ((c ^ (r & 7)) << 4)) = v;
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(q_ready);
        }
        float m = -INFINITY, l = 0.f;
        const float sc = TC_LOG2E * rsqrtf((float)TC_HD);
        const uint32_t t_x = tmem_base + ((uint32_t)(quarter * 32) << 16);
        const uint32_t o_mine = t_x + 128u + (uint32_t)(half * 64);

        for (int j = 0; j < n_tiles; ++j) {
            const int a = j & 1;
            mbar_wait(&s_full[a], (uint32_t)((j >> 1) & 1));
            tc_fence_after();
            uint32_t u[32];
            float s[32];
            tmem_ld_32x32(t_x + (uint32_t)(a * 64 + half * 32), u);
            tmem_ld_wait();
            const int tok0 = j * TC_NT + half * 32;
            float mx = -INFINITY;
            if (tok0 + 32 > lim) {
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float v = (tok0 + i >= lim) ? -INFINITY : __uint_as_float(u[i]);
                    s[i] = v;
                    mx = fmaxf(mx, v);
                }
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    s[i] = __uint_as_float(u[i]);
                    mx = fmaxf(mx, s[i]);
                }
            }
            float* xb = mx_x + a * (2 * TC_ROWS);
            xb[half * TC_ROWS + r] = mx;
            asm volatile("bar.sync 1, %0;" ::"n"(TC_SOFTMAX_THREADS) : "memory");   // both halves hold S_j in registers
            const float tile_mx = fmaxf(mx, xb[(half ^ 1) * TC_ROWS + r]) * sc;
            const float m_new = (tile_mx > m + 8.f) ? tile_mx : m;
            const float corr = ex2_approx(m - m_new);
            m = m_new;
            float sum = 0.f;
            uint32_t pw[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float p0 = ex2_approx(fmaf(s[2 * i], sc, -m_new));
                const float p1 = ex2_approx(fmaf(s[2 * i + 1], sc, -m_new));
                sum += p0 + p1;
                pw[i] = pack_f16(p0, p1);
            }
            l = l * corr + sum;
            tmem_st_32x16(t_x + (uint32_t)(a * 64 + half * 16), pw);       // P_j over the first 32 columns of S[a]
            if (j > 0 && __any_sync(0xffffffffu, corr != 1.f)) {
                mbar_wait(pv_done, (uint32_t)((j - 1) & 1));
                tc_fence_after();
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    tmem_ld_32x32(o_mine + (uint32_t)(c * 32), u);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) u[i] = __float_as_uint(__uint_as_float(u[i]) * corr);
                    tmem_st_32x16(o_mine + (uint32_t)(c * 32), reinterpret_cast<const uint32_t(&)[16]>(u[0]));
                    tmem_st_32x16(o_mine + (uint32_t)(c * 32 + 16), reinterpret_cast<const uint32_t(&)[16]>(u[16]));
                }
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&p_full[a]);
        }
        if (n_tiles > 0) {
            mbar_wait(pv_done, (uint32_t)((n_tiles - 1) & 1));
            tc_fence_after();
            float* xb = mx_x + (n_tiles & 1) * (2 * TC_ROWS);
            xb[half * TC_ROWS + r] = l;
            asm volatile("bar.sync 1, %0;" ::"n"(TC_SOFTMAX_THREADS) : "memory");
            l += xb[(half ^ 1) * TC_ROWS + r];
            const float inv = 1.f / l;
            uint16_t* op = out + ((size_t)(tok_begin + tl) * n_heads + head) * TC_HD + half * 64;
            uint32_t u[32];
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                tmem_ld_32x32(o_mine + (uint32_t)(c * 32), u);
                tmem_ld_wait();
                if (valid) {
#pragma unroll
                    for (int q4 = 0; q4 < 4; ++q4) {
                        uint4 w;
                        w.x = pack_act(__uint_as_float(u[8 * q4 + 0]) * inv, __uint_as_float(u[8 * q4 + 1]) * inv, out_f16);
                        w.y = pack_act(__uint_as_float(u[8 * q4 + 2]) * inv, __uint_as_float(u[8 * q4 + 3]) * inv, out_f16);
                        w.z = pack_act(__uint_as_float(u[8 * q4 + 4]) * inv, __uint_as_float(u[8 * q4 + 5]) * inv, out_f16);
                        w.w = pack_act(__uint_as_float(u[8 * q4 + 6]) * inv, __uint_as_float(u[8 * q4 + 7]) * inv, out_f16);
                        *reinterpret_cast<uint4*>(op + 32 * c + 8 * q4) = w;
                    }
                }
            }
            tc_fence_before();
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 9) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 256);
    }
}

template <int GROUP>
int launch_tcs(const void* q, int q_stride, const void* kp, const void* vp, const int* cu, const int* ctx,
               const int* bt, void* out, int n_seqs, int max_new, int max_blocks, int n_heads, int n_kv_heads,
               int out_f16, cudaStream_t stream) {
    auto kern = attn_prefill_tcs_kernel<GROUP>;
    static bool configured[64] = {};
    int dev = 0;
    KB_CUDA(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        KB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TCS_SMEM));
        configured[dev & 63] = true;
    }
    constexpr int TQ = TC_ROWS / GROUP;
    dim3 grid((max_new + TQ - 1) / TQ, n_kv_heads, n_seqs);
    KB_CUDA(launch_kernel(kern, grid, dim3(TC_THREADS), TCS_SMEM, stream, (const __half*)q, q_stride,
                          (const __half*)kp, (const __half*)vp, cu, ctx, bt, (uint16_t*)out, max_blocks,
                          n_heads, n_kv_heads, out_f16));
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

template <int GROUP>
int launch_tc2(const void* q, int q_stride, const void* kp, const void* vp, const int* cu, const int* ctx,
               const int* bt, void* out, int n_seqs, int max_new, int max_blocks, int n_heads, int n_kv_heads,
               int out_f16, cudaStream_t stream) {
    auto kern = attn_prefill_tc2_kernel<GROUP>;
    static bool configured[64] = {};
    int dev = 0;
    KB_CUDA(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        KB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TC2_SMEM));
        configured[dev & 63] = true;
    }
    constexpr int TQ2 = 2 * TC_ROWS / GROUP;               // query tokens per CTA (two tiles)
    dim3 grid((max_new + TQ2 - 1) / TQ2, n_kv_heads, n_seqs);
    KB_CUDA(launch_kernel(kern, grid, dim3(TC2_THREADS), TC2_SMEM, stream, (const __half*)q, q_stride,
                          (const __half*)kp, (const __half*)vp, cu, ctx, bt, (uint16_t*)out, max_blocks,
                          n_heads, n_kv_heads, out_f16));
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

template <int GROUP>
int launch_tc(const void* q, int q_stride, const void* kp, const void* vp, const int* cu, const int* ctx,
              const int* bt, void* out, int n_seqs, int max_new, int max_blocks, int n_heads, int n_kv_heads,
              int out_f16, cudaStream_t stream) {
    auto kern = attn_prefill_tc_kernel<GROUP>;
    static bool configured[64] = {};
    int dev = 0;
    KB_CUDA(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        KB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM));
        configured[dev & 63] = true;
    }
    constexpr int TQ = TC_ROWS / GROUP;
    dim3 grid((max_new + TQ - 1) / TQ, n_kv_heads, n_seqs);
    KB_CUDA(launch_kernel(kern, grid, dim3(TC_THREADS), TC_SMEM, stream, (const __half*)q, q_stride,
                          (const __half*)kp, (const __half*)vp, cu, ctx, bt, (uint16_t*)out, max_blocks,
                          n_heads, n_kv_heads, out_f16));
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

}  // namespace

// Returns -1 when the shape is not covered (caller falls back to the mma.sync kernel).
int attn_prefill_tc(const void* q, int q_stride, const void* kp, const void* vp, const int* cu, const int* ctx,
                    const int* bt, void* out, int n_seqs, int max_new, int max_blocks, int n_heads, int n_kv_heads,
                    int head_dim, int out_f16, cudaStream_t stream) {
    if (head_dim != TC_HD) return -1;
    const int group = n_heads / n_kv_heads;
    // KB_ATTN_TC2=1: two query tiles per CTA (attn_prefill_tc2_kernel) whenever some sequence can fill more than one
    // (max_new * group > 128 rows). Measured on the turn prefill (512 x 64 new tokens on 2 048 cached,
    // profiles/r02_prefill_attention.md) it is NOT faster than one tile per CTA — 83.7 vs 75.6 ms: the pair of tiles
    // advances at the rate at which the one MMA-issuing thread gets its 24 small MMAs per K/V tile accepted, and with
    // half as many CTAs the last wave is emptier — so it stays an option (read per call: tests cover both kernels).
    const char* old64 = getenv("KB_ATTN_TC64");       // read per call: "1" = one of the 64-key-tile kernels below
    const char* halves = getenv("KB_ATTN_T128H");     // read per call: "1" = 128-key tiles with independent key halves
    if ((old64 == nullptr || old64[0] != '1') && halves != nullptr && halves[0] == '1') {
#define KB_T128H_CASE(G_)                                                                                       \
    if (group == G_)                                                                                            \
        return launch_tc128h<G_>(q, q_stride, kp, vp, cu, ctx, bt, out, n_seqs, max_new, max_blocks, n_heads,  \
                                 n_kv_heads, out_f16, stream);
        KB_T128H_CASE(1) KB_T128H_CASE(2) KB_T128H_CASE(4) KB_T128H_CASE(8)
#undef KB_T128H_CASE
    }
    if (old64 == nullptr || old64[0] != '1') {
#define KB_T128_CASE(G_)                                                                                        \
    if (group == G_)                                                                                            \
        return launch_tc128<G_>(q, q_stride, kp, vp, cu, ctx, bt, out, n_seqs, max_new, max_blocks, n_heads,   \
                                n_kv_heads, out_f16, stream);
        KB_T128_CASE(1) KB_T128_CASE(2) KB_T128_CASE(4) KB_T128_CASE(8)
#undef KB_T128_CASE
    }
    const char* small = getenv("KB_ATTN_TCS");        // read per call: "1" = two small-footprint CTAs per SM
    if (small != nullptr && small[0] == '1') {
#define KB_TCS_CASE(G_)                                                                                         \
    if (group == G_)                                                                                            \
        return launch_tcs<G_>(q, q_stride, kp, vp, cu, ctx, bt, out, n_seqs, max_new, max_blocks, n_heads,     \
                              n_kv_heads, out_f16, stream);
        KB_TCS_CASE(1) KB_TCS_CASE(2) KB_TCS_CASE(4) KB_TCS_CASE(8)
#undef KB_TCS_CASE
    }
    const char* two = getenv("KB_ATTN_TC2");
    if (two != nullptr && two[0] == '1' && max_new * group > TC_ROWS) {
#define KB_TC2_CASE(G_)                                                                                         \
    if (group == G_)                                                                                            \
        return launch_tc2<G_>(q, q_stride, kp, vp, cu, ctx, bt, out, n_seqs, max_new, max_blocks, n_heads,     \
                              n_kv_heads, out_f16, stream);
        KB_TC2_CASE(1) KB_TC2_CASE(2) KB_TC2_CASE(4) KB_TC2_CASE(8)
#undef KB_TC2_CASE
    }
#define KB_TC_CASE(G_)                                                                                          \
    if (group == G_)                                                                                            \
        return launch_tc<G_>(q, q_stride, kp, vp, cu, ctx, bt, out, n_seqs, max_new, max_blocks, n_heads,      \
                             n_kv_heads, out_f16, stream);
    KB_TC_CASE(1) KB_TC_CASE(2) KB_TC_CASE(4) KB_TC_CASE(8)
#undef KB_TC_CASE
    return -1;
}

}  // namespace kb

#ifdef KB_ATTN_TRACE
extern "C" int kb_debug_attn_trace(unsigned long long* host_out512) {
    return cudaMemcpyFromSymbol(host_out512, kb::g_attn_trace, sizeof(unsigned long long) * 512) == cudaSuccess ? 0 : 1;
}
#endif
