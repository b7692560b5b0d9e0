// kb_gemm.cu — D = A[M,K] * W[N,K]^T on the 5th-gen tensor cores (tcgen05), sm_100a.
//
// Takes over the projections (QKV / O / gate-up / down / lm_head) of the completion the reference
// obtains remotely (src/llm/portkey.py:379). There is no reference kernel to translate; the design
// is the Blackwell-native one:
//
//   * persistent CTAs (one per SM), static tile schedule, M-fastest so the CTAs that run
//     concurrently share one weight tile through L2;
//   * warp 0   : TMA producer  (cp.async.bulk.tensor.2d, SWIZZLE_128B boxes, mbarrier tx-count)
//     warp 1   : MMA issuer    (one elected thread, tcgen05.mma.cta_group::1.kind::f16, 128 x BN x 16)
//     warp 2   : TMEM allocator (2 accumulators of BN fp32 columns -> epilogue overlaps next tile)
//     warps 4-7: epilogue      (tcgen05.ld 32x32b.x32 -> registers -> fused epilogue -> global)
//   * three mbarrier pipelines: smem full/empty (TMA <-> MMA), TMEM full/empty (MMA <-> epilogue).
//
// Three kernels share that anatomy (gemm_ex picks per shape):
//   gemm_tcgen05_kernel<BN,EPI>      one CTA per 128 x BN tile (BN = 64 / 128 / 256)
//   gemm_tcgen05_2cta_kernel<EPI>    a CTA pair (cta_group::2) per 256 x bn tile, bn chosen at run time
//   gemm_tcgen05_skinny_kernel<EPI>  M <= 32 tokens: operands swapped, cluster split-K reduced through DSMEM
// and every kernel starts with its private set-up, then griddepcontrol.wait / launch_dependents (PDL).
//
// Roofline: tensor pipe for M >= ~256 rows, weight bandwidth below that.
// Algorithmic FLOPs per launch = 2*M*N*K.
#include "kb_common.cuh"

#include <atomic>
#include <mutex>
#include <unordered_map>
#include <string>
#include <cudaTypedefs.h>
#include <stdlib.h>

namespace kb {

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;  // 64 bf16 = 128 bytes = one SWIZZLE_128B row
constexpr int GEMM_THREADS = 256;

template <int BN>
struct GemmCfg {
    static constexpr int STAGES = (BN == 256) ? 4 : (BN == 128 ? 6 : 8);
    static constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
    static constexpr int B_BYTES = BN * GEMM_BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int TMEM_COLS = (2 * BN < 32) ? 32 : 2 * BN;  // power of two: 128/256/512
    static constexpr int BAR_BYTES = 256;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + BAR_BYTES + 1024;  // +1024: manual alignment
};

// Shared-memory matrix descriptor for a K-major tile stored as rows of 128 bytes with the
// 128-byte swizzle TMA applied (8-row atoms of 1024 bytes).
__device__ __forceinline__ uint64_t make_sw128_kmajor_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);  // start address, 16-byte units
    d |= (uint64_t)1 << 16;                    // leading byte offset: unused for swizzled K-major
    d |= (uint64_t)(1024 >> 4) << 32;          // stride byte offset: next 8-row atom
    d |= (uint64_t)1 << 46;                    // descriptor version 1 (Blackwell)
    d |= (uint64_t)2 << 61;                    // SWIZZLE_128B
    return d;
}

// kind::f16 instruction descriptor: A and B both bf16 or both fp16 (K-major), D = fp32, M = 128, N = BN.
// The descriptor has separate A/B format fields, but sm_100a traps (illegal instruction) when they
// differ — measured, round 1 — so the hot path, which wants fp16 activations for their 10 mantissa bits
// (error budget), also stores its projection weights as fp16 re-encodings of the bf16
// values (exact for |w| >= 2^-14).
__host__ __device__ constexpr uint32_t make_idesc(int m, int n, uint32_t is_bf16) {
    return (1u << 4)                       // D format  fp32
           | (is_bf16 << 7)                // A format  0 = fp16, 1 = bf16
           | (is_bf16 << 10)               // B format  (must equal A's)
           | ((uint32_t)(n >> 3) << 17)    // N / 8
           | ((uint32_t)(m >> 4) << 24);   // M / 16
}

// silu(g) = g / (1 + e^-g) with the approximate reciprocal (MUFU.RCP, 2 ulp; the result is rounded to 16 bits right
// after): an IEEE division costs ~3x the instructions, and the epilogue warps run one per scheduler with nothing
// to hide their latency. Denominators beyond 2^126 (g < -87) give 0 where the exact value is ~ -1e-36.
__device__ __forceinline__ float silu_f(float g) { return __fdividef(g, 1.0f + __expf(-g)); }

template <int EPI>
__device__ __forceinline__ void epilogue_store32(const uint32_t (&r)[32], void* out, const float* resid, int row,
                                                 int col0, int N, int ncols) {
    // 32 consecutive fp32 accumulator columns [col0, col0+32) of one row, of which the first `ncols`
    // (a multiple of 8; of 16 for SwiGLU) belong to this tile and lie inside N.
    if (EPI == KB_EPI_BF16 || EPI == KB_EPI_F16) {
        uint16_t* o = reinterpret_cast<uint16_t*>(out) + (size_t)row * N + col0;
        constexpr int f16 = (EPI == KB_EPI_F16);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (j * 8 < ncols) {
                uint4 v;
                v.x = pack_act(__uint_as_float(r[j * 8 + 0]), __uint_as_float(r[j * 8 + 1]), f16);
                v.y = pack_act(__uint_as_float(r[j * 8 + 2]), __uint_as_float(r[j * 8 + 3]), f16);
                v.z = pack_act(__uint_as_float(r[j * 8 + 4]), __uint_as_float(r[j * 8 + 5]), f16);
                v.w = pack_act(__uint_as_float(r[j * 8 + 6]), __uint_as_float(r[j * 8 + 7]), f16);
                *reinterpret_cast<uint4*>(o + j * 8) = v;
            }
        }
    } else if (EPI == KB_EPI_F32 || EPI == KB_EPI_RESID_F32) {
        float* o = reinterpret_cast<float*>(out) + (size_t)row * N + col0;
        const float* rs = resid + (size_t)row * N + col0;
        if (EPI == KB_EPI_RESID_F32 && rs == o) {
            // in place (x += A W^T, the decode step's O and down projections): let L2 do the add —
            // fire-and-forget vector reductions instead of a load / add / store round trip per chunk,
            // which left the epilogue waiting on L2 latency (9 us of a 27 us kernel, tools/gemm_trace.py).
            // Exactly one thread adds to each element, so the result is the same fp32 sum.
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (j * 4 < ncols)
                    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(o + j * 4),
                                 "f"(__uint_as_float(r[j * 4 + 0])), "f"(__uint_as_float(r[j * 4 + 1])),
                                 "f"(__uint_as_float(r[j * 4 + 2])), "f"(__uint_as_float(r[j * 4 + 3]))
                                 : "memory");
            }
            return;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (j * 4 < ncols) {
                float4 v = make_float4(__uint_as_float(r[j * 4 + 0]), __uint_as_float(r[j * 4 + 1]),
                                       __uint_as_float(r[j * 4 + 2]), __uint_as_float(r[j * 4 + 3]));
                if (EPI == KB_EPI_RESID_F32) {
                    float4 q = *reinterpret_cast<const float4*>(rs + j * 4);
                    v.x += q.x; v.y += q.y; v.z += q.z; v.w += q.w;
                }
                *reinterpret_cast<float4*>(o + j * 4) = v;
            }
        }
    } else {  // KB_EPI_SWIGLU_{BF16,F16}: columns (2j, 2j+1) = (gate_j, up_j) -> out[:, j]
        const int No = N >> 1;
        uint16_t* o = reinterpret_cast<uint16_t*>(out) + (size_t)row * No + (col0 >> 1);
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            if (j * 16 < ncols) {
                float h[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    float g = __uint_as_float(r[j * 16 + 2 * i]);
                    float u = __uint_as_float(r[j * 16 + 2 * i + 1]);
                    h[i] = silu_f(g) * u;
                }
                uint4 v;
                if (EPI == KB_EPI_SWIGLU_F16) {
                    v.x = pack_f16(h[0], h[1]);
                    v.y = pack_f16(h[2], h[3]);
                    v.z = pack_f16(h[4], h[5]);
                    v.w = pack_f16(h[6], h[7]);
                } else {
                    v.x = pack_bf16(h[0], h[1]);
                    v.y = pack_bf16(h[2], h[3]);
                    v.z = pack_bf16(h[4], h[5]);
                    v.w = pack_bf16(h[6], h[7]);
                }
                *reinterpret_cast<uint4*>(o + j * 8) = v;
            }
        }
    }
}

template <int BN, int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, void* out,
                    const float* resid, int M, int N, int K, uint32_t a_is_bf16) {
    using Cfg = GemmCfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::STAGES * Cfg::STAGE_BYTES);
    uint64_t* empty_bar = full_bar + Cfg::STAGES;
    uint64_t* tmem_full_bar = empty_bar + Cfg::STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int num_m = (M + GEMM_BM - 1) / GEMM_BM;
    const int num_n = (N + BN - 1) / BN;
    const int num_tiles = num_m * num_n;
    const int num_k = K / GEMM_BK;

    if (warp == 0) {
        if (elect_one()) {
            tma_prefetch_desc(&tmA);
            tma_prefetch_desc(&tmB);
        }
    } else if (warp == 1) {
        if (elect_one()) {
            for (int s = 0; s < Cfg::STAGES; ++s) {
                mbar_init(&full_bar[s], 1);
                mbar_init(&empty_bar[s], 1);
            }
            for (int a = 0; a < 2; ++a) {
                mbar_init(&tmem_full_bar[a], 1);
                mbar_init(&tmem_empty_bar[a], 4);  // one arrive per epilogue warp
            }
            fence_mbar_init();
        }
    } else if (warp == 2) {
        tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();      // set-up above overlapped the previous kernel; its outputs are visible from here on
    pdl_launch();

    if (warp == 0) {
        // ===== TMA producer =====
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                const int m_blk = tile % num_m;
                const int n_blk = tile / num_m;
                for (int kb = 0; kb < num_k; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * Cfg::STAGE_BYTES;
                    uint8_t* sb = sa + Cfg::A_BYTES;
                    mbar_arrive_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
                    tma_load_2d(sa, &tmA, kb * GEMM_BK, m_blk * GEMM_BM, &full_bar[stage]);
                    tma_load_2d(sb, &tmB, kb * GEMM_BK, n_blk * BN, &full_bar[stage]);
                    if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (elect_one()) {
            const uint32_t idesc = make_idesc(GEMM_BM, BN, a_is_bf16);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1);  // epilogue drained this accumulator
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
                for (int kb = 0; kb < num_k; ++kb) {
                    mbar_wait(&full_bar[stage], phase);  // TMA bytes landed
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * Cfg::STAGE_BYTES);
                    const uint64_t adesc = make_sw128_kmajor_desc(sa);
                    const uint64_t bdesc = make_sw128_kmajor_desc(sa + Cfg::A_BYTES);
#pragma unroll
                    for (int k = 0; k < GEMM_BK / 16; ++k) {
                        // advance 16 bf16 = 32 bytes along K inside the swizzle atom: +2 in 16-byte units
                        umma_bf16(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc,
                                  (uint32_t)((kb | k) != 0));
                    }
                    umma_commit(&empty_bar[stage]);  // smem slot free once these MMAs retire
                    if (kb == num_k - 1) umma_commit(&tmem_full_bar[acc]);
                    if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
                }
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
            }
        }
    } else if (warp >= 4) {
        // ===== epilogue: TMEM -> registers -> global =====
        const int ew = warp - 4;  // == warp % 4: the TMEM lane quarter this warp may read
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            const int m_blk = tile % num_m;
            const int n_blk = tile / num_m;
            mbar_wait(&tmem_full_bar[acc], acc_phase);
            tc_fence_after();
            const int row = m_blk * GEMM_BM + ew * 32 + lane;
            const uint32_t t_row = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(acc * BN);
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
                uint32_t r[32];
                tmem_ld_32x32(t_row + (uint32_t)c, r);
                tmem_ld_wait();
                const int col0 = n_blk * BN + c;
                if (row < M && col0 < N) epilogue_store32<EPI>(r, out, resid, row, col0, N, N - col0);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
    }
}

// ---------------------------------------------------------------------------------------------
// Phase timestamps of cluster 0 (profiling aid; build with `make EXTRA=-DKB_GEMM_TRACE`, read with
// tools/gemm_trace.py through kb_debug_gemm_trace — not part of the ABI, absent from normal builds).
#ifdef KB_GEMM_TRACE
__device__ unsigned long long g_gemm_trace[16];
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define KB_TRACE(slot) do { if (blockIdx.x == 0) g_gemm_trace[slot] = globaltimer_ns(); } while (0)
#else
#define KB_TRACE(slot) do { } while (0)
#endif

// CTA-pair variant (cta_group::2): a cluster of two CTAs owns a 256 x BN output tile.
//
// Why: a 1-CTA 128 x BN tile has to bring 16 KB of A plus BN*128 B of B into shared memory per k-block,
// and an SM takes in only ~97 GB/s by TMA (~50 B/clk at boost clock; profiles/r02_gemm_operand_staging.md) —
// that ingress, not the MMA, bounds the main loop. In a pair each CTA stages its own 128 rows of A but
// only HALF of B (BN/2 rows); the tensor cores of both SMs read both halves, so the ingress per CTA drops
// to 16 KB + BN*64 B for twice the rows: close to balanced with the MMA at BN = 256.
//
//   * both CTAs : TMA producer (own A rows, own half of B), completion bytes counted on the LEADER's
//                 full barrier; epilogue of their own 128 accumulator rows (own TMEM).
//   * leader    : the one MMA-issuing thread (tcgen05.mma.cta_group::2, M = 256); its commits arrive on
//                 the slot-free and accumulator-full barriers of BOTH CTAs (multicast).
//   * accumulator-free: the epilogue warps of both CTAs arrive on the leader's barrier (8 arrivals).
// ---------------------------------------------------------------------------------------------
// The tile width `bn` is a RUN-TIME argument (a multiple of 16, 32..256; the B tensor map's box is bn/2
// rows): the host picks it per shape so that the number of tiles fills whole rounds of the 74 CTA pairs —
// with fixed 256-wide tiles the decode step's gate/up projection (224 tiles) spent a quarter of its time
// in a fourth round that held 2 tiles (tools/gemm_trace.py).
// Split-K for the in-place residual epilogue (x += A W^T: the O and down projections, N = hidden).
//
// At M = 512 those two launches have only 2 x N/256 = 32 tiles of 256 x 256 for 74 CTA pairs; filling the GPU with
// narrower tiles (bn = 112) makes every tile pay 16 KB of A per 7 KB of W, and the main loop runs at the
// TMA intake rate (~97 GB/s per SM), not at the tensor rate. Instead the K range of a wide tile is
// cut into `splits` work units that run on different CTA pairs at the same time. Their partial sums meet in
// the output itself: every unit adds its accumulator to x with red.global.add — but in a FIXED order, unit s
// after unit s - 1, so that the fp32 result does not depend on which unit finished first (the engine's outputs
// are bit-reproducible, tests/test_gpu_full_size.py). The order is kept per (tile, CTA of the pair, epilogue
// warp): warp w of unit s touches exactly the elements warp w of unit s - 1 touched, so one producer warp
// hands over to one consumer warp through one 64-bit word: the producer publishes `tag + s + 1` after its
// additions (atomicMax: words only ever grow, so launches on other streams that share the array can disturb the
// order of two additions but can never make a waiter miss its value), the consumer spins until it sees it.
// `tag` is unique per launch (host counter), nothing is ever reset. Units are scheduled in index order
// (unit = tile * splits + s), every dependence points to a lower index and all clusters of the persistent grid
// are resident together, so the wait cannot deadlock.
constexpr int SPLITK_MAX_TILES = 1024;
static __device__ unsigned long long g_splitk_flags[SPLITK_MAX_TILES * 8];   // [tile][cta of the pair][epilogue warp]

__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// Tiles up to 256 columns wide: ring of 6 x 32 KB stages, two accumulators of 256 tensor-memory columns (the
// epilogue of tile i runs under the main loop of tile i + 1).
// WIDE tiles, 256 < bn <= 512: the accumulator takes all 512 columns and every k-step issues TWO MMAs that share
// the A operand — N = 256 into columns [0, 256) and N = bn - 256 into [256, bn). What it buys at large M: fewer,
// fuller rounds of the 74 CTA pairs (N = 4 096 at M = 4 096: 144 units = 2 rounds instead of 256 = 4, the last one
// 46 % full) and a third less L2 / DRAM traffic. It does not make an MMA faster: a 128 x 256 x 16 MMA takes ~95 ns
// per SM whatever the SM clock reads (1.64 PFLOP/s over the chip, cuBLAS's burst figure), an SM takes in ~97 GB/s
// through TMA, and t(k-block) = max of the two explains the narrow, the 256-wide and the wide tiles alike
// (profiles/r02_gemm_operand_staging.md). The price: one accumulator, so the epilogue no longer hides under the
// next tile's main loop, and a ring of 4 - 5 deeper stages (40 - 48 KB each).
struct Gemm2Cfg {
    static constexpr int BN_MAX = 256;                          // widest single MMA
    static constexpr int BN_WIDE_MAX = 512;
    static constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
    static constexpr int B_BYTES_MAX = (BN_MAX / 2) * GEMM_BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES_MAX;   // per CTA, tiles up to 256 wide
    static constexpr int STAGES = 6;                            // 192 KB
    static constexpr int MAX_STAGES = 6;
    static constexpr int TMEM_COLS = 2 * BN_MAX;                // two accumulators, or one wide one
    static constexpr int BAR_BYTES = 256;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + BAR_BYTES + 1024;
    static constexpr int WIDE_RING_BYTES = 216 * 1024;          // wide tiles: as many stages as fit in this
    static constexpr int SMEM_BYTES_WIDE = WIDE_RING_BYTES + BAR_BYTES + 1024;
    // bytes of one stage of a wide tile: A + this CTA's bn / 2 rows of B, kept 1024-aligned for SWIZZLE_128B
    __host__ __device__ static constexpr int wide_stage_bytes(int bn) { return A_BYTES + (((bn / 2) * GEMM_BK * 2 + 1023) & ~1023); }
    __host__ __device__ static constexpr int wide_stages(int bn) {
        return WIDE_RING_BYTES / wide_stage_bytes(bn) < MAX_STAGES ? WIDE_RING_BYTES / wide_stage_bytes(bn) : MAX_STAGES;
    }
};

template <int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_2cta_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const __grid_constant__ CUtensorMap tmB2, void* out, const float* resid, int M, int N, int K,
                         int bn, uint32_t is_bf16, int splits, unsigned long long tag) {
    using Cfg = Gemm2Cfg;
    const int BN = bn;
    const int half_n = bn >> 1;                       // B rows staged by one CTA
    const bool wide = bn > Cfg::BN_MAX;               // one 512-column accumulator, two MMAs per k-step
    const int n2 = wide ? bn - Cfg::BN_MAX : 0;       // width of the second MMA
    const int stage_bytes = wide ? Cfg::wide_stage_bytes(bn) : Cfg::STAGE_BYTES;
    const int n_stages = wide ? Cfg::wide_stages(bn) : Cfg::STAGES;
    const uint32_t stage_tx = 2u * (uint32_t)(Cfg::A_BYTES + half_n * GEMM_BK * 2);
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + n_stages * stage_bytes);
    uint64_t* empty_bar = full_bar + Cfg::MAX_STAGES;
    uint64_t* tmem_full_bar = empty_bar + Cfg::MAX_STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    if (threadIdx.x == 0) KB_TRACE(0);               // kernel entry
    const uint32_t cta = cluster_ctarank();          // 0 = leader
    const int cluster_id = blockIdx.x >> 1;
    const int num_clusters = gridDim.x >> 1;
    const int num_mp = (M + 2 * GEMM_BM - 1) / (2 * GEMM_BM);
    const int num_n = (N + BN - 1) / BN;
    const int num_tiles = num_mp * num_n;
    const int num_k = K / GEMM_BK;
    const int num_units = num_tiles * splits;         // unit = tile * splits + (which part of K)
    const int kps = num_k / splits;                   // k-blocks per unit (host: num_k % splits == 0)

    if (warp == 0) {
        if (elect_one()) {
            tma_prefetch_desc(&tmA);
            tma_prefetch_desc(&tmB);
            if (wide) tma_prefetch_desc(&tmB2);
        }
    } else if (warp == 1) {
        if (elect_one()) {
            for (int s = 0; s < n_stages; ++s) {
                mbar_init(&full_bar[s], 1);      // leader's producer arms it; both CTAs' TMA bytes land on it
                mbar_init(&empty_bar[s], 1);     // one multicast commit per use
            }
            for (int a = 0; a < 2; ++a) {
                mbar_init(&tmem_full_bar[a], 1);
                mbar_init(&tmem_empty_bar[a], 8);    // 4 epilogue warps x 2 CTAs (only the leader's is waited on)
            }
            fence_mbar_init();
        }
    } else if (warp == 2) {
        tmem_alloc_2cta(tmem_slot, Cfg::TMEM_COLS);
        tmem_relinquish_2cta();
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                              // peer barriers are initialised before anyone signals them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();      // set-up above overlapped the previous kernel; its outputs are visible from here on
    pdl_launch();
    if (threadIdx.x == 0) KB_TRACE(1);               // barriers, TMEM, cluster sync done

    if (warp == 0) {
        // ===== TMA producer (both CTAs) =====
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int unit = cluster_id; unit < num_units; unit += num_clusters) {
                const int tile = unit / splits;
                const int kb_lo = (unit - tile * splits) * kps;
                const int mp = tile % num_mp;
                const int n_blk = tile / num_mp;
                const int row0 = mp * 2 * GEMM_BM + (int)cta * GEMM_BM;        // this CTA's 128 rows of A
                // this CTA's half of B. Wide tile: its half of the first 256 columns (128 rows) and, behind them,
                // its half of the remaining n2 — accumulator column j is output column n_blk * BN + j either way
                const int nrow0 = n_blk * BN + (int)cta * (wide ? Cfg::BN_MAX / 2 : half_n);
                const int nrow1 = n_blk * BN + Cfg::BN_MAX + (int)cta * (n2 >> 1);
                for (int kb = kb_lo; kb < kb_lo + kps; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * stage_bytes;
                    uint8_t* sb = sa + Cfg::A_BYTES;
                    if (cta == 0) mbar_arrive_expect_tx(&full_bar[stage], stage_tx);
                    tma_load_2d_2cta(sa, &tmA, kb * GEMM_BK, row0, &full_bar[stage]);
                    tma_load_2d_2cta(sb, &tmB, kb * GEMM_BK, nrow0, &full_bar[stage]);
                    if (wide) tma_load_2d_2cta(sb + Cfg::B_BYTES_MAX, &tmB2, kb * GEMM_BK, nrow1, &full_bar[stage]);
                    if (++stage == n_stages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (leader CTA only) =====
        if (cta == 0 && elect_one()) {
            const uint32_t idesc = make_idesc(2 * GEMM_BM, wide ? Cfg::BN_MAX : BN, is_bf16);
            const uint32_t idesc2 = make_idesc(2 * GEMM_BM, wide ? n2 : BN, is_bf16);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int unit = cluster_id; unit < num_units; unit += num_clusters) {
                mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * Cfg::BN_MAX);
                for (int kb = 0; kb < kps; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    if (kb == 0) KB_TRACE(2);        // first operands landed
                    if (kb == kps - 1) KB_TRACE(3);  // last operands landed
                    const uint32_t sa = smem_u32(smem + stage * stage_bytes);
                    const uint64_t adesc = make_sw128_kmajor_desc(sa);
                    const uint64_t bdesc = make_sw128_kmajor_desc(sa + Cfg::A_BYTES);
                    if (!wide) {
#pragma unroll
                        for (int k = 0; k < GEMM_BK / 16; ++k) {
                            umma_f16_2cta(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc,
                                          (uint32_t)((kb | k) != 0));
                        }
                    } else {
                        const uint64_t bdesc2 = make_sw128_kmajor_desc(sa + Cfg::A_BYTES + Cfg::B_BYTES_MAX);
#pragma unroll
                        for (int k = 0; k < GEMM_BK / 16; ++k)
                            umma_f16_2cta(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc,
                                          (uint32_t)((kb | k) != 0));
#pragma unroll
                        for (int k = 0; k < GEMM_BK / 16; ++k)
                            umma_f16_2cta(d_tmem + (uint32_t)Cfg::BN_MAX, adesc + (uint64_t)(k * 2),
                                          bdesc2 + (uint64_t)(k * 2), idesc2, (uint32_t)((kb | k) != 0));
                    }
                    umma_commit_2cta(&empty_bar[stage]);
                    if (kb == kps - 1) umma_commit_2cta(&tmem_full_bar[acc]);
                    if (++stage == n_stages) { stage = 0; phase ^= 1; }
                }
                if (wide) {                          // one accumulator: the same barrier pair every unit
                    acc_phase ^= 1;
                } else {
                    acc ^= 1;
                    if (acc == 0) acc_phase ^= 1;
                }
            }
        }
    } else if (warp >= 4) {
        // ===== epilogue (both CTAs): own 128 rows =====
        const int ew = warp - 4;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int unit = cluster_id; unit < num_units; unit += num_clusters) {
            const int tile = unit / splits;
            const int part = unit - tile * splits;
            const int mp = tile % num_mp;
            const int n_blk = tile / num_mp;
            mbar_wait(&tmem_full_bar[acc], acc_phase);
            tc_fence_after();
            if (threadIdx.x == 128) KB_TRACE(4);     // accumulator complete
            const int row = mp * 2 * GEMM_BM + (int)cta * GEMM_BM + ew * 32 + lane;
            const uint32_t t_row = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(acc * Cfg::BN_MAX);
            // Two register sets: the tcgen05.ld of chunk c + 1 is in flight while chunk c is converted and stored
            // (with one set the load of the next chunk queued behind the 8 vector reductions of the previous one).
            uint32_t ra[32], rb[32];
            tmem_ld_32x32(t_row, ra);
            unsigned long long* flag = nullptr;
            if (splits > 1) {
                // ordered split-K (see g_splitk_flags): wait until the unit that covers the K range before ours
                // has added its partial sums to these rows
                flag = &g_splitk_flags[((size_t)tile * 2 + cta) * 4 + ew];
                if (part > 0) {
                    if (lane == 0)
                        while (ld_acquire_u64(flag) < tag + (unsigned long long)part) { }
                    __syncwarp();
                }
            }
#pragma unroll 1
            for (int c = 0; c < BN; c += 64) {
                tmem_ld_wait();                                           // ra = chunk c
                if (c + 32 < BN) tmem_ld_32x32(t_row + (uint32_t)(c + 32), rb);
                {
                    const int col0 = n_blk * BN + c;
                    const int ncols = (N - col0 < BN - c) ? N - col0 : BN - c;
                    if (row < M && ncols > 0) epilogue_store32<EPI>(ra, out, resid, row, col0, N, ncols);
                }
                if (c + 32 < BN) {
                    tmem_ld_wait();                                       // rb = chunk c + 32
                    if (c + 64 < BN) tmem_ld_32x32(t_row + (uint32_t)(c + 64), ra);
                    const int col0 = n_blk * BN + c + 32;
                    const int ncols = (N - col0 < BN - c - 32) ? N - col0 : BN - c - 32;
                    if (row < M && ncols > 0) epilogue_store32<EPI>(rb, out, resid, row, col0, N, ncols);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(&tmem_empty_bar[acc], 0);   // the leader's barrier
            if (splits > 1 && part + 1 < splits) {
                __threadfence();                     // this lane's additions are performed before the hand-over
                __syncwarp();
                if (lane == 0) atomicMax(flag, tag + (unsigned long long)part + 1ull);
            }
            if (threadIdx.x == 128) KB_TRACE(5);     // this warp's rows stored
            if (wide) {
                acc_phase ^= 1;
            } else {
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                              // nobody leaves while the peer may still touch its smem / TMEM
    if (threadIdx.x == 0) KB_TRACE(6);
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc_2cta(tmem_base, Cfg::TMEM_COLS);
    }
    if (threadIdx.x == 64) KB_TRACE(7);              // TMEM released
}

// ---------------------------------------------------------------------------------------------
// Small-M variant ("skinny": M <= 128 tokens — decode at low concurrency, short prefill chunks).
//
// With few tokens the projections are pure weight streaming, and the 128 x BN kernel above wastes its
// shared-memory ingress on a 128-row activation tile that is mostly zero fill (16 KB per k-block against
// 8 KB of weights: ~2 TB/s of weights at best, tools/gemm_trace.py / bench --threads 1). Here the
// operands trade places:  D^T[128 weight rows, T tokens] = W_tile[128, K] * X[T, K]^T
//   * A operand = 128 weight rows (16 KB per k-block, all of it useful), B operand = the T-token
//     activation tile (T = 16 / 32 / 64 / 128 >= M; rows beyond M are TMA zero fill: 2..16 KB),
//     accumulator = T fp32 columns of TMEM;
//   * the epilogue stores transposed: lane = weight row, so for one token a warp writes 32 consecutive
//     outputs (coalesced); the SwiGLU pair (gate_j, up_j) sits in adjacent lanes;
//   * N / 128 tiles are too few CTAs to pull the weights at HBM speed when N = hidden, so a cluster of
//     `splits` CTAs shares one tile, each streaming 1/splits of K; the partial accumulators are added in
//     rank order by the leader through distributed shared memory — deterministic, no atomics, no scratch.
// ---------------------------------------------------------------------------------------------
struct SkinnyCfg {
    static constexpr int W_BYTES = GEMM_BM * GEMM_BK * 2;   // 16 KB
    static constexpr int MAX_STAGES = 10;
    static constexpr int BAR_BYTES = 512;
    static constexpr int SMEM_BUDGET = 200 * 1024;          // ring + reduction buffers
};

__device__ __forceinline__ void mbar_arrive_release_cluster(uint64_t* bar, uint32_t cta) {
    asm volatile(
        "{\n"
        ".reg .b32 ra;\n"
        "mapa.shared::cluster.u32 ra, %0, %1;\n"
        "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(cta)
        : "memory");
}
__device__ __forceinline__ void mbar_wait_acquire_cluster(uint64_t* bar, uint32_t parity) {
    const uint32_t a = smem_u32(bar);
    uint32_t done = 0;
    while (!done) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(a), "r"(parity)
            : "memory");
    }
}
__device__ __forceinline__ void st_cluster_f32x4(uint32_t local_saddr, uint32_t cta, float4 v) {
    asm volatile(
        "{\n"
        ".reg .b32 ra;\n"
        "mapa.shared::cluster.u32 ra, %0, %1;\n"
        "st.shared::cluster.v4.f32 [ra], {%2, %3, %4, %5};\n"
        "}\n" ::"r"(local_saddr),
        "r"(cta), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
        : "memory");
}

// r[j] = D[n (this lane's weight row)][token c + j]; stores the tokens < M of output column n.
template <int EPI>
__device__ __forceinline__ void skinny_store32(const float (&r)[32], void* out, const float* resid, int n, int c,
                                               int M, int N, int lane) {
    if (EPI == KB_EPI_RESID_F32) {
        float* o = reinterpret_cast<float*>(out);
        if (n >= N) return;
        if (resid == o) {
            // in place: fire-and-forget reductions in L2 (one adder per element: same fp32 sum), so the
            // loop does not wait for one L2 round trip per token
#pragma unroll
            for (int j = 0; j < 32; ++j)
                if (c + j < M)
                    asm volatile("red.global.add.f32 [%0], %1;" ::"l"(o + (size_t)(c + j) * N + n), "f"(r[j]) : "memory");
        } else {
            float q[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) q[j] = (c + j < M) ? resid[(size_t)(c + j) * N + n] : 0.f;   // independent loads
#pragma unroll
            for (int j = 0; j < 32; ++j)
                if (c + j < M) o[(size_t)(c + j) * N + n] = q[j] + r[j];
        }
        return;
    }
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const int token = c + j;
        if (token >= M) break;                      // warp-uniform
        const float v = r[j];
        if (EPI == KB_EPI_SWIGLU_BF16 || EPI == KB_EPI_SWIGLU_F16) {
            const float up = __shfl_down_sync(0xffffffffu, v, 1);
            if (!(lane & 1) && n < N) {
                const float h = silu_f(v) * up;
                uint16_t* o = reinterpret_cast<uint16_t*>(out) + (size_t)token * (N >> 1) + (n >> 1);
                *o = (EPI == KB_EPI_SWIGLU_F16) ? to_f16_sat(h) : __bfloat16_as_ushort(__float2bfloat16(h));
            }
        } else if (n < N) {
            const size_t idx = (size_t)token * N + n;
            if (EPI == KB_EPI_BF16) reinterpret_cast<__nv_bfloat16*>(out)[idx] = __float2bfloat16(v);
            else if (EPI == KB_EPI_F16) reinterpret_cast<uint16_t*>(out)[idx] = to_f16_sat(v);
            else reinterpret_cast<float*>(out)[idx] = v;   // KB_EPI_F32
        }
    }
}

template <int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_skinny_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX, void* out,
                           const float* resid, int M, int N, int K, int T, int stages, int splits, uint32_t is_bf16) {
    using Cfg = SkinnyCfg;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int stage_bytes = Cfg::W_BYTES + T * GEMM_BK * 2;
    const int red_bytes_per_peer = GEMM_BM * T * 4;
    float* red_buf = reinterpret_cast<float*>(smem + stages * stage_bytes);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + stages * stage_bytes + (splits - 1) * red_bytes_per_peer);
    uint64_t* empty_bar = full_bar + Cfg::MAX_STAGES;
    uint64_t* tmem_full_bar = empty_bar + Cfg::MAX_STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint64_t* red_bar = tmem_empty_bar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(red_bar + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = splits > 1 ? cluster_ctarank() : 0u;
    const int num_tiles = (N + GEMM_BM - 1) / GEMM_BM;
    const int num_k = K / GEMM_BK;
    const int kb_lo = (int)rank * (num_k / splits);
    const int kb_hi = kb_lo + num_k / splits;
    const int acc_stride = T < 32 ? 32 : T;                  // TMEM columns between the two accumulators
    const uint32_t tmem_cols = (uint32_t)(2 * acc_stride);   // 64 / 64 / 128 / 256
    // persistent over tiles without split-K; exactly one tile per cluster with it
    const int tile0 = splits > 1 ? (int)blockIdx.x / splits : (int)blockIdx.x;
    const int tile_step = splits > 1 ? num_tiles : (int)gridDim.x;

    if (warp == 0) {
        if (elect_one()) {
            tma_prefetch_desc(&tmW);
            tma_prefetch_desc(&tmX);
        }
    } else if (warp == 1) {
        if (elect_one()) {
            for (int s = 0; s < stages; ++s) {
                mbar_init(&full_bar[s], 1);
                mbar_init(&empty_bar[s], 1);
            }
            for (int a = 0; a < 2; ++a) {
                mbar_init(&tmem_full_bar[a], 1);
                mbar_init(&tmem_empty_bar[a], 4);
            }
            mbar_init(red_bar, (uint32_t)((splits - 1) * 4));   // one arrive per epilogue warp of every peer
            fence_mbar_init();
        }
    } else if (warp == 2) {
        tmem_alloc(tmem_slot, tmem_cols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    if (splits > 1) cluster_sync_all();          // the leader's reduction barrier exists before a peer arrives on it
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();      // set-up above overlapped the previous kernel; its outputs are visible from here on
    pdl_launch();

    if (warp == 0) {
        // ===== TMA producer =====
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = tile0; tile < num_tiles; tile += tile_step) {
                for (int kb = kb_lo; kb < kb_hi; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sw = smem + stage * stage_bytes;
                    mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)stage_bytes);
                    tma_load_2d(sw, &tmW, kb * GEMM_BK, tile * GEMM_BM, &full_bar[stage]);
                    tma_load_2d(sw + Cfg::W_BYTES, &tmX, kb * GEMM_BK, 0, &full_bar[stage]);
                    if (++stage == stages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: A = weights (M = 128), B = tokens (N = T) =====
        if (elect_one()) {
            const uint32_t idesc = make_idesc(GEMM_BM, T, is_bf16);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int tile = tile0; tile < num_tiles; tile += tile_step) {
                mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * acc_stride);
                for (int kb = kb_lo; kb < kb_hi; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t sw = smem_u32(smem + stage * stage_bytes);
                    const uint64_t adesc = make_sw128_kmajor_desc(sw);
                    const uint64_t bdesc = make_sw128_kmajor_desc(sw + Cfg::W_BYTES);
#pragma unroll
                    for (int k = 0; k < GEMM_BK / 16; ++k)
                        umma_bf16(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc,
                                  (uint32_t)((kb > kb_lo) | (k != 0)));
                    umma_commit(&empty_bar[stage]);
                    if (kb == kb_hi - 1) umma_commit(&tmem_full_bar[acc]);
                    if (++stage == stages) { stage = 0; phase ^= 1; }
                }
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
            }
        }
    } else if (warp >= 4) {
        // ===== epilogue: TMEM -> (cluster reduction) -> transposed store =====
        const int ew = warp - 4;
        const int row = ew * 32 + lane;              // weight row inside the tile = TMEM lane
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int tile = tile0; tile < num_tiles; tile += tile_step) {
            mbar_wait(&tmem_full_bar[acc], acc_phase);
            tc_fence_after();
            const int n = tile * GEMM_BM + row;
            const uint32_t t_row = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(acc * acc_stride);
            if (rank != 0) {
                // peer: ship the partial accumulator to the leader's reduction buffer
                const uint32_t dst = smem_u32(red_buf) + (uint32_t)(((int)rank - 1) * red_bytes_per_peer) +
                                     (uint32_t)(row * T * 4);
#pragma unroll 1
                for (int c = 0; c < T; c += 32) {
                    uint32_t u[32];
                    tmem_ld_32x32(t_row + (uint32_t)c, u);
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        if (c + j < T)
                            st_cluster_f32x4(dst + (uint32_t)((c + j) * 4), 0,
                                             make_float4(__uint_as_float(u[j]), __uint_as_float(u[j + 1]),
                                                         __uint_as_float(u[j + 2]), __uint_as_float(u[j + 3])));
                }
                __syncwarp();
                if (lane == 0) mbar_arrive_release_cluster(red_bar, 0);
            } else {
                if (splits > 1) mbar_wait_acquire_cluster(red_bar, 0);
#pragma unroll 1
                for (int c = 0; c < T; c += 32) {
                    uint32_t u[32];
                    tmem_ld_32x32(t_row + (uint32_t)c, u);
                    tmem_ld_wait();
                    float r[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) r[j] = __uint_as_float(u[j]);
                    for (int p = 0; p + 1 < splits; ++p) {      // partial sums in rank order: deterministic
                        const float4* src = reinterpret_cast<const float4*>(
                            reinterpret_cast<const uint8_t*>(red_buf) + p * red_bytes_per_peer + row * T * 4 + c * 4);
#pragma unroll
                        for (int j = 0; j < 32; j += 4) {
                            if (c + j < T) {
                                const float4 q = src[j >> 2];
                                r[j] += q.x; r[j + 1] += q.y; r[j + 2] += q.z; r[j + 3] += q.w;
                            }
                        }
                    }
                    skinny_store32<EPI>(r, out, resid, n, c, M, N, lane);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, tmem_cols);
    }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;
static std::mutex g_map_mu;

struct MapKey {
    const void* ptr;
    uint64_t rows, cols;
    uint32_t box_rows;
    bool operator==(const MapKey& o) const {
        return ptr == o.ptr && rows == o.rows && cols == o.cols && box_rows == o.box_rows;
    }
};
struct MapKeyHash {
    size_t operator()(const MapKey& k) const {
        size_t h = std::hash<const void*>()(k.ptr);
        h ^= std::hash<uint64_t>()(k.rows * 1000003ull + k.cols) + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2);
        h ^= std::hash<uint32_t>()(k.box_rows) + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2);
        return h;
    }
};
static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> g_maps;

static int get_tensor_map(const void* ptr, uint64_t rows, uint64_t cols, uint32_t box_rows, CUtensorMap* out) {
    std::lock_guard<std::mutex> lk(g_map_mu);
    if (!g_encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        KB_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
        KB_CHECK(fn != nullptr && qres == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled not available");
        g_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
    }
    MapKey key{ptr, rows, cols, box_rows};
    auto it = g_maps.find(key);
    if (it != g_maps.end()) {
        *out = it->second;
        return KB_OK;
    }
    if (g_maps.size() > 65536) g_maps.clear();
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * 2};
    cuuint32_t box[2] = {(cuuint32_t)GEMM_BK, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUtensorMap m;
    CUresult r = g_encode(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    KB_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d) rows=%llu cols=%llu box=%u", (int)r,
             (unsigned long long)rows, (unsigned long long)cols, box_rows);
    g_maps.emplace(key, m);
    *out = m;
    return KB_OK;
}

static int g_num_sms = 0;
static int num_sms() {
    if (g_num_sms == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
        if (g_num_sms <= 0) g_num_sms = 148;
    }
    return g_num_sms;
}

template <int BN, int EPI>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, void* out, const float* resid, int M, int N,
                       int K, uint32_t a_is_bf16, cudaStream_t stream) {
    using Cfg = GemmCfg<BN>;
    static bool configured[64] = {};
    auto kern = gemm_tcgen05_kernel<BN, EPI>;
    int dev = 0;
    KB_CUDA(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        KB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
        configured[dev & 63] = true;
    }
    const int num_tiles = ((M + GEMM_BM - 1) / GEMM_BM) * ((N + BN - 1) / BN);
    int grid = num_tiles < num_sms() ? num_tiles : num_sms();
    KB_CUDA(launch_kernel(kern, dim3(grid), dim3(GEMM_THREADS), Cfg::SMEM_BYTES, stream, tmA, tmB, out, resid, M, N, K,
                          a_is_bf16));
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

template <int BN>
static int dispatch_epi(const CUtensorMap& tmA, const CUtensorMap& tmB, void* out, const float* resid, int M, int N,
                        int K, int epi, uint32_t ab, cudaStream_t stream) {
    switch (epi) {
        case KB_EPI_BF16: return launch_gemm<BN, KB_EPI_BF16>(tmA, tmB, out, resid, M, N, K, ab, stream);
        case KB_EPI_F32: return launch_gemm<BN, KB_EPI_F32>(tmA, tmB, out, resid, M, N, K, ab, stream);
        case KB_EPI_RESID_F32: return launch_gemm<BN, KB_EPI_RESID_F32>(tmA, tmB, out, resid, M, N, K, ab, stream);
        case KB_EPI_SWIGLU_BF16: return launch_gemm<BN, KB_EPI_SWIGLU_BF16>(tmA, tmB, out, resid, M, N, K, ab, stream);
        case KB_EPI_SWIGLU_F16: return launch_gemm<BN, KB_EPI_SWIGLU_F16>(tmA, tmB, out, resid, M, N, K, ab, stream);
        case KB_EPI_F16: return launch_gemm<BN, KB_EPI_F16>(tmA, tmB, out, resid, M, N, K, ab, stream);
    }
    set_error("kb_gemm: unknown epilogue %d", epi);
    return KB_ERR_INVALID;
}

static std::atomic<unsigned long long> g_splitk_epoch{1};   // tag of a split-K launch = epoch * 16

template <int EPI>
static int launch_gemm_2cta(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmB2, void* out,
                            const float* resid, int M, int N, int K, int bn, int splits, uint32_t is_bf16,
                            cudaStream_t stream) {
    using Cfg = Gemm2Cfg;
    static bool configured[64] = {};
    auto kern = gemm_tcgen05_2cta_kernel<EPI>;
    int dev = 0;
    KB_CUDA(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        KB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES_WIDE));
        configured[dev & 63] = true;
    }
    const int smem_bytes = bn > Cfg::BN_MAX ? Cfg::SMEM_BYTES_WIDE : Cfg::SMEM_BYTES;
    const int num_tiles = ((M + 2 * GEMM_BM - 1) / (2 * GEMM_BM)) * ((N + bn - 1) / bn);
    if (splits > 1) {
        KB_CHECK(EPI == KB_EPI_RESID_F32 && resid == out, "kb_gemm: split-K needs the in-place residual epilogue");
        KB_CHECK((K / GEMM_BK) % splits == 0 && splits <= 8, "kb_gemm: K / 64 = %d k-blocks do not divide into %d parts",
                 K / GEMM_BK, splits);
        KB_CHECK(num_tiles <= SPLITK_MAX_TILES, "kb_gemm: %d tiles exceed the split-K hand-over table (%d)", num_tiles,
                 SPLITK_MAX_TILES);
    }
    const unsigned long long tag = splits > 1 ? g_splitk_epoch.fetch_add(1) * 16ull : 0ull;
    const int num_units = num_tiles * splits;
    const int max_clusters = num_sms() / 2;
    const int clusters = num_units < max_clusters ? num_units : max_clusters;
    KB_CUDA(launch_kernel(kern, dim3(2 * clusters), dim3(GEMM_THREADS), smem_bytes, stream, tmA, tmB, tmB2, out, resid,
                          M, N, K, bn, is_bf16, splits, tag));
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

static int dispatch_epi_2cta(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmB2, void* out,
                             const float* resid, int M, int N, int K, int bn, int splits, int epi, uint32_t ab,
                             cudaStream_t stream) {
    switch (epi) {
        case KB_EPI_BF16: return launch_gemm_2cta<KB_EPI_BF16>(tmA, tmB, tmB2, out, resid, M, N, K, bn, 1, ab, stream);
        case KB_EPI_F32: return launch_gemm_2cta<KB_EPI_F32>(tmA, tmB, tmB2, out, resid, M, N, K, bn, 1, ab, stream);
        case KB_EPI_RESID_F32:
            return launch_gemm_2cta<KB_EPI_RESID_F32>(tmA, tmB, tmB2, out, resid, M, N, K, bn, splits, ab, stream);
        case KB_EPI_SWIGLU_BF16:
            return launch_gemm_2cta<KB_EPI_SWIGLU_BF16>(tmA, tmB, tmB2, out, resid, M, N, K, bn, 1, ab, stream);
        case KB_EPI_SWIGLU_F16:
            return launch_gemm_2cta<KB_EPI_SWIGLU_F16>(tmA, tmB, tmB2, out, resid, M, N, K, bn, 1, ab, stream);
        case KB_EPI_F16: return launch_gemm_2cta<KB_EPI_F16>(tmA, tmB, tmB2, out, resid, M, N, K, bn, 1, ab, stream);
    }
    set_error("kb_gemm: unknown epilogue %d", epi);
    return KB_ERR_INVALID;
}

// Small-M path: T = padded token count, splits = CTAs per tile (cluster split-K), chosen so that at least
// ~96 CTAs stream weights and the reduction buffers fit beside a >= 4-stage ring.
static void skinny_plan(int M, int N, int K, int want_splits, int* T_out, int* splits_out, int* stages_out) {
    const int T = M <= 16 ? 16 : (M <= 32 ? 32 : (M <= 64 ? 64 : 128));
    const int tiles = (N + GEMM_BM - 1) / GEMM_BM;
    const int num_k = K / GEMM_BK;
    const int stage_bytes = SkinnyCfg::W_BYTES + T * GEMM_BK * 2;
    int splits = 1;
    if (want_splits > 0) {
        splits = want_splits;
    } else {
        while (splits < 4 && tiles * splits < 96) splits *= 2;
    }
    while (splits > 1 && (num_k % splits != 0 || tiles * splits > 65535 ||
                          (SkinnyCfg::SMEM_BUDGET - (splits - 1) * GEMM_BM * T * 4) / stage_bytes < 4))
        splits /= 2;
    int stages = (SkinnyCfg::SMEM_BUDGET - (splits - 1) * GEMM_BM * T * 4) / stage_bytes;
    if (stages > SkinnyCfg::MAX_STAGES) stages = SkinnyCfg::MAX_STAGES;
    *T_out = T;
    *splits_out = splits;
    *stages_out = stages;
}

template <int EPI>
static int launch_skinny(const void* a, const void* w, void* out, const float* resid, int M, int N, int K,
                         int want_splits, uint32_t is_bf16, cudaStream_t stream) {
    int T, splits, stages;
    skinny_plan(M, N, K, want_splits, &T, &splits, &stages);
    CUtensorMap tmW, tmX;
    int rc = get_tensor_map(w, (uint64_t)N, (uint64_t)K, GEMM_BM, &tmW);
    if (rc) return rc;
    rc = get_tensor_map(a, (uint64_t)M, (uint64_t)K, (uint32_t)T, &tmX);
    if (rc) return rc;
    const size_t smem = (size_t)stages * (SkinnyCfg::W_BYTES + T * GEMM_BK * 2) + (size_t)(splits - 1) * GEMM_BM * T * 4 +
                        SkinnyCfg::BAR_BYTES + 1024;
    static bool configured[64] = {};
    auto kern = gemm_tcgen05_skinny_kernel<EPI>;
    int dev = 0;
    KB_CUDA(cudaGetDevice(&dev));
    if (!configured[dev & 63]) {
        KB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     SkinnyCfg::SMEM_BUDGET + SkinnyCfg::BAR_BYTES + 1024));
        configured[dev & 63] = true;
    }
    const int tiles = (N + GEMM_BM - 1) / GEMM_BM;
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    int n_attr = 0;
    if (pdl_enabled()) {
        attr[n_attr].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n_attr].val.programmaticStreamSerializationAllowed = 1;
        ++n_attr;
    }
    if (splits > 1) {
        cfg.gridDim = dim3(tiles * splits);
        attr[n_attr].id = cudaLaunchAttributeClusterDimension;
        attr[n_attr].val.clusterDim.x = splits;
        attr[n_attr].val.clusterDim.y = 1;
        attr[n_attr].val.clusterDim.z = 1;
        ++n_attr;
    } else {
        cfg.gridDim = dim3(tiles < num_sms() ? tiles : num_sms());
    }
    cfg.attrs = attr;
    cfg.numAttrs = n_attr;
    KB_CUDA(cudaLaunchKernelEx(&cfg, kern, tmW, tmX, out, resid, M, N, K, T, stages, splits, is_bf16));
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

static int dispatch_skinny(const void* a, const void* w, void* out, const float* resid, int M, int N, int K, int epi,
                           int want_splits, uint32_t ab, cudaStream_t stream) {
    switch (epi) {
        case KB_EPI_BF16: return launch_skinny<KB_EPI_BF16>(a, w, out, resid, M, N, K, want_splits, ab, stream);
        case KB_EPI_F32: return launch_skinny<KB_EPI_F32>(a, w, out, resid, M, N, K, want_splits, ab, stream);
        case KB_EPI_RESID_F32: return launch_skinny<KB_EPI_RESID_F32>(a, w, out, resid, M, N, K, want_splits, ab, stream);
        case KB_EPI_SWIGLU_BF16:
            return launch_skinny<KB_EPI_SWIGLU_BF16>(a, w, out, resid, M, N, K, want_splits, ab, stream);
        case KB_EPI_SWIGLU_F16:
            return launch_skinny<KB_EPI_SWIGLU_F16>(a, w, out, resid, M, N, K, want_splits, ab, stream);
        case KB_EPI_F16: return launch_skinny<KB_EPI_F16>(a, w, out, resid, M, N, K, want_splits, ab, stream);
    }
    set_error("kb_gemm: unknown epilogue %d", epi);
    return KB_ERR_INVALID;
}

// Tile choice: minimise the time of the busiest CTA (or CTA pair) under the static persistent schedule.
// Returns BN for the 1-CTA kernel, -BN for the CTA-pair kernel; *splits_out > 1 = split-K work units (only
// offered when `can_split`: the in-place residual epilogue, see g_splitk_flags).
static int pick_bn(int M, int N, int K, bool can_split, bool allow_wide, int* splits_out) {
    *splits_out = 1;
    static int forced = -1, no_split = -1;
    if (forced < 0) {
        const char* s = getenv("KB_GEMM_BN");
        forced = s ? atoi(s) : 0;
        no_split = getenv("KB_GEMM_NO_SPLITK") != nullptr;
    }
    if (forced == 64 || forced == 128 || forced == 256 || (forced <= -32 && forced >= -256 && forced % 16 == 0) ||
        (forced < -256 && forced >= -512 && forced % 16 == 0 && M > GEMM_BM))
        return forced;
    const int sms = num_sms();
    const int num_m = (M + GEMM_BM - 1) / GEMM_BM;
    const int num_k = K / GEMM_BK;
    // cycles per k-block of one CTA: `rows` 128-byte rows of operands staged by TMA at ~97 GB/s per SM (~50 B/clk at the boost clock)
    // (measured: tools/gemm_trace.py, profiles/r01_gemm_bench.md) against 2 * bn cycles of MMA
    auto kblock = [](int rows, int bn) { const long ing = (long)rows * 128 * 100 / 5000; return ing > 2L * bn ? ing : 2L * bn; };
    auto epilogue = [](int bn) { return 1200L + 6L * bn; };
    long best_cost = -1;
    int best = 128;
    const int cands[3] = {256, 128, 64};
    for (int i = 0; i < 3; ++i) {
        const int bn = cands[i];
        const long tiles = (long)num_m * ((N + bn - 1) / bn);
        const long per_cta = (tiles + sms - 1) / sms;
        // a 1-CTA tile brings 128 rows of A + bn rows of B per k-block: wide tiles win unless they leave SMs idle
        const long cost = per_cta * (num_k * kblock(128 + bn, bn) + epilogue(bn));
        if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = bn; }
    }
    if (M > GEMM_BM) {
        // CTA pair, 256 x bn tile: each CTA stages 128 rows of A and bn / 2 rows of B per k-block.
        // ... any multiple of 16 (run-time tile width): take the width whose unit count wastes the least of the
        // last round; wider wins ties (less A re-read). With split-K a unit is 1 / splits of a tile's K range and
        // the epilogues of a tile's units run one after the other.
        const int num_mp = (M + 2 * GEMM_BM - 1) / (2 * GEMM_BM);
        const int pairs = sms / 2;
        for (int splits = 1; splits <= (can_split && !no_split ? 4 : 1); splits *= 2) {
            // measured (tools/gemm_bench.py, M = 512, profiles/r02_gemm_splitk.md): splitting pays when a unit
            // still has a long main loop to amortise the second, ordered epilogue — the down projection
            // (224 k-blocks: 58.9 -> 54.5 us); it loses on the O projection (64 k-blocks: 19.7 -> 24.4 us)
            if (num_k % splits != 0 || (splits > 1 && num_k / splits < 96)) break;
            for (int bn = 256; bn >= 64; bn -= 16) {
                const long tiles = (long)num_mp * ((N + bn - 1) / bn);
                if (splits > 1 && tiles > SPLITK_MAX_TILES) continue;
                // ... and only while the tiles alone cannot fill the machine: with 256 tiles on 74 pairs
                // (M = 4 096) two parts per tile cost 423 us against 386 without (tools/gemm_bench.py)
                if (splits > 1 && tiles >= pairs) continue;
                const long rounds = (tiles * splits + pairs - 1) / pairs;
                const long cost = rounds * ((num_k / splits) * kblock(128 + bn / 2, bn) + epilogue(bn)) +
                                  (splits - 1) * epilogue(bn);
                if (cost < best_cost) { best_cost = cost; best = -bn; *splits_out = splits; }
            }
        }
        // Wide tiles (256 < bn <= 512, one accumulator, two MMAs per k-step; see Gemm2Cfg). Per k-block both kinds
        // are MMA-bound at these widths (~95 ns per 256 columns: 398 ns measured at bn = 256, 752 ns at bn = 480);
        // (256 + 2 bn) and (384 + 2 bn) below are those costs in units that only need to be comparable with each
        // other — a wide tile wins through fewer, fuller rounds, not through faster math.
        // Sustained: the down projection (K = 14 336) gains 7 - 12 % from M = 2 048 up (806 -> 737 us at M = 8 192),
        // loses at M = 1 024; K = 4 096 shapes gain nothing (the epilogue of a wide tile does not hide under the next
        // main loop), SwiGLU loses.
        if (allow_wide && best < 0 && *splits_out == 1 && num_k >= 128 && num_mp >= 8) {
            const int nb = -best;
            const long n_rounds = ((long)num_mp * ((N + nb - 1) / nb) + pairs - 1) / pairs;
            const long narrow = n_rounds * (256 + 2 * nb);
            long wbest = -1;
            int wbn = 0;
            for (int bn = 512; bn >= 448; bn -= 16) {
                const long rounds = ((long)num_mp * ((N + bn - 1) / bn) + pairs - 1) / pairs;
                const long c = rounds * (384 + 2 * bn);
                if (wbest < 0 || c < wbest) { wbest = c; wbn = bn; }
            }
            if (wbest * 93 < narrow * 100) best = -wbn;
        }
    }
    return best;
}

int gemm_ex(const void* a, const void* w, void* out, const float* resid, int M, int N, int K, int epi_flags,
            int tile_n, cudaStream_t stream) {
    const int epi = epi_flags & 0xff;
    const uint32_t a_is_bf16 = (epi_flags & KB_GEMM_F16) ? 0u : 1u;  // both operands share the format
    KB_CHECK(M > 0 && N > 0 && K > 0, "kb_gemm: bad shape M=%d N=%d K=%d", M, N, K);
    KB_CHECK(K % GEMM_BK == 0, "kb_gemm: K=%d must be a multiple of %d", K, GEMM_BK);
    KB_CHECK(N % 8 == 0, "kb_gemm: N=%d must be a multiple of 8", N);
    KB_CHECK((epi != KB_EPI_SWIGLU_BF16 && epi != KB_EPI_SWIGLU_F16) || N % 16 == 0,
             "kb_gemm: SwiGLU needs N %% 16 == 0");
    KB_CHECK(epi != KB_EPI_RESID_F32 || resid != nullptr, "kb_gemm: residual epilogue needs resid");
    KB_CHECK((reinterpret_cast<uintptr_t>(a) & 15) == 0 && (reinterpret_cast<uintptr_t>(w) & 15) == 0 &&
                 (reinterpret_cast<uintptr_t>(out) & 15) == 0,
             "kb_gemm: pointers must be 16-byte aligned");
    int forced_splits = 1;
    if (tile_n <= -1024) {          // -(1024 * splits + bn): CTA-pair tile of width bn with split-K (tests, tuning)
        forced_splits = (-tile_n) / 1024;
        tile_n = -((-tile_n) % 1024);
        KB_CHECK(forced_splits == 2 || forced_splits == 4, "kb_gemm: split-K takes 2 or 4 parts (got %d)", forced_splits);
    }
    KB_CHECK(tile_n == 0 || tile_n == 64 || tile_n == 128 || tile_n == 256 ||
                 (tile_n <= -32 && tile_n >= -Gemm2Cfg::BN_WIDE_MAX && tile_n % 16 == 0) ||
                 (tile_n >= 1 && tile_n <= 4 && tile_n != 3),
             "kb_gemm: tile_n=%d not in {0, 64, 128, 256}, a negative multiple of 16 in [-512, -32], or 1 / 2 / 4 "
             "(small-M kernel with that many CTAs per tile)", tile_n);
    static int skinny_off = -1;
    if (skinny_off < 0) skinny_off = getenv("KB_GEMM_NO_SKINNY") != nullptr;
    // measured in the decode step (bench.py --threads 32 / 64 / 128, with and without KB_GEMM_NO_SKINNY):
    // the small-M kernel wins up to 32 tokens (6.1 vs 6.6 ms per step), ties at 64, loses at 128
    // ... and between 33 and 64 tokens it still wins where the weight matrix is large in N or K (gate/up, down,
    // lm_head: many tiles or a long K to split), and loses on the square QKV / O projections
    const bool skinny_mid = M <= 64 && (N >= 16384 || K >= 8192);
    if ((tile_n >= 1 && tile_n <= 4) || (tile_n == 0 && (M <= 32 || skinny_mid) && !skinny_off)) {
        KB_CHECK(M <= GEMM_BM, "kb_gemm: the small-M kernel takes at most %d rows (got %d)", GEMM_BM, M);
        // tile_n = 0: CTAs per tile chosen by skinny_plan; 1 / 2 / 4: forced (benchmarks, tests)
        return dispatch_skinny(a, w, out, resid, M, N, K, epi, tile_n, a_is_bf16, stream);
    }
    int splits = forced_splits;
    const bool in_place = epi == KB_EPI_RESID_F32 && resid == reinterpret_cast<const float*>(out);
    KB_CHECK(forced_splits == 1 || (in_place && tile_n < 0),
             "kb_gemm: split-K needs the CTA-pair kernel and the in-place residual epilogue (KB_EPI_RESID_F32, out == resid)");
    const bool swiglu = epi == KB_EPI_SWIGLU_BF16 || epi == KB_EPI_SWIGLU_F16;
    const int bn = tile_n ? tile_n : pick_bn(M, N, K, in_place, !swiglu, &splits);
    CUtensorMap tmA, tmB;
    int rc = get_tensor_map(a, (uint64_t)M, (uint64_t)K, GEMM_BM, &tmA);
    if (rc) return rc;
    if (bn < 0) {   // CTA pair: every CTA stages half of the tile's B rows
        const int w_ = -bn;
        if (w_ > Gemm2Cfg::BN_MAX) {   // wide tile: half of the first 256 columns and half of the rest, two loads
            CUtensorMap tmB2;
            rc = get_tensor_map(w, (uint64_t)N, (uint64_t)K, (uint32_t)(Gemm2Cfg::BN_MAX / 2), &tmB);
            if (rc) return rc;
            rc = get_tensor_map(w, (uint64_t)N, (uint64_t)K, (uint32_t)((w_ - Gemm2Cfg::BN_MAX) / 2), &tmB2);
            if (rc) return rc;
            return dispatch_epi_2cta(tmA, tmB, tmB2, out, resid, M, N, K, w_, splits, epi, a_is_bf16, stream);
        }
        rc = get_tensor_map(w, (uint64_t)N, (uint64_t)K, (uint32_t)(w_ / 2), &tmB);
        if (rc) return rc;
        return dispatch_epi_2cta(tmA, tmB, tmB, out, resid, M, N, K, w_, splits, epi, a_is_bf16, stream);
    }
    rc = get_tensor_map(w, (uint64_t)N, (uint64_t)K, (uint32_t)bn, &tmB);
    if (rc) return rc;
    if (bn == 256) return dispatch_epi<256>(tmA, tmB, out, resid, M, N, K, epi, a_is_bf16, stream);
    if (bn == 128) return dispatch_epi<128>(tmA, tmB, out, resid, M, N, K, epi, a_is_bf16, stream);
    return dispatch_epi<64>(tmA, tmB, out, resid, M, N, K, epi, a_is_bf16, stream);
}

int gemm(const void* a, const void* w, void* out, const float* resid, int M, int N, int K, int epi,
         cudaStream_t stream) {
    return gemm_ex(a, w, out, resid, M, N, K, epi, 0, stream);
}

}  // namespace kb

extern "C" int kb_gemm_ex(const void* a_bf16, const void* w_bf16, void* out, const float* resid, int32_t M,
                          int32_t N, int32_t K, int32_t epilogue, int32_t tile_n, kb_stream stream) {
    return kb::gemm_ex(a_bf16, w_bf16, out, resid, M, N, K, epilogue, tile_n, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int kb_gemm(const void* a_bf16, const void* w_bf16, void* out, const float* resid, int32_t M, int32_t N,
                       int32_t K, int32_t epilogue, kb_stream stream) {
    return kb::gemm(a_bf16, w_bf16, out, resid, M, N, K, epilogue, reinterpret_cast<cudaStream_t>(stream));
}

#ifdef KB_GEMM_TRACE
extern "C" int kb_debug_gemm_trace(unsigned long long* host_out16) {
    return cudaMemcpyFromSymbol(host_out16, kb::g_gemm_trace, sizeof(unsigned long long) * 16) == cudaSuccess ? 0 : 1;
}
#endif
