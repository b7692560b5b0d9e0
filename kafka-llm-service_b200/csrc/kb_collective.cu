// kb_collective.cu — the tensor-parallel all-reduce as ONE kernel over peer memory (NVLink 5 / NVSwitch),
// used by the engine after the two row-sharded projections of every layer (O and down).
//
// Every rank owns 1/world of the [T, hidden] fp32 buffer. The kernel
//   1. meets the same block of every other rank at a flag barrier (all partial sums are written),
//   2. reduces ITS slice over all ranks — with a multicast mapping one `multimem.ld_reduce` per 16 bytes
//      makes the switch fetch and add the eight copies (NVLS); without one it loads the slice from each
//      peer's mapping and adds in rank order,
//   3. writes the sums to every rank (`multimem.st`, or one store per peer),
//   4. meets the other ranks again (release / acquire at system scope: the sums are visible everywhere).
// Per GPU this moves 1/world of the buffer in and 1x out instead of the ring's 2 x (world-1)/world
// both ways, and it needs no second kernel or host round trip between the phases.
//
// Barriers: only block 0 of a rank talks to the other ranks (one 4-byte slot per source rank in each
// rank's flag area, holding monotonically increasing epochs — two per call, never reset, so a fast rank
// entering the next call cannot confuse a slow one). The other blocks of the grid wait for block 0 on a
// LOCAL flag at the start and report to it through a local counter at the end, so the NVLink flag
// traffic does not grow with the grid. Grids stay at or below KB_TP_MAX_BLOCKS (< SM count): all blocks
// of all ranks are co-resident, the condition for a spinning barrier. A rank that waits longer than
// ~2 s traps (the process fails loudly instead of hanging the box).
//
// Reference counterpart: none (the reference's models sit behind a gateway, src/llm/portkey.py:379).
#include "kb_common.cuh"

#include <stdlib.h>

namespace kb {

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 multimem_ld_reduce_f32x4(const void* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st_f32x4(void* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y),
                 "f"(v.z), "f"(v.w)
                 : "memory");
}
// four f16x2 words, added in fp32 inside the switch
__device__ __forceinline__ float4 multimem_ld_reduce_f16x8(const void* mc) {
    uint32_t a, b, c, d;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(a), "=r"(b), "=r"(c), "=r"(d)
                 : "l"(mc)
                 : "memory");
    return make_float4(__uint_as_float(a), __uint_as_float(b), __uint_as_float(c), __uint_as_float(d));
}
__device__ __forceinline__ float4 ld_relaxed_sys_f32x4(const void* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_sys_f32x4(void* p, float4 v) {
    asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w)
                 : "memory");
}

__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// flag area of a rank (uint32 indices): [0, 8) epoch written by rank i | [32] "go" for this rank's own
// blocks | [64] arrival counter of this rank's own blocks  (separate 128-byte lines)
#define KB_FLAG_GO 32
#define KB_FLAG_COUNT 64

__device__ __forceinline__ uint32_t* flag_area(const kb_tp_peers& P, int r) {
    return reinterpret_cast<uint32_t*>(static_cast<char*>(P.peer_base[r]) + P.flags_offset);
}

__device__ __noinline__ void barrier_timeout(int rank, int peer, uint32_t epoch) {
    printf("kb_collective: rank %d timed out waiting for %s %d (epoch %u)\n", rank, peer >= 0 ? "rank" : "block",
           peer, epoch);
    __trap();
}

// Thread `peer` (< world) of block 0: tell rank `peer` that this rank reached `epoch`, wait for its word.
__device__ __forceinline__ void meet_ranks(const kb_tp_peers& P, int rank, int world, uint32_t epoch) {
    if (threadIdx.x < (unsigned)world) {
        const int peer = threadIdx.x;
        st_release_sys(flag_area(P, peer) + rank, epoch);
        const uint32_t* mine = flag_area(P, rank) + peer;
        const long long t0 = clock64();
        while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0)
            if (clock64() - t0 > (long long)P.timeout_cycles) barrier_timeout(rank, peer, epoch);   // a peer never arrived
    }
}

// Start of a collective: every rank's inputs are in place (and its outputs of the previous call consumed).
__device__ __forceinline__ void barrier_begin(const kb_tp_peers& P, int rank, int world, uint32_t epoch) {
    uint32_t* mine = flag_area(P, rank);
    if (blockIdx.x == 0) {
        meet_ranks(P, rank, world, epoch);
        __syncthreads();
        if (threadIdx.x == 0) st_release_gpu(mine + KB_FLAG_GO, epoch);
    } else {
        if (threadIdx.x == 0) {
            const long long t0 = clock64();
            while ((int32_t)(ld_acquire_gpu(mine + KB_FLAG_GO) - epoch) < 0)
                if (clock64() - t0 > (long long)P.timeout_cycles) barrier_timeout(rank, -1, epoch);
        }
        __syncthreads();
    }
}

// End of a collective: everything every block of every rank wrote is visible everywhere. Blocks other
// than 0 only report and leave; block 0 collects them, then meets the other ranks (the kernel cannot
// complete before that).
__device__ __forceinline__ void barrier_end(const kb_tp_peers& P, int rank, int world, uint32_t epoch) {
    uint32_t* mine = flag_area(P, rank);
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        atomicAdd(mine + KB_FLAG_COUNT, 1u);
    }
    if (blockIdx.x != 0) return;
    if (threadIdx.x == 0) {
        const long long t0 = clock64();
        while (ld_acquire_gpu(mine + KB_FLAG_COUNT) < gridDim.x)
            if (clock64() - t0 > (long long)P.timeout_cycles) barrier_timeout(rank, -2, epoch);
        mine[KB_FLAG_COUNT] = 0;   // all blocks have arrived; the next call starts after this kernel
        __threadfence_system();
    }
    __syncthreads();
    meet_ranks(P, rank, world, epoch);
    __syncthreads();
}

__device__ __forceinline__ float4 add_vec(float4 a, float4 b, int f16) {
    if (!f16) return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
    float4 r;
    const __half2* ha = reinterpret_cast<const __half2*>(&a);
    const __half2* hb = reinterpret_cast<const __half2*>(&b);
    __half2* hr = reinterpret_cast<__half2*>(&r);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float2 x = __half22float2(ha[k]), y = __half22float2(hb[k]);
        hr[k] = __floats2half2_rn(x.x + y.x, x.y + y.y);
    }
    return r;
}

// F16 == false: 16-byte vectors of 4 fp32;  F16 == true: of 8 fp16 (payload moved as raw 32-bit lanes)
template <bool MULTICAST, bool F16>
__global__ void __launch_bounds__(512) allreduce_kernel(const kb_tp_peers P, int rank, int world,
                                                            size_t offset_bytes, size_t n_vec, uint32_t epoch) {
    barrier_begin(P, rank, world, epoch);         // every rank's partial sums are in place

    const size_t per = (n_vec + world - 1) / world;   // 16-byte vectors per rank
    const size_t lo = per * rank;
    const size_t hi = lo + per < n_vec ? lo + per : n_vec;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    if (MULTICAST) {
        char* mc = static_cast<char*>(P.mc_base) + offset_bytes;
        size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
        for (; i + 3 * stride < hi; i += 4 * stride) {    // four independent reductions in flight per thread
            float4 v0 = F16 ? multimem_ld_reduce_f16x8(mc + 16 * i) : multimem_ld_reduce_f32x4(mc + 16 * i);
            float4 v1 = F16 ? multimem_ld_reduce_f16x8(mc + 16 * (i + stride))
                            : multimem_ld_reduce_f32x4(mc + 16 * (i + stride));
            float4 v2 = F16 ? multimem_ld_reduce_f16x8(mc + 16 * (i + 2 * stride))
                            : multimem_ld_reduce_f32x4(mc + 16 * (i + 2 * stride));
            float4 v3 = F16 ? multimem_ld_reduce_f16x8(mc + 16 * (i + 3 * stride))
                            : multimem_ld_reduce_f32x4(mc + 16 * (i + 3 * stride));
            multimem_st_f32x4(mc + 16 * i, v0);
            multimem_st_f32x4(mc + 16 * (i + stride), v1);
            multimem_st_f32x4(mc + 16 * (i + 2 * stride), v2);
            multimem_st_f32x4(mc + 16 * (i + 3 * stride), v3);
        }
        for (; i < hi; i += stride)
            multimem_st_f32x4(mc + 16 * i,
                              F16 ? multimem_ld_reduce_f16x8(mc + 16 * i) : multimem_ld_reduce_f32x4(mc + 16 * i));
    } else {
        for (size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += stride) {
            float4 part[KB_TP_MAX_RANKS];
#pragma unroll
            for (int r = 0; r < KB_TP_MAX_RANKS; ++r)
                if (r < world)
                    part[r] = ld_relaxed_sys_f32x4(static_cast<char*>(P.peer_base[r]) + offset_bytes + 16 * i);
            float4 s = part[0];
#pragma unroll
            for (int r = 1; r < KB_TP_MAX_RANKS; ++r)
                if (r < world) s = add_vec(s, part[r], F16);
#pragma unroll
            for (int r = 0; r < KB_TP_MAX_RANKS; ++r)
                if (r < world) st_relaxed_sys_f32x4(static_cast<char*>(P.peer_base[r]) + offset_bytes + 16 * i, s);
        }
    }

    barrier_end(P, rank, world, epoch + 1);       // every slice has been written to every rank
}

// The region descriptor as the kernels get it: the caller's, with the timeout filled in.
static kb_tp_peers with_timeout(const kb_tp_peers* P) {
    static unsigned long long dflt = 0;
    if (dflt == 0) {
        const char* s = getenv("KB_TP_TIMEOUT_S");
        double sec = s ? atof(s) : 30.0;
        if (!(sec > 0.0)) sec = 30.0;
        dflt = (unsigned long long)(sec * 2.0e9);      // SM clock <= ~2 GHz: at least `sec` seconds
    }
    kb_tp_peers Q = *P;
    if (Q.timeout_cycles == 0) Q.timeout_cycles = dflt;
    return Q;
}

int allreduce_peers(const kb_tp_peers* P, int rank, int world, size_t offset_bytes, size_t n, int f16, uint32_t epoch,
                    int blocks, cudaStream_t stream) {
    KB_CHECK(P != nullptr && world >= 2 && world <= KB_TP_MAX_RANKS && rank >= 0 && rank < world,
             "kb_allreduce_peers: bad rank %d of %d", rank, world);
    const size_t bytes = n * (f16 ? 2 : 4);
    KB_CHECK(n > 0 && bytes % 16 == 0 && offset_bytes % 16 == 0, "kb_allreduce_peers: size and offset must be 16-byte multiples");
    KB_CHECK(offset_bytes + bytes <= P->flags_offset && P->flags_offset + KB_TP_FLAG_BYTES <= P->bytes,
             "kb_allreduce_peers: [%zu, +%zu) overlaps the flag area at %zu", offset_bytes, bytes, P->flags_offset);
    for (int r = 0; r < world; ++r) KB_CHECK(P->peer_base[r] != nullptr, "kb_allreduce_peers: no mapping of rank %d", r);
    // measured on 8 B200 (tools/allreduce_bench.py): the switch saturates with 8 blocks of multimem
    // traffic; plain peer loads want as many blocks as are allowed
    if (blocks <= 0) blocks = P->mc_base != nullptr ? (64 / world > 8 ? 64 / world : 8) : KB_TP_MAX_BLOCKS;
    if (blocks > KB_TP_MAX_BLOCKS) blocks = KB_TP_MAX_BLOCKS;
    const size_t n_vec = bytes / 16;
    const bool mc = P->mc_base != nullptr;
    const kb_tp_peers Q = with_timeout(P);
    if (mc && f16)
        allreduce_kernel<true, true><<<blocks, 512, 0, stream>>>(Q, rank, world, offset_bytes, n_vec, epoch);
    else if (mc)
        allreduce_kernel<true, false><<<blocks, 512, 0, stream>>>(Q, rank, world, offset_bytes, n_vec, epoch);
    else if (f16)
        allreduce_kernel<false, true><<<blocks, 512, 0, stream>>>(Q, rank, world, offset_bytes, n_vec, epoch);
    else
        allreduce_kernel<false, false><<<blocks, 512, 0, stream>>>(Q, rank, world, offset_bytes, n_vec, epoch);
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

// ---------------------------------------------------------------------------------------------
// reduce-scatter + residual add + RMSNorm + all-gather in ONE kernel (sequence-parallel residual).
//
// After a row-sharded projection every rank holds fp16 partial sums of the whole [T, hidden] output.
// Instead of all-reducing them and letting every rank repeat the same residual add and RMSNorm on all
// T rows, rank r owns rows [r*ceil(T/world), ...) of the fp32 residual stream:
//   d    = sum over ranks of partial[row]            (multimem.ld_reduce, fp32 accumulate in the switch)
//   x[row] += d                                      (local HBM, only the owner keeps these rows current)
//   xn[row] = fp16(x[row] * rsqrt(mean(x^2) + eps) * w)   -> multimem.st to EVERY rank's xn buffer
// so the wire carries the same bytes as the all-reduce (fp16 in, fp16 out), the norm is computed once
// per row instead of `world` times, and the next GEMM finds its A operand ready.
// One row per block iteration, VPT 16-byte vectors per thread, the next row's reductions are issued
// before the current row is normalised.
// ---------------------------------------------------------------------------------------------
template <bool MULTICAST>
__device__ __forceinline__ uint4 reduce_f16x8(const kb_tp_peers& P, int world, size_t byte_off) {
    if (MULTICAST) {
        const float4 v = multimem_ld_reduce_f16x8(static_cast<const char*>(P.mc_base) + byte_off);
        return make_uint4(__float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z), __float_as_uint(v.w));
    }
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int r = 0; r < KB_TP_MAX_RANKS; ++r) {
        if (r < world) {
            const float4 p = ld_relaxed_sys_f32x4(static_cast<const char*>(P.peer_base[r]) + byte_off);
            const __half2* h = reinterpret_cast<const __half2*>(&p);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float2 f = __half22float2(h[k]);
                acc[2 * k] += f.x;
                acc[2 * k + 1] += f.y;
            }
        }
    }
    return make_uint4(pack_f16(acc[0], acc[1]), pack_f16(acc[2], acc[3]), pack_f16(acc[4], acc[5]),
                      pack_f16(acc[6], acc[7]));
}

template <bool MULTICAST, int VPT>
__global__ void __launch_bounds__(256)
tp_reduce_norm_gather_kernel(const kb_tp_peers P, int rank, int world, size_t partial_off, size_t xn_off,
                             float* __restrict__ x, const __nv_bfloat16* __restrict__ w, int T, int hidden, float eps,
                             uint32_t epoch) {
    __shared__ float red[2][8];
    barrier_begin(P, rank, world, epoch);         // every rank's partial sums are in place, xn is free

    const int rows_per = (T + world - 1) / world;
    const int row_lo = rank * rows_per;
    const int row_hi = row_lo + rows_per < T ? row_lo + rows_per : T;
    const int vecs = hidden >> 3;                 // 16-byte vectors (8 fp16) per row
    const int tid = threadIdx.x;
    const size_t row_bytes = (size_t)hidden * 2;

    uint4 d[VPT];
    int row = row_lo + blockIdx.x;
    if (row < row_hi) {
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            const int idx = tid + k * 256;
            if (idx < vecs) d[k] = reduce_f16x8<MULTICAST>(P, world, partial_off + row * row_bytes + (size_t)idx * 16);
        }
    }
    int it = 0;
    for (; row < row_hi; row += gridDim.x, it ^= 1) {
        // issue the next row's reductions before working on this one
        uint4 dn[VPT];
        const int nrow = row + gridDim.x;
        if (nrow < row_hi) {
#pragma unroll
            for (int k = 0; k < VPT; ++k) {
                const int idx = tid + k * 256;
                if (idx < vecs)
                    dn[k] = reduce_f16x8<MULTICAST>(P, world, partial_off + nrow * row_bytes + (size_t)idx * 16);
            }
        }
        float v[VPT][8];
        float ss = 0.f;
        float* xrow = x + (size_t)row * hidden;
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            const int idx = tid + k * 256;
            if (idx < vecs) {
                const float4 a = *reinterpret_cast<const float4*>(xrow + idx * 8);
                const float4 b = *reinterpret_cast<const float4*>(xrow + idx * 8 + 4);
                const __half2* h = reinterpret_cast<const __half2*>(&d[k]);
                const float2 h0 = __half22float2(h[0]), h1 = __half22float2(h[1]), h2 = __half22float2(h[2]),
                             h3 = __half22float2(h[3]);
                v[k][0] = a.x + h0.x; v[k][1] = a.y + h0.y; v[k][2] = a.z + h1.x; v[k][3] = a.w + h1.y;
                v[k][4] = b.x + h2.x; v[k][5] = b.y + h2.y; v[k][6] = b.z + h3.x; v[k][7] = b.w + h3.y;
                *reinterpret_cast<float4*>(xrow + idx * 8) = make_float4(v[k][0], v[k][1], v[k][2], v[k][3]);
                *reinterpret_cast<float4*>(xrow + idx * 8 + 4) = make_float4(v[k][4], v[k][5], v[k][6], v[k][7]);
#pragma unroll
                for (int j = 0; j < 8; ++j) ss += v[k][j] * v[k][j];
            }
        }
        ss = warp_sum(ss);
        if ((tid & 31) == 0) red[it][tid >> 5] = ss;
        __syncthreads();
        float tot = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) tot += red[it][k];
        const float inv = rsqrtf(tot / (float)hidden + eps);
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            const int idx = tid + k * 256;
            if (idx < vecs) {
                const uint4 wv = *reinterpret_cast<const uint4*>(w + idx * 8);
                const __nv_bfloat162* wb = reinterpret_cast<const __nv_bfloat162*>(&wv);
                uint4 o;
                uint32_t* op = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 wf = __bfloat1622float2(wb[j]);
                    op[j] = pack_f16(v[k][2 * j] * inv * wf.x, v[k][2 * j + 1] * inv * wf.y);
                }
                const size_t off = xn_off + row * row_bytes + (size_t)idx * 16;
                const float4 of = make_float4(__uint_as_float(o.x), __uint_as_float(o.y), __uint_as_float(o.z),
                                              __uint_as_float(o.w));
                if (MULTICAST) {
                    multimem_st_f32x4(static_cast<char*>(P.mc_base) + off, of);
                } else {
#pragma unroll
                    for (int r = 0; r < KB_TP_MAX_RANKS; ++r)
                        if (r < world) st_relaxed_sys_f32x4(static_cast<char*>(P.peer_base[r]) + off, of);
                }
            }
        }
#pragma unroll
        for (int k = 0; k < VPT; ++k) d[k] = dn[k];
    }

    barrier_end(P, rank, world, epoch + 1);       // every row of xn has reached every rank
}

// ---------------------------------------------------------------------------------------------
// Vocabulary-sharded lm_head: every rank has sampled inside ITS slice of the vocabulary and left, per
// sequence, the pair (winning value, global id) in its own part of the region (sample_kernel, pair_out).
// One block per rank: meet the other ranks (their pairs are written), read the `world` pairs of every
// sequence straight from the peers' mappings, keep the largest value (ties: lowest id — the rule of the
// unsharded kernel, so the result is the token the unsharded kernel picks), write the token. Every rank
// computes the same answer; 8 bytes per (rank, sequence) cross NVLink instead of the rank's logits.
// The pairs are double-buffered by the caller (two consecutive exchanges never share a buffer), and at
// least one full-barrier collective of the next forward pass lies between two uses of the same buffer.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint2 ld_relaxed_sys_u32x2(const void* p) {
    uint2 v;
    asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(1024)
tp_argmax_kernel(const kb_tp_peers P, int rank, int world, size_t pairs_off, int n_seqs, uint32_t epoch,
                 int* __restrict__ out_tokens, int* __restrict__ ctx_lens_advance) {
    __threadfence_system();                       // the pairs (previous kernel on this stream) before the flag
    meet_ranks(P, rank, world, epoch);
    __syncthreads();
    for (int seq = threadIdx.x; seq < n_seqs; seq += blockDim.x) {
        float best = -INFINITY;
        uint32_t best_i = 0x7fffffffu;
#pragma unroll
        for (int r = 0; r < KB_TP_MAX_RANKS; ++r) {
            if (r < world) {
                const uint2 pr = ld_relaxed_sys_u32x2(static_cast<const char*>(P.peer_base[r]) + pairs_off + (size_t)seq * 8);
                const float v = __uint_as_float(pr.x);
                if (v > best || (v == best && pr.y < best_i)) { best = v; best_i = pr.y; }
            }
        }
        out_tokens[seq] = best_i == 0x7fffffffu ? 0 : (int)best_i;
        if (ctx_lens_advance) ctx_lens_advance[seq] += 1;
    }
}

int tp_argmax(const kb_tp_peers* P, int rank, int world, size_t pairs_off, int n_seqs, uint32_t epoch, int* out_tokens,
              int* ctx_advance, cudaStream_t stream) {
    KB_CHECK(P != nullptr && world >= 2 && world <= KB_TP_MAX_RANKS && rank >= 0 && rank < world,
             "kb_tp_argmax: bad rank %d of %d", rank, world);
    KB_CHECK(n_seqs > 0 && out_tokens != nullptr, "kb_tp_argmax: bad arguments");
    KB_CHECK(pairs_off % 8 == 0 && pairs_off + (size_t)n_seqs * 8 <= P->flags_offset,
             "kb_tp_argmax: pairs at %zu (+%zu) leave the region's payload (%zu)", pairs_off, (size_t)n_seqs * 8,
             P->flags_offset);
    for (int r = 0; r < world; ++r) KB_CHECK(P->peer_base[r] != nullptr, "kb_tp_argmax: no mapping of rank %d", r);
    const kb_tp_peers Q = with_timeout(P);
    const int threads = n_seqs < 32 ? 32 : (n_seqs > 1024 ? 1024 : ((n_seqs + 31) & ~31));
    tp_argmax_kernel<<<1, threads, 0, stream>>>(Q, rank, world, pairs_off, n_seqs, epoch, out_tokens, ctx_advance);
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

__global__ void __launch_bounds__(256)
gather_last_rows_kernel(const uint4* __restrict__ src, const int* __restrict__ cu_seqlens, uint4* __restrict__ dst,
                        int vecs) {
    const int row = cu_seqlens[blockIdx.x + 1] - 1;
    for (int j = threadIdx.x; j < vecs; j += blockDim.x) dst[(size_t)blockIdx.x * vecs + j] = src[(size_t)row * vecs + j];
}

int gather_last_rows(const void* src_f16, const int* cu_seqlens, void* dst_f16, int n_seqs, int hidden,
                     cudaStream_t stream) {
    KB_CHECK(n_seqs > 0 && hidden % 8 == 0 && cu_seqlens != nullptr, "gather_last_rows: bad arguments");
    gather_last_rows_kernel<<<n_seqs, 256, 0, stream>>>((const uint4*)src_f16, cu_seqlens, (uint4*)dst_f16, hidden / 8);
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

int tp_reduce_norm_gather(const kb_tp_peers* P, int rank, int world, size_t partial_off, size_t xn_off, float* x,
                          const void* norm_w, int T, int hidden, float eps, uint32_t epoch, int blocks,
                          cudaStream_t stream) {
    KB_CHECK(P != nullptr && world >= 2 && world <= KB_TP_MAX_RANKS && rank >= 0 && rank < world,
             "kb_tp_reduce_norm_gather: bad rank %d of %d", rank, world);
    KB_CHECK(T > 0 && hidden % 8 == 0 && hidden <= 8192, "kb_tp_reduce_norm_gather: hidden=%d must be a multiple of 8, "
             "at most 8192", hidden);
    const size_t bytes = (size_t)T * hidden * 2;
    KB_CHECK(partial_off % 16 == 0 && xn_off % 16 == 0 && partial_off + bytes <= P->flags_offset &&
                 xn_off + bytes <= P->flags_offset && (partial_off + bytes <= xn_off || xn_off + bytes <= partial_off),
             "kb_tp_reduce_norm_gather: buffers at %zu / %zu (+%zu) overlap or leave the region", partial_off, xn_off,
             bytes);
    KB_CHECK(x != nullptr && norm_w != nullptr, "kb_tp_reduce_norm_gather: null x / weight");
    const int rows_per = (T + world - 1) / world;
    if (blocks <= 0) blocks = rows_per;            // one row per block: the row's latency is paid once
    if (blocks > rows_per) blocks = rows_per;
    if (blocks > KB_TP_MAX_BLOCKS) blocks = KB_TP_MAX_BLOCKS;
    const bool mc = P->mc_base != nullptr;
    const int vpt = (hidden / 8 + 255) / 256;
    const __nv_bfloat16* w = static_cast<const __nv_bfloat16*>(norm_w);
    const kb_tp_peers Q = with_timeout(P);
#define KB_LAUNCH_RNG(MC, V)                                                                                     \
    tp_reduce_norm_gather_kernel<MC, V><<<blocks, 256, 0, stream>>>(Q, rank, world, partial_off, xn_off, x, w, T, \
                                                                     hidden, eps, epoch)
    if (vpt <= 1) { if (mc) KB_LAUNCH_RNG(true, 1); else KB_LAUNCH_RNG(false, 1); }
    else if (vpt == 2) { if (mc) KB_LAUNCH_RNG(true, 2); else KB_LAUNCH_RNG(false, 2); }
    else { if (mc) KB_LAUNCH_RNG(true, 4); else KB_LAUNCH_RNG(false, 4); }
#undef KB_LAUNCH_RNG
    count_launch();
    KB_LAUNCH_CHECK();
    return KB_OK;
}

}  // namespace kb

extern "C" int kb_allreduce_peers(const kb_tp_peers* peers, int32_t rank, int32_t world, size_t offset_bytes,
                                  size_t n, int32_t f16, uint32_t epoch, int32_t blocks, kb_stream stream) {
    return kb::allreduce_peers(peers, rank, world, offset_bytes, n, f16, epoch, blocks,
                               reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int kb_tp_argmax_pairs(const kb_tp_peers* peers, int32_t rank, int32_t world, size_t pairs_offset,
                                  int32_t n_seqs, uint32_t epoch, int32_t* out_tokens, kb_stream stream) {
    return kb::tp_argmax(peers, rank, world, pairs_offset, n_seqs, epoch, out_tokens, nullptr,
                         reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int kb_tp_reduce_norm_gather(const kb_tp_peers* peers, int32_t rank, int32_t world, size_t partial_offset,
                                        size_t xn_offset, float* x, const void* norm_w_bf16, int32_t n_rows,
                                        int32_t hidden, float eps, uint32_t epoch, int32_t blocks, kb_stream stream) {
    return kb::tp_reduce_norm_gather(peers, rank, world, partial_offset, xn_offset, x, norm_w_bf16, n_rows, hidden, eps,
                                     epoch, blocks, reinterpret_cast<cudaStream_t>(stream));
}
