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
// Flags are monotonically increasing epochs (two per call), one 4-byte slot per (block, source rank) in
// each rank's flag area; nothing is ever reset, so a fast rank entering the next call cannot confuse a
// slow one. The grid is at most KB_TP_MAX_BLOCKS blocks, far below the SM count, so all blocks of all
// ranks are co-resident — the condition for a spinning barrier. A rank that waits longer than ~2 s
// traps (the process fails loudly instead of hanging the box).
//
// Reference counterpart: none (the reference's models sit behind a gateway, src/llm/portkey.py:379).
#include "kb_common.cuh"

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

// Block `blockIdx.x` of this rank meets block `blockIdx.x` of every other rank.
__device__ __forceinline__ void rank_barrier(const kb_tp_peers& P, int rank, int world, uint32_t epoch) {
    if (threadIdx.x < (unsigned)world) {
        const int peer = threadIdx.x;
        uint32_t* theirs = reinterpret_cast<uint32_t*>(static_cast<char*>(P.peer_base[peer]) + P.flags_offset) +
                           blockIdx.x * KB_TP_MAX_RANKS + rank;
        const uint32_t* mine = reinterpret_cast<const uint32_t*>(static_cast<char*>(P.peer_base[rank]) +
                                                                 P.flags_offset) +
                               blockIdx.x * KB_TP_MAX_RANKS + peer;
        st_release_sys(theirs, epoch);
        long long t0 = clock64();
        while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
            if (clock64() - t0 > 4000000000LL) {   // ~2 s at 2 GHz: a peer never arrived
                printf("kb_allreduce: rank %d block %d timed out waiting for rank %d (epoch %u)\n", rank,
                       (int)blockIdx.x, peer, epoch);
                __trap();
            }
        }
    }
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
    __syncthreads();
    rank_barrier(P, rank, world, epoch);          // every rank's partial sums are in place
    __syncthreads();

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

    __syncthreads();
    rank_barrier(P, rank, world, epoch + 1);      // every slice has been written to every rank
    __syncthreads();
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
    if (blocks <= 0) blocks = 32;
    if (blocks > KB_TP_MAX_BLOCKS) blocks = KB_TP_MAX_BLOCKS;
    const size_t n_vec = bytes / 16;
    const bool mc = P->mc_base != nullptr;
    if (mc && f16)
        allreduce_kernel<true, true><<<blocks, 512, 0, stream>>>(*P, rank, world, offset_bytes, n_vec, epoch);
    else if (mc)
        allreduce_kernel<true, false><<<blocks, 512, 0, stream>>>(*P, rank, world, offset_bytes, n_vec, epoch);
    else if (f16)
        allreduce_kernel<false, true><<<blocks, 512, 0, stream>>>(*P, rank, world, offset_bytes, n_vec, epoch);
    else
        allreduce_kernel<false, false><<<blocks, 512, 0, stream>>>(*P, rank, world, offset_bytes, n_vec, epoch);
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
