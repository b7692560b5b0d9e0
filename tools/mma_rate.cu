// mma_rate.cu — how fast does ONE thread get tcgen05.mma instructions of the shapes the attention and GEMM kernels
// use through the tensor pipe of one SM?  (Measurement aid, not part of the library.)
//
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I kafka-llm-service_b200/csrc tools/mma_rate.cu \
//        -o tools/_bin/mma_rate -lcudart   &&   tools/_bin/mma_rate
//
// One CTA per SM (so the chip's power state is that of a busy GPU), thread 0 of each issues `n` MMAs (M = 128,
// cta_group::1, K = 16, kind::f16) into tensor memory, commits them to an mbarrier and waits; cycles = clock64
// around issue + wait, reported per MMA for chains of 8 / 64 / 512 instructions (the slope is the pipe's rate, the
// intercept the commit / barrier round trip). Operand values are irrelevant (shared memory is zero).
#include "kb_common.cuh"

using namespace kb;

__device__ __forceinline__ uint64_t desc_k(uint32_t saddr) {       // K-major, SWIZZLE_128B
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) |
           ((uint64_t)2 << 61);
}
__device__ __forceinline__ uint64_t desc_mn(uint32_t saddr, uint32_t lbo, uint32_t sbo) {   // MN-major, SWIZZLE_128B
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) |
           ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
__device__ __forceinline__ uint32_t idesc(int n, int b_mn) {
    return (1u << 4) | ((uint32_t)b_mn << 16) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t id, uint32_t acc) {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(d),
                 "l"(a), "l"(b), "r"(id), "r"(acc) : "memory");
}
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t id, uint32_t acc) {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n}\n" ::"r"(d),
                 "r"(a), "l"(b), "r"(id), "r"(acc) : "memory");
}

// mode: 0 = SS (A, B K-major in shared memory), 1 = TS (A in tensor memory, B K-major), 2 = TS with B MN-major,
//       3 = SS alternating between two accumulators, 4 = the attention tile: 8 x SS N=64 then 4 x TS N=128 MN-major
struct Case { int mode, n_cols, chain; };

__global__ void __launch_bounds__(128, 1) mma_rate_kernel(const Case* cases, int n_cases, long long* out, int reps) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* a_s = smem;                       // 128 x 64 fp16, 16 KB
    uint8_t* b_s = smem + 16384;               // up to 256 x 64 fp16, 32 KB (or a 64 x 128 V tile)
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384 + 32768);
    uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 4);
    for (int i = threadIdx.x; i < (16384 + 32768) / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
    if (threadIdx.x == 0) { mbar_init(bar, 1); mbar_init(bar + 2, 1); slot[1] = 0; slot[2] = 0; fence_mbar_init(); }
    if (threadIdx.x < 32) { tmem_alloc(slot, 512); tmem_relinquish(); }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = *slot;
    uint64_t* bar2 = bar + 2;
    if (threadIdx.x == 32) {          // second issuing thread (mode 8 only): its own accumulator and barrier
        uint32_t phase2 = 0;
        const uint64_t ad = desc_k(smem_u32(a_s)), bk = desc_k(smem_u32(b_s));
        for (int c = 0; c < n_cases; ++c) {
            const Case cs = cases[c];
            if (cs.mode != 8) continue;
            const uint32_t id = idesc(cs.n_cols, 0);
            for (int r = 0; r < reps; ++r) {
                while (*reinterpret_cast<volatile int*>(slot + 1) != c * 1000 + r + 1) { }      // start together
                for (int i = 0; i < cs.chain; i += 8) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) mma_ss(tm + 256, ad + (uint64_t)((k & 3) * 2), bk + (uint64_t)((k & 3) * 2), id, 1);
                }
                umma_commit(bar2);
                mbar_wait(bar2, phase2);
                phase2 ^= 1;
                *reinterpret_cast<volatile int*>(slot + 2) = c * 1000 + r + 1;                    // done
            }
        }
    }
    if (threadIdx.x == 0) {
        uint32_t phase = 0;
        const uint64_t ad = desc_k(smem_u32(a_s)), bk = desc_k(smem_u32(b_s)), bm = desc_mn(smem_u32(b_s), 64 * 128, 1024);
        for (int c = 0; c < n_cases; ++c) {
            const Case cs = cases[c];
            long long best = 1LL << 60;
            for (int r = 0; r < reps; ++r) {
                const long long t0 = clock64();
                if (cs.mode == 4) {
                    for (int i = 0; i < cs.chain; ++i) {          // one attention tile: S = Q K^T (8), O += P V (4)
                        for (int k = 0; k < 8; ++k)
                            mma_ss(tm, ad + (uint64_t)((k & 3) * 2), bk + (uint64_t)((k & 3) * 2), idesc(64, 0), k != 0);
                        for (int k = 0; k < 4; ++k)
                            mma_ts(tm + 128, tm + 384 + k * 8, bm + (uint64_t)((k * 2048) >> 4), idesc(128, 1), 1);
                    }
                } else if (cs.mode == 5) {                        // chains of 8 SS MMAs, the first of each chain overwrites
                    const uint32_t id = idesc(cs.n_cols, 0);
                    for (int i = 0; i < cs.chain; ++i)
                        mma_ss(tm, ad + (uint64_t)((i & 3) * 2), bk + (uint64_t)((i & 3) * 2), id, (i & 7) != 0);
                } else if (cs.mode == 6) {                        // SS, always overwrite (no dependence on the old D)
                    const uint32_t id = idesc(cs.n_cols, 0);
                    for (int i = 0; i < cs.chain; ++i)
                        mma_ss(tm, ad + (uint64_t)((i & 3) * 2), bk + (uint64_t)((i & 3) * 2), id, 0);
                } else if (cs.mode == 8) {                        // like mode 7, while thread 32 issues the same chain too
                    const uint32_t id = idesc(cs.n_cols, 0);
                    *reinterpret_cast<volatile int*>(slot + 1) = c * 1000 + r + 1;
                    for (int i = 0; i < cs.chain; i += 8) {
#pragma unroll
                        for (int k = 0; k < 8; ++k) mma_ss(tm, ad + (uint64_t)((k & 3) * 2), bk + (uint64_t)((k & 3) * 2), id, 1);
                    }
                } else if (cs.mode == 7) {                        // SS accumulate, fully unrolled by 8 (no loop overhead)
                    const uint32_t id = idesc(cs.n_cols, 0);
                    for (int i = 0; i < cs.chain; i += 8) {
#pragma unroll
                        for (int k = 0; k < 8; ++k) mma_ss(tm, ad + (uint64_t)((k & 3) * 2), bk + (uint64_t)((k & 3) * 2), id, 1);
                    }
                } else {
                    const uint32_t id = idesc(cs.n_cols, cs.mode == 2);
                    for (int i = 0; i < cs.chain; ++i) {
                        const uint64_t ko = (uint64_t)((i & 3) * 2);
                        if (cs.mode == 0) mma_ss(tm, ad + ko, bk + ko, id, 1);
                        else if (cs.mode == 1) mma_ts(tm, tm + 384 + (i & 3) * 8, bk + ko, id, 1);
                        else if (cs.mode == 2) mma_ts(tm, tm + 384 + (i & 3) * 8, bm + (uint64_t)(((i & 3) * 2048) >> 4), id, 1);
                        else mma_ss(tm + (i & 1) * 256, ad + ko, bk + ko, id, 1);
                    }
                }
                umma_commit(bar);
                mbar_wait(bar, phase);
                phase ^= 1;
                if (cs.mode == 8) while (*reinterpret_cast<volatile int*>(slot + 2) != c * 1000 + r + 1) { }
                const long long dt = clock64() - t0;
                best = dt < best ? dt : best;
            }
            if (blockIdx.x == 0) out[c] = best;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

int main() {
    const char* names[] = {"SS  (A, B K-major in smem)", "TS  (A in TMEM, B K-major)", "TS  (A in TMEM, B MN-major)",
                           "SS  two accumulators alternating", "attention tile: 8 x SS N=64 + 4 x TS N=128 MN-major",
                           "SS  chains of 8, first overwrites", "SS  always overwrite", "SS  accumulate, unrolled x8",
                           "SS  TWO threads, a chain each (per thread)"};
    Case h[96];
    int n = 0;
    const int chains[3] = {8, 64, 512};
    for (int mode = 0; mode < 4; ++mode)
        for (int nc : {64, 128, 256}) {
            if (mode == 2 && nc != 128) continue;
            for (int ch : chains) h[n++] = {mode, nc, ch};
        }
    for (int ch : {1, 8, 64}) h[n++] = {4, 0, ch};
    for (int mode = 5; mode < 9; ++mode)
        for (int nc : {64, 128, 256}) h[n++] = {mode, nc, 512};
    Case* d;
    long long* out;
    cudaMalloc(&d, sizeof(h));
    cudaMalloc(&out, 96 * sizeof(long long));
    cudaMemcpy(d, h, sizeof(h), cudaMemcpyHostToDevice);
    const int smem = 16384 + 32768 + 64 + 1024;
    cudaFuncSetAttribute(mma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    mma_rate_kernel<<<sms, 128, smem>>>(d, n, out, 20);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
    long long r[96];
    cudaMemcpy(r, out, sizeof(r), cudaMemcpyDeviceToHost);
    printf("M = 128, cta_group::1, K = 16 per instruction, kind::f16; best of 20; cycles for the whole chain incl. commit + wait\n");
    for (int i = 0; i < n; ++i) {
        if (h[i].mode == 4)
            printf("%-52s tiles %4d : %8lld cycles = %7.1f per tile (12 MMAs, floor 512)\n", names[4], h[i].chain, r[i],
                   (double)r[i] / h[i].chain);
        else
            printf("%-36s N %3d chain %4d : %8lld cycles = %7.1f per MMA (floor N/2 = %d)\n", names[h[i].mode], h[i].n_cols,
                   h[i].chain, r[i], (double)r[i] / h[i].chain, h[i].n_cols / 2);
    }
    return 0;
}
