// tma_ingress.cu — how many bytes per clock does an SM take in through TMA when every SM of the chip streams
// 16 KB operand tiles out of an L2-resident matrix, and does sharing a tile between the CTAs of a cluster
// (the same tile requested twice, or one request multicast to the cluster) change that?
// (Measurement aid, not part of the library. The matrix is this process's own cudaMalloc memory on one GPU.)
//
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I kafka-llm-service_b200/csrc tools/tma_ingress.cu \
//        -o tools/_bin/tma_ingress -lcudart   &&   tools/_bin/tma_ingress
//
// Background: at M = 512 the projections' main loop takes ~370 cycles per 64-wide k-block on 256 x 112 tiles where
// the MMAs need ~224 (profiles/r02_gemm_splitk.md): every CTA stages 16 KB of A per k-block, and ~37 CTAs stage the
// SAME A tile. This probe isolates the staging: one producer thread per CTA keeps a ring of 8 x 16 KB slots full,
// one consumer thread frees a slot as soon as it has landed; no MMA, no epilogue.
//
// Modes (tile = 128 rows x 64 fp16 columns, SWIZZLE_128B, out of a 512 x 4096 matrix = 4 MB, L2 resident):
//   0  every CTA walks its own k sequence (CTA i starts at k-block 7 i): tiles requested by different SMs at the
//      same time are mostly different
//   1  like the GEMM: all CTAs with the same row block request the same tile at about the same time
//   2  cluster of 2, both CTAs request the same tile (two unicast loads of the whole tile)
//   3  cluster of 2, each CTA requests HALF of the tile (64 rows) and multicasts it to both
//   4  cluster of 4, each CTA requests a QUARTER (32 rows) and multicasts it to all four
// Reported: bytes landed in shared memory per clock per SM, and chip-wide.
#include "kb_common.cuh"

#include <cuda.h>
#include <cudaTypedefs.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

using namespace kb;

constexpr int STAGES = 8;
constexpr int TILE_BYTES = 128 * 64 * 2;

__device__ __forceinline__ void tma_load_2d_mc(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar,
                                               uint16_t mask) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, "
        "%4}], [%2], %5;" ::"r"(smem_u32(smem_dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask)
        : "memory");
}

// tm_full: box 128 rows; tm_half: box 64 rows; tm_quarter: box 32 rows
__global__ void __launch_bounds__(64, 1)
ingress_kernel(const __grid_constant__ CUtensorMap tm_full, const __grid_constant__ CUtensorMap tm_half,
               const __grid_constant__ CUtensorMap tm_quarter, int mode, int csize, int iters, int n_kblocks,
               long long* cycles) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * TILE_BYTES);
    uint64_t* empty = full + STAGES;
    const uint32_t rank = csize > 1 ? cluster_ctarank() : 0;
    const int cluster_id = blockIdx.x / csize;
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(full + s, 1);
            mbar_init(empty + s, mode >= 3 ? csize : 1);    // multicast: every CTA of the cluster must have freed it
        }
        fence_mbar_init();
    }
    __syncthreads();
    if (csize > 1) cluster_sync_all();
    // which tile this CTA (or its cluster) wants at iteration `it`
    const int unit = (mode >= 2) ? cluster_id : (int)blockIdx.x;
    const int row_block = unit & 3;
    const int k0 = (mode == 0) ? unit * 7 : 0;
    long long t0 = 0;
    if (threadIdx.x == 0) {                 // producer
        t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            const int s = it % STAGES;
            if (it >= STAGES) mbar_wait(empty + s, ((it / STAGES) - 1) & 1);
            const int kb = (k0 + it) % n_kblocks;
            uint8_t* dst = smem + s * TILE_BYTES;
            mbar_arrive_expect_tx(full + s, TILE_BYTES);
            if (mode <= 2) {
                tma_load_2d(dst, &tm_full, kb * 64, row_block * 128, full + s);
            } else if (mode == 3) {
                tma_load_2d_mc(dst + rank * (TILE_BYTES / 2), &tm_half, kb * 64, row_block * 128 + rank * 64, full + s,
                               (uint16_t)0x3);
            } else {
                tma_load_2d_mc(dst + rank * (TILE_BYTES / 4), &tm_quarter, kb * 64, row_block * 128 + rank * 32,
                               full + s, (uint16_t)0xF);
            }
        }
    } else if (threadIdx.x == 32) {         // consumer: hand the slot back as soon as its bytes are there
        for (int it = 0; it < iters; ++it) {
            const int s = it % STAGES;
            mbar_wait(full + s, (it / STAGES) & 1);
            if (mode >= 3) {
                for (int c = 0; c < csize; ++c) mbar_arrive_cluster(empty + s, c);
            } else {
                mbar_arrive(empty + s);
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
    if (csize > 1) cluster_sync_all();      // nobody leaves while a peer may still write into its shared memory
}

static PFN_cuTensorMapEncodeTiled_v12000 get_encode() {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn) {
        fprintf(stderr, "cuTensorMapEncodeTiled not available\n");
        exit(1);
    }
    return reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
}

static CUtensorMap make_map(PFN_cuTensorMapEncodeTiled_v12000 enc, void* p, uint64_t rows, uint64_t cols, uint32_t box_rows) {
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * 2};
    cuuint32_t box[2] = {64, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUtensorMap m;
    CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, p, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { fprintf(stderr, "encode failed %d\n", (int)r); exit(1); }
    return m;
}

int main() {
    const int M = 512, K = 4096, iters = 4096;
    void* a = nullptr;
    cudaMalloc(&a, (size_t)M * K * 2);
    cudaMemset(a, 0, (size_t)M * K * 2);
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    auto enc = get_encode();
    const CUtensorMap tf = make_map(enc, a, M, K, 128), th = make_map(enc, a, M, K, 64), tq = make_map(enc, a, M, K, 32);
    long long* d_cyc = nullptr;
    cudaMalloc(&d_cyc, sizeof(long long) * 256);
    const size_t smem = STAGES * TILE_BYTES + 1024 + 256;
    cudaFuncSetAttribute(ingress_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(ingress_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    const char* names[5] = {"own k sequence per CTA", "same tile per row block (GEMM-like)", "cluster 2, same tile twice",
                            "cluster 2, halves multicast", "cluster 4, quarters multicast"};
    printf("%d SMs, tile %d B, ring of %d, %d tiles per CTA\n", sms, TILE_BYTES, STAGES, iters);
    for (int grid_kind = 0; grid_kind < 2; ++grid_kind) {
        for (int mode = 0; mode < 5; ++mode) {
            const int csize = mode < 2 ? 1 : (mode == 4 ? 4 : 2);
            int grid = grid_kind == 0 ? sms : sms / 2;
            grid -= grid % csize;
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(grid);
            cfg.blockDim = dim3(64);
            cfg.dynamicSmemBytes = smem;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = csize;
            at[0].val.clusterDim.y = 1;
            at[0].val.clusterDim.z = 1;
            cfg.attrs = at;
            cfg.numAttrs = 1;
            double best = 0, best_ms = 0;
            for (int rep = 0; rep < 3; ++rep) {
                cudaEvent_t e0, e1;
                cudaEventCreate(&e0);
                cudaEventCreate(&e1);
                cudaEventRecord(e0);
                cudaError_t err = cudaLaunchKernelEx(&cfg, ingress_kernel, tf, th, tq, mode, csize, iters, K / 64, d_cyc);
                cudaEventRecord(e1);
                if (err != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
                    fprintf(stderr, "mode %d failed: %s\n", mode, cudaGetErrorString(cudaGetLastError()));
                    return 1;
                }
                float ms = 0;
                cudaEventElapsedTime(&ms, e0, e1);
                std::vector<long long> cyc(grid);
                cudaMemcpy(cyc.data(), d_cyc, sizeof(long long) * grid, cudaMemcpyDeviceToHost);
                double mean = 0;
                for (long long c : cyc) mean += (double)c;
                mean /= grid;
                const double bpc = (double)iters * TILE_BYTES / mean;
                if (bpc > best) { best = bpc; best_ms = ms; }
            }
            printf("grid %3d  mode %d (%-36s): %6.1f B/clk per SM, %7.0f B/clk chip, %.3f ms -> %.2f TB/s landed\n", grid, mode,
                   names[mode], best, best * grid, best_ms, (double)grid * iters * TILE_BYTES / (best_ms * 1e-3) / 1e12);
        }
    }
    return 0;
}
