// Micro-benchmark: how fast can TMA 2-D tile loads / stores move [32 rows x W floats] boxes of a [R, T] fp32 array,
// as a function of the box's inner extent W (bytes per row request) and swizzle?  (profiles/r02_tma_rowcost.txt)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_rowcost tma_rowcost.cu -lcudart_static -ldl -lrt -lpthread
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include "../../microbeast_b200/csrc/common.cuh"
namespace mb { void set_error(const char*, ...) {} int cuda_fail(cudaError_t, const char*) { return 2; } int sm_count(int) { return 148; } }
using namespace mb;

template <int NSTAGE>
__global__ void __launch_bounds__(64) tile_copy(const __grid_constant__ CUtensorMap src, const __grid_constant__ CUtensorMap dst,
                                                int rows, int T, int W, int box_rows, int do_store) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full[NSTAGE];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int stage_bytes = box_rows * W * 4;
    if (threadIdx.x == 0) {
        for (int s = 0; s < NSTAGE; ++s) mbar_init(&full[s], 1);
        fence_mbar_init();
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    const int n_groups = rows / box_rows, nt = T / W;
    const long long total = (long long)((n_groups - int(blockIdx.x) + int(gridDim.x) - 1) / int(gridDim.x)) * nt;
    long long lq = 0;
    auto issue = [&]() {
        const int s = int(lq % NSTAGE);
        const int g = int(lq / nt), k = int(lq % nt);
        mbar_arrive_expect_tx(&full[s], uint32_t(stage_bytes));
        tma_load_2d(smem + size_t(s) * stage_bytes, &src, k * W, (int(blockIdx.x) + g * int(gridDim.x)) * box_rows, &full[s]);
        ++lq;
    };
    for (int i = 0; i < NSTAGE && lq < total; ++i) issue();
    for (long long q = 0; q < total; ++q) {
        const int s = int(q % NSTAGE);
        mbar_wait(&full[s], uint32_t((q / NSTAGE) & 1), nullptr);
        if (do_store) {
            const int g = int(q / nt), k = int(q % nt);
            tma_store_2d(&dst, k * W, (int(blockIdx.x) + g * int(gridDim.x)) * box_rows, smem + size_t(s) * stage_bytes);
            bulk_commit();
            bulk_wait_read<0>();
        }
        if (lq < total) issue();
    }
    bulk_wait_all<0>();
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    void* fp = nullptr; cudaDriverEntryPointQueryResult qr;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &qr);
    EncodeTiledFn enc = (EncodeTiledFn)fp;
    const int R = 65536 * 8, T = 512;
    float *a, *b; uint8_t* flush;
    cudaMalloc(&a, size_t(R) * T * 4); cudaMalloc(&b, size_t(R) * T * 4); cudaMalloc(&flush, 512 << 20);
    cudaMemset(a, 1, size_t(R) * T * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    printf("W_floats box_rows swizzle stages ctas_per_sm store  us  GB/s  clk_per_row_request_per_SM\n");
    for (int do_store = 0; do_store < 2; ++do_store)
    for (int cfg = 0; cfg < 9; ++cfg) {
        // W floats per row request, rows per box, swizzle
        const int Ws[9] = {32, 32, 64, 128, 256, 16, 64, 128, 8};
        const int BRs[9] = {32, 32, 32, 32, 16, 32, 16, 8, 32};
        const CUtensorMapSwizzle SWs[9] = {CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_SWIZZLE_64B,
                                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_SWIZZLE_32B};
        const int W = Ws[cfg], BR = BRs[cfg];
        CUtensorMap ms, md;
        const cuuint64_t gdim[2] = {cuuint64_t(T), cuuint64_t(R)}; const cuuint64_t gstr[1] = {cuuint64_t(T) * 4};
        const cuuint32_t box[2] = {cuuint32_t(W), cuuint32_t(BR)}; const cuuint32_t es[2] = {1, 1};
        if (enc(&ms, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, a, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, SWs[cfg],
                CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS ||
            enc(&md, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, b, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, SWs[cfg],
                CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
            printf("encode failed for cfg %d\n", cfg); continue;
        }
        const int stage = BR * W * 4;
        for (int per_sm = 2; per_sm <= 6; per_sm += 2) {
            constexpr int NST = 4;
            const size_t smem = size_t(NST) * stage + 1024;
            if (smem * per_sm > 220 * 1024) continue;
            cudaFuncSetAttribute(tile_copy<NST>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
            float best = 1e9f;
            for (int it = 0; it < 5; ++it) {
                cudaMemsetAsync(flush, it, 512 << 20);
                cudaEventRecord(e0);
                tile_copy<NST><<<148 * per_sm, 64, smem>>>(ms, md, R, T, W, BR, do_store);
                cudaEventRecord(e1); cudaEventSynchronize(e1);
                float ms_; cudaEventElapsedTime(&ms_, e0, e1);
                if (it > 0 && ms_ < best) best = ms_;
            }
            cudaError_t err = cudaGetLastError();
            if (err != cudaSuccess) { printf("cuda error %s\n", cudaGetErrorString(err)); return 1; }
            const double bytes = double(R) * T * 4 * (do_store ? 2 : 1);
            const double row_reqs = double(R) * T / W * (do_store ? 2 : 1);
            printf("%4d %4d %2d %d %d %d  %8.1f %8.1f  %6.2f\n", W, BR, int(SWs[cfg]), NST, per_sm, do_store, best * 1e3,
                   bytes / (best * 1e-3) / 1e9, best * 1e-3 * 1.965e9 * 148 / row_reqs);
        }
    }
    return 0;
}
