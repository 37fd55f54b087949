// Micro-benchmark 2: the V-trace load pattern in isolation.  NARR fp32 arrays [R, T] (+ optionally one u8 array) are
// loaded as [32 x 32-step] 128B-swizzled TMA tiles, all tiles of a stage on ONE mbarrier, walking time forwards or
// backwards, persistent CTAs.  Which of these ingredients costs bandwidth?
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include "../../microbeast_b200/csrc/common.cuh"
namespace mb { void set_error(const char*, ...) {} int cuda_fail(cudaError_t, const char*) { return 2; } int sm_count(int) { return 148; } }
using namespace mb;

struct Maps { CUtensorMap f[5]; CUtensorMap u8; CUtensorMap o[4]; };

template <int NSTAGE>
__global__ void __launch_bounds__(64) multi_load(const __grid_constant__ Maps maps, int rows, int T, int narr, int with_u8,
                                                 int backwards, int two_warps, int pair, int nstore) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full[NSTAGE];
    __shared__ __align__(8) uint64_t done[NSTAGE];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int stage_bytes = narr * 4096 + 1024;
    if (threadIdx.x == 0) {
        for (int s = 0; s < NSTAGE; ++s) { mbar_init(&full[s], 1); mbar_init(&done[s], 1); }
        fence_mbar_init();
    }
    __syncthreads();
    const int n_groups = rows / 32, nt = T / 32;
    const long long total = (long long)((n_groups - int(blockIdx.x) + int(gridDim.x) - 1) / int(gridDim.x)) * nt;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (two_warps && warp == 0) {          // consumer warp: wait full, hand the stage back
        for (long long q = 0; q < total; ++q) {
            const int s = int(q % NSTAGE);
            mbar_wait(&full[s], uint32_t((q / NSTAGE) & 1), nullptr);
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(&done[s]);
        }
        return;
    }
    if (threadIdx.x != (two_warps ? 32 : 0)) return;
    if (nstore > 0 && !two_warps) return;
    long long lq = 0;
    auto issue = [&]() {
        const int s = int(lq % NSTAGE);
        const int g = int(lq / nt), k = int(lq % nt);
        const int kb = k / pair, ki = k % pair;        // blocks of `pair` tiles walked backwards, ascending inside a block
        const int t0 = (backwards ? ((nt / pair - 1 - kb) * pair + ki) : k) * 32;
        const int row0 = (int(blockIdx.x) + g * int(gridDim.x)) * 32;
        uint8_t* st = smem + size_t(s) * stage_bytes;
        mbar_arrive_expect_tx(&full[s], uint32_t(narr * 4096 + (with_u8 ? 1024 : 0)));
        for (int a = 0; a < narr; ++a) tma_load_2d(st + a * 4096, &maps.f[a], t0, row0, &full[s]);
        if (with_u8) tma_load_2d(st + narr * 4096, &maps.u8, t0, row0, &full[s]);
        ++lq;
    };
    for (int i = 0; i < NSTAGE && lq < total; ++i) issue();
    for (long long q = 0; q < total; ++q) {
        const int s = int(q % NSTAGE);
        if (two_warps) mbar_wait(&done[s], uint32_t((q / NSTAGE) & 1), nullptr);
        else mbar_wait(&full[s], uint32_t((q / NSTAGE) & 1), nullptr);
        if (nstore > 0) {
            const int g = int(q / nt), k = int(q % nt);
            const int kb = k / pair, ki = k % pair;
            const int t0 = (backwards ? ((nt / pair - 1 - kb) * pair + ki) : k) * 32;
            const int row0 = (int(blockIdx.x) + g * int(gridDim.x)) * 32;
            for (int a = 0; a < nstore; ++a) tma_store_2d(&maps.o[a], t0, row0, smem + size_t(s) * stage_bytes + a * 4096);
            bulk_commit();
            if (q >= 1 && lq < total) { bulk_wait_read<1>(); issue(); }
        } else if (two_warps) { if (q >= 1 && lq < total) issue(); }      // same refill rule as vtrace_tma_kernel
        else if (lq < total) issue();
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv) {
    void* fp = nullptr; cudaDriverEntryPointQueryResult qr;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &qr);
    EncodeTiledFn enc = (EncodeTiledFn)fp;
    const int T = 512;
    uint8_t *flush, *flush2; cudaMalloc(&flush, 512 << 20); cudaMalloc(&flush2, 512 << 20);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    printf("rows narr u8 backwards two_warps stages ctas_per_sm   us   GB/s\n");
    for (int R : {65536, 262144}) {
        float* arr[5]; uint8_t* u8;
        for (int a = 0; a < 5; ++a) { cudaMalloc(&arr[a], size_t(R) * T * 4); cudaMemset(arr[a], 1, size_t(R) * T * 4); }
        cudaMalloc(&u8, size_t(R) * T); cudaMemset(u8, 0, size_t(R) * T);
        float* outp[4]; for (int a = 0; a < 4; ++a) cudaMalloc(&outp[a], size_t(R) * T * 4);
        Maps maps;
        const cuuint64_t gdim[2] = {cuuint64_t(T), cuuint64_t(R)}; const cuuint32_t box[2] = {32, 32}; const cuuint32_t es[2] = {1, 1};
        for (int a = 0; a < 5; ++a) {
            const cuuint64_t gstr[1] = {cuuint64_t(T) * 4};
            enc(&maps.f[a], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, arr[a], gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        }
        { const cuuint64_t gstr[1] = {cuuint64_t(T)};
          enc(&maps.u8, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, u8, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE); }
        struct Cfg { int narr, u8, back, two, per_sm, pair, promo, nstore; };
        const Cfg cfgs[] = {{4,1,1,1,3,2,256,0},{4,1,1,1,3,2,256,2},{4,1,1,1,3,2,256,4},{4,1,0,1,3,1,256,4},{4,1,1,1,2,2,256,4},
                            {4,1,1,1,3,1,256,4},{4,1,1,1,3,4,256,4},{2,0,1,1,3,2,256,2},{4,1,1,1,1,2,256,4}};
        for (const Cfg& c : cfgs) {
            constexpr int NST = 4;
            const CUtensorMapL2promotion pr = c.promo >= 256 ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : c.promo >= 128 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B : CU_TENSOR_MAP_L2_PROMOTION_NONE;
            for (int a = 0; a < 5; ++a) {
                const cuuint64_t gstr[1] = {cuuint64_t(T) * 4};
                enc(&maps.f[a], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, arr[a], gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, pr, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                if (a < 4) enc(&maps.o[a], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, outp[a], gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, pr, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            }
            const size_t smem = size_t(NST) * (c.narr * 4096 + 1024) + 1024;
            if (smem * c.per_sm > 225 * 1024) continue;
            cudaFuncSetAttribute(multi_load<NST>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
            float best = 1e9f;
            for (int it = 0; it < 5; ++it) {
                cudaMemsetAsync(flush, it, 512 << 20);
                cudaMemcpyAsync(flush, flush2, 256 << 20, cudaMemcpyDeviceToDevice);    // leave clean lines in L2
                cudaEventRecord(e0);
                multi_load<NST><<<148 * c.per_sm, 64, smem>>>(maps, R, T, c.narr, c.u8, c.back, c.two, c.pair, c.nstore);
                cudaEventRecord(e1); cudaEventSynchronize(e1);
                float ms_; cudaEventElapsedTime(&ms_, e0, e1);
                if (it > 0 && ms_ < best) best = ms_;
            }
            cudaError_t err = cudaGetLastError();
            if (err != cudaSuccess) { printf("cuda error %s\n", cudaGetErrorString(err)); return 1; }
            const double bytes = double(R) * T * (4.0 * c.narr + (c.u8 ? 1 : 0) + 4.0 * c.nstore);
            printf("%7d %d %d %d %d %d %d pair%d promo%3d st%d %8.1f %8.1f\n", R, c.narr, c.u8, c.back, c.two, NST, c.per_sm, c.pair, c.promo, c.nstore, best * 1e3, bytes / (best * 1e-3) / 1e9);
        }
        for (int a = 0; a < 5; ++a) cudaFree(arr[a]);
        cudaFree(u8); for (int a = 0; a < 4; ++a) cudaFree(outp[a]);
    }
    return 0;
}
