// How fast is the legacy warp-level tensor-core path (mma.sync.m16n8k8 tf32) on B200?  (profiles/r02_micro_mma_sync.txt)
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>
__global__ void __launch_bounds__(256) mma_loop(float* out, int iters) {
    uint32_t a[4] = {0x3f800000u, 0x3f800000u, 0x3f800000u, 0x3f800000u}, b[2] = {0x3f800000u, 0x3f800000u};
    float c[8][4];
    for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) c[i][j] = 0.f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
            asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                         : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
    }
    float s = 0.f;
    for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) s += c[i][j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void __launch_bounds__(256) ffma_loop(float* out, int iters) {
    float c[32]; for (int i = 0; i < 32; ++i) c[i] = float(i);
    float a = 1.0001f, b = 0.5f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 32; ++i) c[i] = fmaf(c[i], a, b);
    }
    float s = 0.f; for (int i = 0; i < 32; ++i) s += c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
    float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int blocks_per_sm = 1; blocks_per_sm <= 8; blocks_per_sm *= 2) {
        const int iters = 20000;
        mma_loop<<<148 * blocks_per_sm, 256>>>(out, 100);
        cudaEventRecord(e0); mma_loop<<<148 * blocks_per_sm, 256>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double flop = double(148 * blocks_per_sm) * 8 /*warps*/ * iters * 8 * (16.0 * 8 * 8 * 2);
        printf("mma.sync m16n8k8 tf32: %d CTAs/SM x 8 warps: %.1f TFLOP/s\n", blocks_per_sm, flop / (ms * 1e-3) / 1e12);
        ffma_loop<<<148 * blocks_per_sm, 256>>>(out, 100);
        cudaEventRecord(e0); ffma_loop<<<148 * blocks_per_sm, 256>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        const double f2 = double(148 * blocks_per_sm) * 256 * iters * 32 * 2.0;
        printf("fp32 FFMA:             %d CTAs/SM x 8 warps: %.1f TFLOP/s\n", blocks_per_sm, f2 / (ms * 1e-3) / 1e12);
    }
    return 0;
}
