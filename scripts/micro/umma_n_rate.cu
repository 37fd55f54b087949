// How does the issue rate of tcgen05.mma.kind::tf32 (M = 128, K = 8 per instruction) depend on N?
// The encoder's convolutions are GEMMs with N = 16 or 32 output channels; the policy head runs N = 240.  Operands sit
// in shared memory (zeros: rate does not depend on the values), one CTA per SM, one thread issues 4 k-steps x ITERS
// back to back into one TMEM accumulator, then commits and waits.  (profiles/r02_micro_umma_n.txt)
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -I microbeast_b200/csrc scripts/micro/umma_n_rate.cu -o gpurun_out/umma_n_rate
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>

#include "tc.cuh"

using namespace mb;

template <int N>
__global__ void __launch_bounds__(128) umma_rate(long long* cycles, int iters) {
    extern __shared__ __align__(1024) unsigned char sm[];
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    unsigned char* base = (unsigned char*)((uintptr_t(sm) + 1023) & ~uintptr_t(1023));
    float* a = (float*)base;                        // 128 rows x 32 fp32, SW128 K-major
    float* b = a + 128 * 32;                        // 256 rows x 32 fp32
    for (int i = threadIdx.x; i < (128 + 256) * 32; i += 128) a[i] = 0.f;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
    }
    __syncthreads();
    if (threadIdx.x < 32) tmem_alloc(&slot, 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = slot;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    long long t0 = 0, t1 = 0;
    if (threadIdx.x == 0) {
        const uint32_t idesc = tf32_idesc(128, N);
        const uint64_t da = make_kmajor_sw128_desc(smem_u32(a)), db = make_kmajor_sw128_desc(smem_u32(b));
        t0 = clock64();
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int k = 0; k < 4; ++k) mma_tf32(tm, da + (k * 32 >> 4), db + (k * 32 >> 4), idesc, 1u);
        }
        tc_commit(&bar);
        mbar_wait(&bar, 0, nullptr);
        t1 = clock64();
        cycles[blockIdx.x] = t1 - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc(tm, 256);
}

template <int N>
void run(long long* d_cycles, int sms, int iters) {
    const int smem = (128 + 256) * 128 + 1024;
    cudaFuncSetAttribute(umma_rate<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    umma_rate<N><<<sms, 128, smem>>>(d_cycles, 64);                      // warm
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    cudaEventRecord(e0);
    umma_rate<N><<<sms, 128, smem>>>(d_cycles, iters);
    cudaEventRecord(e1);
    cudaError_t err = cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    long long h[256];
    cudaMemcpy(h, d_cycles, sms * sizeof(long long), cudaMemcpyDeviceToHost);
    const double per = double(h[0]) / (4.0 * iters);
    const double flop = 2.0 * 128 * N * 8 * 4.0 * iters * sms;
    printf("N=%3d  %-8s clk/instr %6.1f  (ideal at 8192 flop/clk/SM: %5.1f)  %7.1f TFLOP/s over %d SMs  A+B smem bytes/instr %d\n", N,
           cudaGetErrorString(err), per, 2.0 * 128 * N * 8 / 8192.0, flop / (ms * 1e-3) / 1e12, sms, 128 * 32 + N * 32);
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    long long* d;
    cudaMalloc(&d, 256 * sizeof(long long));
    const int iters = 20000;
    run<16>(d, sms, iters);
    run<32>(d, sms, iters);
    run<64>(d, sms, iters);
    run<128>(d, sms, iters);
    run<240>(d, sms, iters);
    run<256>(d, sms, iters);
    return 0;
}
