// Shared device/host helpers for the microbeast_b200 kernels (sm_100a only).
//
// PTX wrappers for the Blackwell/Hopper async machinery the kernels use:
// mbarrier (full/empty pipelines), 1-D bulk copies (cp.async.bulk, SASS UBLKCP),
// 2-D tensor-map TMA (cp.async.bulk.tensor, SASS UTMALDG/UTMASTG) and the
// generic->async proxy fence.  No CUTLASS dependency.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace mb {

// ---- policy-head geometry (reference microbeast.py:145: nvec = [6,4,4,4,4,7,49]*C) ----
constexpr int kLogitsPerCell = 78;
constexpr int kCompsPerCell = 7;
constexpr float kMaskFill = -1e8f;          // reference model.py:43
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

// status bits written (atomicOr) to the optional device status word
constexpr int kStatusBadAction = 1;         // action outside [0, n_k)  (torch validate_args, Q13)
constexpr int kStatusTimeout = 2;           // an mbarrier wait gave up (pipeline bug guard)

// ---- error plumbing (host) ----
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

enum : int { MB_OK = 0, MB_ERR_INVALID = 1, MB_ERR_CUDA = 2, MB_ERR_UNSUPPORTED = 3 };

#define MB_CUDA(call)                                                     \
    do {                                                                  \
        cudaError_t _e = (call);                                          \
        if (_e != cudaSuccess) return ::mb::cuda_fail(_e, #call);         \
    } while (0)

// ---- device helpers ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}

__device__ __forceinline__ void fence_mbar_init() {
    // make the inits visible to the async proxy (TMA completes on these barriers)
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(
                     smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}

// try_wait suspends the warp in hardware until the phase completes or the time hint (ns) runs
// out, so a waiting warp does not burn issue slots that computing warps need.
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity), "r"(0x989680u)
        : "memory");
    return ok != 0;
}

__device__ __forceinline__ uint64_t global_timer_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// Bounded wait: a pipeline bug must trap instead of hanging the GPU (about 2 s of wall clock).
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity, int* status) {
    if (mbar_try_wait(bar, parity)) return true;
    uint64_t t0 = 0;
    uint32_t spins = 0;
#pragma unroll 1
    while (!mbar_try_wait(bar, parity)) {
        if ((++spins & 63u) == 0) {
            const uint64_t now = global_timer_ns();
            if (t0 == 0) t0 = now;
            if (now - t0 > 2000000000ull) {
                if (status) atomicOr(status, kStatusTimeout);
                __trap();
            }
        }
    }
    return true;
}

__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// global -> shared 1-D bulk copy, completes `bytes` on `bar`.  16-B aligned, bytes % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// shared -> global 1-D bulk copy (bulk async-group completion).
__device__ __forceinline__ void bulk_s2g(void* gmem_dst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst),
                 "r"(smem_u32(smem_src)), "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }

template <int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_all() {
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

// 2-D tiled TMA through a CUtensorMap (passed as a __grid_constant__ kernel parameter).
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const void* tmap, int c0, int c1, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::
            "r"(smem_u32(smem_dst)),
        "l"(tmap), "r"(c0), "r"(c1), "r"(smem_u32(bar))
        : "memory");
}

__device__ __forceinline__ void tma_store_2d(const void* tmap, int c0, int c1, const void* smem_src) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%1, %2}], [%3];" ::"l"(tmap),
                 "r"(c0), "r"(c1), "r"(smem_u32(smem_src))
                 : "memory");
}

__device__ __forceinline__ void tma_prefetch_desc(const void* tmap) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float lg2_approx(float x) {
    float y;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

int sm_count(int device);

}  // namespace mb
