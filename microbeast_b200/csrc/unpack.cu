// Wire-format expansion of the rollout store: the store ships action masks as bits
// (78*C bits per frame, bit i of byte j = entry 8j+i) and actions as uint8; the policy-head kernels
// take the reference's dtypes (u8 mask, i64 action: libs/utils.py:43-44), so one streaming kernel
// widens both on the device.  9.75 + 7 B read and 78 + 56 B written per cell: HBM-bound and small
// next to the 446 B/cell the head forward reads; the point is the 8x smaller host->device copy.
#include "common.cuh"

#include "../../include/mbeast.h"

namespace mb {

__global__ void __launch_bounds__(256) unpack_mask_kernel(const uint8_t* __restrict__ bits, uint8_t* __restrict__ mask,
                                                          long long n_frames, int width, int bytes_per_frame) {
    const long long total = n_frames * bytes_per_frame;
    const bool vec = (width % 8 == 0) && ((reinterpret_cast<uintptr_t>(mask) & 7) == 0);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long f = i / bytes_per_frame;
        const int j = int(i - f * bytes_per_frame);
        const uint32_t b = bits[i];
        uint8_t* out = mask + f * width + 8 * j;
        if (vec) {
            uint64_t v = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) v |= uint64_t((b >> k) & 1u) << (8 * k);
            *reinterpret_cast<uint64_t*>(out) = v;
        } else {
            const int cnt = min(8, width - 8 * j);
            for (int k = 0; k < cnt; ++k) out[k] = uint8_t((b >> k) & 1u);
        }
    }
}

__global__ void __launch_bounds__(256) widen_action_kernel(const uint8_t* __restrict__ a8, int64_t* __restrict__ a64,
                                                           long long total) {
    const bool vec = ((reinterpret_cast<uintptr_t>(a8) & 7) == 0) && ((reinterpret_cast<uintptr_t>(a64) & 15) == 0);
    const long long groups = vec ? total / 8 : 0;
    const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x, stride = (long long)gridDim.x * blockDim.x;
    for (long long g = tid; g < groups; g += stride) {
        const uint64_t v = *reinterpret_cast<const uint64_t*>(a8 + 8 * g);
        longlong2* out = reinterpret_cast<longlong2*>(a64 + 8 * g);
#pragma unroll
        for (int k = 0; k < 4; ++k)
            out[k] = make_longlong2((long long)((v >> (16 * k)) & 0xffu), (long long)((v >> (16 * k + 8)) & 0xffu));
    }
    for (long long i = groups * 8 + tid; i < total; i += stride) a64[i] = a8[i];
}

// obs: the 27 feature planes of a cell are 0/1 (five one-hot groups), so a cell travels as one
// uint32 (bit c = plane c) and is expanded straight into the NCHW fp32 tensor the cuDNN
// convolutions consume -- replacing the uint8->float cast and the NHWC->NCHW copy on the device.
__global__ void __launch_bounds__(256) unpack_obs_kernel(const uint32_t* __restrict__ bits, float* __restrict__ out,
                                                         long long n_frames, int cells, int planes) {
    const long long total = n_frames * cells;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long f = i / cells;
        const int c = int(i - f * cells);
        const uint32_t b = bits[i];
        float* o = out + f * (long long)planes * cells + c;
        for (int k = 0; k < planes; ++k) o[(long long)k * cells] = float((b >> k) & 1u);
    }
}

// Which cells have any valid action (the cell list of Agent.sparse_head): one byte out per 78 in.
// A warp stages 32 cells (2496 B = 156 x 16 B, coalesced) in shared memory, then every lane ORs
// its own cell's 39 half-words.  78 B read + 1 B written per cell.
__global__ void __launch_bounds__(256) active_cells_kernel(const uint8_t* __restrict__ mask, uint8_t* __restrict__ active,
                                                           long long n_cells) {
    constexpr int kGroupBytes = 32 * kLogitsPerCell, kVecs = kGroupBytes / 16;
    static_assert(kGroupBytes % 16 == 0 && kLogitsPerCell % 2 == 0, "a 32-cell group is whole 16-byte vectors");
    __shared__ __align__(16) uint8_t stage[8][kGroupBytes];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const long long warp = blockIdx.x * 8ll + wib, n_warps = gridDim.x * 8ll;
    const bool vec = (reinterpret_cast<uintptr_t>(mask) & 15) == 0;
    const long long groups = vec ? n_cells / 32 : 0;
    for (long long g = warp; g < groups; g += n_warps) {
        const uint4* src = reinterpret_cast<const uint4*>(mask + g * kGroupBytes);
        uint4* dst = reinterpret_cast<uint4*>(stage[wib]);
#pragma unroll
        for (int k = 0; k < (kVecs + 31) / 32; ++k)
            if (lane + 32 * k < kVecs) dst[lane + 32 * k] = __ldg(src + lane + 32 * k);
        __syncwarp();
        const uint16_t* mine = reinterpret_cast<const uint16_t*>(stage[wib] + lane * kLogitsPerCell);
        uint32_t any = 0;
#pragma unroll
        for (int k = 0; k < kLogitsPerCell / 2; ++k) any |= mine[k];
        active[g * 32 + lane] = any ? 1 : 0;
        __syncwarp();
    }
    for (long long c = groups * 32 + warp * 32 + lane; c < n_cells; c += n_warps * 32) {
        uint32_t any = 0;
        for (int k = 0; k < kLogitsPerCell; ++k) any |= mask[c * kLogitsPerCell + k];
        active[c] = any ? 1 : 0;
    }
}

}  // namespace mb

extern "C" int mb_unpack_obs(const uint32_t* obs_bits, float* obs_out, int64_t n_frames, int32_t cells,
                             int32_t planes, void* stream) {
    if (n_frames < 0 || cells <= 0 || planes <= 0 || planes > 32) {
        mb::set_error("mb_unpack_obs: n_frames=%lld cells=%d planes=%d out of range (planes <= 32)",
                      (long long)n_frames, cells, planes);
        return MB_ERR_INVALID;
    }
    if (n_frames == 0) return MB_OK;
    if (!obs_bits || !obs_out) {
        mb::set_error("mb_unpack_obs: null pointer");
        return MB_ERR_INVALID;
    }
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    mb::unpack_obs_kernel<<<mb::sm_count(dev) * 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        obs_bits, obs_out, n_frames, cells, planes);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}

extern "C" int mb_unpack_rollout(const uint8_t* mask_bits, const uint8_t* action_u8, uint8_t* mask_out,
                                 int64_t* action_out, int64_t n_frames, int32_t cells, void* stream) {
    if (n_frames < 0 || cells <= 0) {
        mb::set_error("mb_unpack_rollout: n_frames=%lld cells=%d out of range", (long long)n_frames, cells);
        return MB_ERR_INVALID;
    }
    if ((mask_bits == nullptr) != (mask_out == nullptr) || (action_u8 == nullptr) != (action_out == nullptr)) {
        mb::set_error("mb_unpack_rollout: give both pointers of a pair (packed input and output) or neither");
        return MB_ERR_INVALID;
    }
    if (n_frames == 0) return MB_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    const int grid = mb::sm_count(dev) * 8;
    if (mask_bits) {
        const int width = cells * mb::kLogitsPerCell;
        mb::unpack_mask_kernel<<<grid, 256, 0, st>>>(mask_bits, mask_out, n_frames, width, (width + 7) / 8);
        MB_CUDA(cudaGetLastError());
    }
    if (action_u8) {
        mb::widen_action_kernel<<<grid, 256, 0, st>>>(action_u8, action_out,
                                                      (long long)n_frames * cells * mb::kCompsPerCell);
        MB_CUDA(cudaGetLastError());
    }
    return MB_OK;
}

extern "C" int mb_active_cells(const uint8_t* mask, uint8_t* active, int64_t n_cells, void* stream) {
    if (n_cells < 0) {
        mb::set_error("mb_active_cells: n_cells=%lld out of range", (long long)n_cells);
        return MB_ERR_INVALID;
    }
    if (n_cells == 0) return MB_OK;
    if (!mask || !active) {
        mb::set_error("mb_active_cells: null pointer");
        return MB_ERR_INVALID;
    }
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    mb::active_cells_kernel<<<mb::sm_count(dev) * 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(mask, active,
                                                                                                 n_cells);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}
