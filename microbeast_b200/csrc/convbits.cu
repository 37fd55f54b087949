// First encoder convolution straight from bit-packed observations, for sm_100a.
//
// Reference: model.py:88 / model.py:121 -- ConvSequence(27 -> 16).conv applied to obs.permute(0, 3, 1, 2)
// (model.py:156-160).  gym-microRTS observations are 27 binary feature planes; the rollout store ships them as one
// int32 per cell (bit c = plane c, csrc/unpack.cu).  A 3x3 convolution over 0/1 inputs is a sum of weight columns:
//     out[n, co, y, x] = bias[co] + sum over the 9 taps, over the SET bits c of cell (y+ky-1, x+kx-1): W[co, c, ky, kx]
// -- 5 set bits per cell for one-hot planes, i.e. 45 additions of a 16-vector per output pixel instead of 243 FMAs of
// one (and no [N, 27, h, w] fp32 tensor has to be written and read back for the forward pass).  Any bit pattern is
// handled (the loop runs over the set bits, whatever their number); out-of-image taps see the zero halo.
//
// One thread per output pixel, 16 accumulators (the 16 output channels).  The weights sit in shared memory as
// [tap][plane][16] rows padded to 80 bytes, so that lanes whose cells differ read different bank groups while the
// common case -- neighbouring cells with the same planes set -- is a broadcast.
#include "common.cuh"

#include "../../include/mbeast.h"

namespace mb {

int sm_count(int device);

namespace {

constexpr int kBitsCout = 16, kBitsCin = 27;
constexpr int kRowPitch = 20;                       // floats per [tap][plane] row: 16 + 4 padding

template <int H>
__global__ void __launch_bounds__(256) conv1_bits_fwd_kernel(const int* __restrict__ obs_bits, const float* __restrict__ weight,
                                                             const float* __restrict__ bias, float* __restrict__ out,
                                                             long long n_frames) {
    constexpr int W = H, HW = H * W;
    constexpr int FPC = 256 / HW > 0 ? 256 / HW : 1;                  // frames per CTA pass (8x8: 4, 16x16: 1)
    constexpr int TP = W + 2;                                          // tile pitch with halo
    __shared__ __align__(16) float wt[9 * kBitsCin * kRowPitch];
    __shared__ float bs[kBitsCout];
    __shared__ int tile[FPC][(H + 2) * TP];
    // torch layout [cout][cin][3][3] -> [tap][cin][cout]
    for (int i = threadIdx.x; i < 9 * kBitsCin * kBitsCout; i += 256) {
        const int co = i % kBitsCout, ci = (i / kBitsCout) % kBitsCin, tap = i / (kBitsCout * kBitsCin);
        wt[(tap * kBitsCin + ci) * kRowPitch + co] = weight[(co * kBitsCin + ci) * 9 + tap];
    }
    if (threadIdx.x < kBitsCout) bs[threadIdx.x] = bias ? bias[threadIdx.x] : 0.0f;
    for (int i = threadIdx.x; i < FPC * (H + 2) * TP; i += 256) (&tile[0][0])[i] = 0;      // the halo stays zero
    __syncthreads();
    const int slot = threadIdx.x / HW, pix = threadIdx.x % HW;
    const int y = pix / W, x = pix % W;
    const long long n_pass = (n_frames + FPC - 1) / FPC;
    for (long long pass = blockIdx.x; pass < n_pass; pass += gridDim.x) {
        const long long f = pass * FPC + slot;
        const bool live = slot < FPC && f < n_frames;
        if (slot < FPC) tile[slot][(y + 1) * TP + x + 1] = live ? __ldg(obs_bits + f * HW + pix) : 0;
        __syncthreads();
        if (live) {
            float acc[kBitsCout];
#pragma unroll
            for (int c = 0; c < kBitsCout; ++c) acc[c] = bs[c];
#pragma unroll
            for (int tap = 0; tap < 9; ++tap) {
                unsigned word = unsigned(tile[slot][(y + tap / 3) * TP + x + tap % 3]) & ((1u << kBitsCin) - 1u);
                const float* base = wt + tap * kBitsCin * kRowPitch;
                while (word) {
                    const int ci = __ffs(word) - 1;
                    word &= word - 1;
                    const float4* r = reinterpret_cast<const float4*>(base + ci * kRowPitch);
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const float4 v = r[q];
                        acc[4 * q] += v.x; acc[4 * q + 1] += v.y; acc[4 * q + 2] += v.z; acc[4 * q + 3] += v.w;
                    }
                }
            }
            float* o = out + f * (kBitsCout * HW) + pix;
#pragma unroll
            for (int c = 0; c < kBitsCout; ++c) o[c * HW] = acc[c];
        }
        __syncthreads();
    }
}

template <int H>
int launch_bits(const int* obs_bits, const float* weight, const float* bias, float* out, long long n_frames, cudaStream_t st) {
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    constexpr int FPC = 256 / (H * H) > 0 ? 256 / (H * H) : 1;
    const long long passes = (n_frames + FPC - 1) / FPC, cap = (long long)sm_count(dev) * 6;
    conv1_bits_fwd_kernel<H><<<int(passes < cap ? passes : cap), 256, 0, st>>>(obs_bits, weight, bias, out, n_frames);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}

}  // namespace
}  // namespace mb

extern "C" int mb_conv1_bits_fwd(const int32_t* obs_bits, const float* weight, const float* bias, float* out, int64_t n_frames,
                                 int32_t planes, int32_t cout, int32_t h, int32_t w, void* stream) {
    if (n_frames < 0 || planes != mb::kBitsCin || cout != mb::kBitsCout || h != w || (h != 8 && h != 16)) {
        mb::set_error("mb_conv1_bits_fwd: covers 27 planes -> 16 channels on 8x8 and 16x16 maps (got planes=%d cout=%d %dx%d)",
                      planes, cout, h, w);
        return MB_ERR_UNSUPPORTED;
    }
    if (n_frames == 0) return MB_OK;
    if (!obs_bits || !weight || !out) {
        mb::set_error("mb_conv1_bits_fwd: null pointer");
        return MB_ERR_INVALID;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (h == 8) return mb::launch_bits<8>(obs_bits, weight, bias, out, n_frames, st);
    return mb::launch_bits<16>(obs_bits, weight, bias, out, n_frames, st);
}
