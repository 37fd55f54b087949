// First encoder convolution straight from bit-packed observations, for sm_100a.
//
// Reference: model.py:88 / model.py:121 -- ConvSequence(27 -> 16).conv applied to obs.permute(0, 3, 1, 2)
// (model.py:156-160).  gym-microRTS observations are 27 binary feature planes; the rollout store ships them as one
// int32 per cell (bit c = plane c, csrc/unpack.cu).  A 3x3 convolution over 0/1 inputs is a sum of weight columns:
//     out[n, co, y, x] = bias[co] + sum over the 9 taps, over the SET bits c of cell (y+ky-1, x+kx-1): W[co, c, ky, kx]
// -- and no [N, 27, h, w] fp32 tensor has to be written and read back.  The kernel is bound by the shared-memory
// reads of those 16-float columns (64 bytes into registers per column and pixel), so it counts EXCEPTIONS instead of
// set bits: gym-microRTS leaves a cell empty by setting plane 0 of each of its five one-hot groups (planes 0, 5, 10,
// 13, 21), which is what ~90 % of the cells look like.  Per pixel it starts from bias + the sum of those five columns
// over the taps inside the image (nine precomputed border classes), then adds the column of every plane a cell has
// beyond that pattern and subtracts the column of every pattern plane the cell lacks; for a one-hot group the two
// are one precomputed "swap" column W[p] - W[default] (see the table below): ~0.5 columns per cell instead of 5.
// Any bit pattern is handled exactly the same way (the pattern is only where the counting starts; fp32 additions,
// another summation order); out-of-image taps carry the pattern in the halo, i.e. no exception.
//
// One thread per output pixel, 16 accumulators (the 16 output channels).  The weights sit in shared memory as
// [tap][plane][16] rows padded to 80 bytes.
#include "common.cuh"

#include "../../include/mbeast.h"

namespace mb {

int sm_count(int device);

namespace {

constexpr int kBitsCout = 16, kBitsCin = 27;
constexpr int kRowPitch = 20;                       // floats per [tap][plane] row: 16 + 4 padding

// Virtual columns of the table a pixel adds up (all with sign +):
//   [0, 27)   "swap" columns  W[p] - W[d(g)]  : group g of cell holds exactly the one plane p != its default d(g)
//   [27, 54)  plain columns   W[p]            : a non-default plane of a group that is not one-hot
//   [54, 59)  "drop" columns  -W[d(g)]        : a group that is not one-hot and lacks its default plane
// A one-hot cell (every real observation) costs ONE column per group that departs from the default, half of the
// set-minus / plus pairs of a plain exception count; any other bit pattern is still summed exactly.
constexpr int kGroups = 5;
// first plane of each group (= its default, empty-cell plane) and the group sizes: hp, resources, owner, unit type,
// action (model.py:156-160 / gym-microRTS observation layout)
__host__ __device__ constexpr int group_base(int g) { return g == 0 ? 0 : g == 1 ? 5 : g == 2 ? 10 : g == 3 ? 13 : 21; }
__host__ __device__ constexpr int group_len(int g) { return g == 0 ? 5 : g == 1 ? 5 : g == 2 ? 3 : g == 3 ? 8 : 6; }
constexpr int kVirtCols = 2 * kBitsCin + kGroups;              // 59

__device__ __forceinline__ unsigned long long virt_mask(unsigned word) {
    unsigned long long v = 0ull;
#pragma unroll
    for (int g = 0; g < kGroups; ++g) {
        const unsigned bits = (word >> group_base(g)) & ((1u << group_len(g)) - 1u);
        if (bits == 1u) continue;                              // the default plane alone: nothing to add
        if (__popc(bits) == 1) {
            v |= (unsigned long long)bits << group_base(g);    // swap column of the one plane that is set
        } else {
            v |= (unsigned long long)(bits & ~1u) << (kBitsCin + group_base(g));       // plain columns of the other planes
            if (!(bits & 1u)) v |= 1ull << (2 * kBitsCin + g);                         // and the default plane is missing
        }
    }
    return v;
}

template <int H>
__global__ void __launch_bounds__(256) conv1_bits_fwd_kernel(const int* __restrict__ obs_bits, const float* __restrict__ weight,
                                                             const float* __restrict__ bias, float* __restrict__ out,
                                                             long long n_frames) {
    constexpr int W = H, HW = H * W;
    constexpr int FPC = 256 / HW > 0 ? 256 / HW : 1;                  // frames per CTA pass (8x8: 4, 16x16: 1)
    constexpr int TP = W + 2;                                          // tile pitch with halo
    __shared__ __align__(16) float wt[9 * kVirtCols * kRowPitch];     // [tap][virtual column][16 + 4]
    __shared__ __align__(16) float start[9][kBitsCout];               // bias + empty-cell columns of the taps inside, per border class
    __shared__ unsigned long long tile[FPC][(H + 2) * TP];             // per cell: which virtual columns it contributes
    // torch layout [cout][cin][3][3] -> the virtual column table
    for (int i = threadIdx.x; i < 9 * kVirtCols * kBitsCout; i += 256) {
        const int co = i % kBitsCout, vc = (i / kBitsCout) % kVirtCols, tap = i / (kBitsCout * kVirtCols);
        auto wv = [&](int ci) { return weight[(co * kBitsCin + ci) * 9 + tap]; };
        float v;
        if (vc < 2 * kBitsCin) {
            const int p = vc % kBitsCin;
            int g = 0;
#pragma unroll
            for (int q = 1; q < kGroups; ++q) g += (p >= group_base(q)) ? 1 : 0;
            v = vc < kBitsCin ? wv(p) - wv(group_base(g)) : wv(p);
        } else {
            v = -wv(group_base(vc - 2 * kBitsCin));
        }
        wt[(tap * kVirtCols + vc) * kRowPitch + co] = v;
    }
    for (int i = threadIdx.x; i < FPC * (H + 2) * TP; i += 256) (&tile[0][0])[i] = 0ull;       // halo: empty cells, no column
    if (threadIdx.x < 9 * kBitsCout) {
        const int co = threadIdx.x % kBitsCout, cls = threadIdx.x / kBitsCout, cy = cls / 3, cx = cls % 3;   // 0 first, 1 inner, 2 last
        float v = bias ? bias[co] : 0.0f;
        for (int tap = 0; tap < 9; ++tap) {
            const int ky = tap / 3, kx = tap % 3;
            if ((cy == 0 && ky == 0) || (cy == 2 && ky == 2) || (cx == 0 && kx == 0) || (cx == 2 && kx == 2)) continue;
            for (int g = 0; g < kGroups; ++g) v += weight[(co * kBitsCin + group_base(g)) * 9 + tap];
        }
        start[cls][co] = v;
    }
    __syncthreads();
    const int slot = threadIdx.x / HW, pix = threadIdx.x % HW;
    const int y = pix / W, x = pix % W;
    const int cls = (y == 0 ? 0 : y == H - 1 ? 2 : 1) * 3 + (x == 0 ? 0 : x == W - 1 ? 2 : 1);
    const long long n_pass = (n_frames + FPC - 1) / FPC;
    for (long long pass = blockIdx.x; pass < n_pass; pass += gridDim.x) {
        const long long f = pass * FPC + slot;
        const bool live = slot < FPC && f < n_frames;
        if (slot < FPC)
            tile[slot][(y + 1) * TP + x + 1] = live ? virt_mask(unsigned(__ldg(obs_bits + f * HW + pix)) & ((1u << kBitsCin) - 1u)) : 0ull;
        __syncthreads();
        if (live) {
            float acc[kBitsCout];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float4 v = reinterpret_cast<const float4*>(start[cls])[q];
                acc[4 * q] = v.x, acc[4 * q + 1] = v.y, acc[4 * q + 2] = v.z, acc[4 * q + 3] = v.w;
            }
#pragma unroll
            for (int tap = 0; tap < 9; ++tap) {
                const unsigned long long vm = tile[slot][(y + tap / 3) * TP + x + tap % 3];
                const float* base = wt + tap * kVirtCols * kRowPitch;
                unsigned lo = unsigned(vm), hi = unsigned(vm >> 32);
                while (lo | hi) {                              // one trip per departing group of a one-hot cell
                    const bool in_lo = lo != 0u;
                    const unsigned w32 = in_lo ? lo : hi;
                    const int vc = __ffs(w32) - 1 + (in_lo ? 0 : 32);
                    if (in_lo) lo &= lo - 1; else hi &= hi - 1;
                    const float4* r = reinterpret_cast<const float4*>(base + vc * kRowPitch);
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

// ---------------------------------------------------------------------------------------------
// Weight (+ bias) gradient of that first convolution, again straight from the bits:
//     dW[co, c, ky, kx] = sum over frames and over the cells (y', x') whose bit c is set of dY[n, co, y'-ky+1, x'-kx+1]
// Work follows the EXCEPTIONS of each (plane, frame): its set cells, or -- when more than half of the frame's cells
// are set, the "empty" plane of a one-hot group -- its unset cells with the opposite sign, on top of the frame's
// rectangle sums T[co, tap] = sum over all cells of dY[co, y'-ky+1, x'-kx+1].  About 10 % of the dense product at
// microRTS densities; any bit pattern gives the exact sums (fp32, another summation order).
//
// A CTA of nine warps walks frames; warp v owns planes v, v+9, v+18, and a lane owns (output channel, one of two
// exception slots): 27 accumulators.  The warp finds a plane's exceptions with ballots over the frame's cells (no
// list in memory), and every exception is nine shared-memory reads of 16 consecutive channels -- dY frames are staged
// channel-innermost with a zero halo (registers -> shared memory, one frame ahead), so taps need no range checks and
// a warp's read touches two 64-byte runs.  The last warp also forms T and the bias gradient.  One barrier per frame.
// ---------------------------------------------------------------------------------------------
namespace mb {
namespace {

template <int H>
struct BwCfg {
    static constexpr int W = H, HW = H * W;
    static constexpr int PW = W + 2;                          // cells per staged row, halo included
    static constexpr int CP = 20;                             // floats per staged cell: 16 channels + 4 (bank spread)
    static constexpr int FRAME = (H + 2) * PW * CP;           // floats of a staged dY frame
    static constexpr int CHUNKS = HW / 32;                    // 32-cell ballots per plane
    static constexpr int TASKS = HW / 8;                      // staging: a warp moves 2 float4 positions x 16 channels
    static constexpr int WARPS = 9, THREADS = 32 * WARPS;
    static constexpr int ROUNDS = (TASKS + WARPS - 1) / WARPS;
    static constexpr int SMEM = 2 * FRAME * 4;                // double buffered
};

template <int H>
__global__ void __launch_bounds__(BwCfg<H>::THREADS, 3) conv1_bits_wgrad_kernel(const int* __restrict__ obs_bits, const float* __restrict__ dy,
                                                                               float* __restrict__ total, long long n_frames) {
    using C = BwCfg<H>;
    constexpr int W = C::W, HW = C::HW, PW = C::PW, CP = C::CP;
    extern __shared__ __align__(16) float bw_smem[];
    __shared__ float s_T[2][9][kBitsCout];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, co = lane & 15, half = lane >> 4;
    for (int i = threadIdx.x; i < 2 * C::FRAME; i += C::THREADS) bw_smem[i] = 0.0f;            // halos stay zero for good

    float acc[3][9];
#pragma unroll
    for (int j = 0; j < 3; ++j)
#pragma unroll
        for (int k = 0; k < 9; ++k) acc[j][k] = 0.0f;
    float bsum = 0.0f;                                        // last warp, lanes < 16: bias gradient of channel `lane`

    float4 pre[C::ROUNDS];
    unsigned wcur[C::CHUNKS], wnext[C::CHUNKS];
    auto load_frame = [&](long long f, unsigned (&wd)[C::CHUNKS]) {           // dY into registers, the frame's bits too
#pragma unroll
        for (int k = 0; k < C::ROUNDS; ++k) {
            const int task = warp + C::WARPS * k;
            if (task < C::TASKS) pre[k] = __ldg(reinterpret_cast<const float4*>(dy + (f * kBitsCout + co) * HW) + 2 * task + half);
        }
#pragma unroll
        for (int c = 0; c < C::CHUNKS; ++c) wd[c] = unsigned(__ldg(obs_bits + f * HW + 32 * c + lane));
    };
    auto store_frame = [&](float* buf) {                      // registers -> channel-innermost halo frame
#pragma unroll
        for (int k = 0; k < C::ROUNDS; ++k) {
            const int task = warp + C::WARPS * k;
            if (task < C::TASKS) {
                const int pos = 2 * task + half, y = pos / (W / 4), x0 = (pos % (W / 4)) * 4;
                float* d = buf + ((y + 1) * PW + x0 + 1) * CP + co;
                d[0] = pre[k].x, d[CP] = pre[k].y, d[2 * CP] = pre[k].z, d[3 * CP] = pre[k].w;
            }
        }
    };

    long long f = blockIdx.x;
    __syncthreads();
    if (f < n_frames) {
        load_frame(f, wcur);
        store_frame(bw_smem);
    }
    __syncthreads();
    int b = 0;
    for (; f < n_frames; f += gridDim.x) {
        const long long nf = f + gridDim.x;
        const bool has_next = nf < n_frames;
        if (has_next) load_frame(nf, wnext);
        const float* fr = bw_smem + b * C::FRAME;
        bool dense[3];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            const int p = warp + C::WARPS * j;
            unsigned bal[C::CHUNKS];
            int cnt = 0;
#pragma unroll
            for (int c = 0; c < C::CHUNKS; ++c) {
                bal[c] = __ballot_sync(0xffffffffu, (wcur[c] >> p) & 1u);
                cnt += __popc(bal[c]);
            }
            dense[j] = 2 * cnt > HW;
            const float sign = dense[j] ? -1.0f : 1.0f;
#pragma unroll
            for (int c = 0; c < C::CHUNKS; ++c) {
                unsigned m = dense[j] ? ~bal[c] : bal[c];
                while (m) {                                   // two exceptions per trip: one per half-warp
                    const int c0 = __ffs(m) - 1;
                    m &= m - 1;
                    const int c1 = __ffs(m) - 1;              // -1 when none is left
                    m &= m - 1;                               // (0 & anything stays 0)
                    const int mine = half ? c1 : c0;
                    if (mine >= 0) {
                        const int cell = 32 * c + mine, y = cell / W, x = cell % W;
                        const float* q = fr + ((y + 2) * PW + x + 2) * CP + co;     // dY[y - ky + 1][x - kx + 1] = q[-(ky*PW + kx)*CP]
#pragma unroll
                        for (int ky = 0; ky < 3; ++ky)
#pragma unroll
                            for (int kx = 0; kx < 3; ++kx)
                                acc[j][ky * 3 + kx] = fmaf(sign, q[-(ky * PW + kx) * CP], acc[j][ky * 3 + kx]);
                    }
                }
            }
        }
        if (warp == C::WARPS - 1) {                           // rectangle sums and the bias gradient
            float s_all = 0.f, r_first = 0.f, r_last = 0.f, c_first = 0.f, c_last = 0.f;
#pragma unroll
            for (int y = 0; y < H; ++y)
#pragma unroll
                for (int jx = 0; jx < W / 2; ++jx) {
                    const float v = fr[((y + 1) * PW + 2 * jx + half + 1) * CP + co];
                    s_all += v;
                    if (y == 0) r_first += v;
                    if (y == H - 1) r_last += v;
                    if (jx == 0) c_first += half ? 0.f : v;
                    if (jx == W / 2 - 1) c_last += half ? v : 0.f;
                }
            s_all += __shfl_xor_sync(0xffffffffu, s_all, 16);
            r_first += __shfl_xor_sync(0xffffffffu, r_first, 16);
            r_last += __shfl_xor_sync(0xffffffffu, r_last, 16);
            c_first += __shfl_xor_sync(0xffffffffu, c_first, 16);
            c_last += __shfl_xor_sync(0xffffffffu, c_last, 16);
            if (half == 0) {
                const float k00 = fr[(1 * PW + 1) * CP + co], k0L = fr[(1 * PW + W) * CP + co];
                const float kL0 = fr[(H * PW + 1) * CP + co], kLL = fr[(H * PW + W) * CP + co];
#pragma unroll
                for (int ky = 0; ky < 3; ++ky)
#pragma unroll
                    for (int kx = 0; kx < 3; ++kx) {
                        // tap ky sees every dY row but the first (ky = 0) / the last (ky = 2); columns likewise
                        float v = s_all - (ky == 0 ? r_first : ky == 2 ? r_last : 0.f) - (kx == 0 ? c_first : kx == 2 ? c_last : 0.f);
                        if (ky == 0 && kx == 0) v += k00;
                        if (ky == 0 && kx == 2) v += k0L;
                        if (ky == 2 && kx == 0) v += kL0;
                        if (ky == 2 && kx == 2) v += kLL;
                        s_T[b][ky * 3 + kx][co] = v;
                    }
                bsum += s_all;
            }
        }
        if (has_next) store_frame(bw_smem + (b ^ 1) * C::FRAME);
        __syncthreads();
        if (half == 0) {
#pragma unroll
            for (int j = 0; j < 3; ++j)
                if (dense[j]) {
#pragma unroll
                    for (int k = 0; k < 9; ++k) acc[j][k] += s_T[b][k][co];
                }
        }
#pragma unroll
        for (int c = 0; c < C::CHUNKS; ++c) wcur[c] = wnext[c];
        b ^= 1;
    }
    // this CTA's share into the (caller-zeroed) result: dW [16][27][9], then the bias gradient [16]
#pragma unroll
    for (int j = 0; j < 3; ++j)
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            const float v = acc[j][k] + __shfl_xor_sync(0xffffffffu, acc[j][k], 16);
            if (half == 0) atomicAdd(total + (co * kBitsCin + warp + C::WARPS * j) * 9 + k, v);
        }
    if (warp == C::WARPS - 1 && half == 0) atomicAdd(total + kBitsCout * kBitsCin * 9 + co, bsum);
}

template <int H>
int launch_bits_wgrad(const int* obs_bits, const float* dy, float* total, long long n_frames, cudaStream_t st) {
    using C = BwCfg<H>;
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    static bool attr_set[64] = {};
    static int per_sm[64] = {};
    const int d = (dev >= 0 && dev < 64) ? dev : 0;
    if (!attr_set[d]) {
        MB_CUDA(cudaFuncSetAttribute(conv1_bits_wgrad_kernel<H>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
        int n = 0;
        MB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, conv1_bits_wgrad_kernel<H>, C::THREADS, C::SMEM));
        per_sm[d] = n > 0 ? n : 1;
        attr_set[d] = true;
    }
    const long long cap = (long long)sm_count(dev) * per_sm[d];
    conv1_bits_wgrad_kernel<H><<<int(n_frames < cap ? n_frames : cap), C::THREADS, C::SMEM, st>>>(obs_bits, dy, total, n_frames);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}

}  // namespace
}  // namespace mb

extern "C" int mb_conv1_bits_wgrad(const int32_t* obs_bits, const float* dy, float* accumulate_out, int64_t n_frames,
                                   int32_t planes, int32_t cout, int32_t h, int32_t w, void* stream) {
    if (n_frames < 0 || planes != mb::kBitsCin || cout != mb::kBitsCout || h != w || (h != 8 && h != 16)) {
        mb::set_error("mb_conv1_bits_wgrad: covers 27 planes -> 16 channels on 8x8 and 16x16 maps (got planes=%d cout=%d %dx%d)",
                      planes, cout, h, w);
        return MB_ERR_UNSUPPORTED;
    }
    if (n_frames == 0) return MB_OK;
    if (!obs_bits || !dy || !accumulate_out || (reinterpret_cast<uintptr_t>(dy) & 15)) {
        mb::set_error("mb_conv1_bits_wgrad: null pointer, or dy not 16-byte aligned");
        return MB_ERR_INVALID;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (h == 8) return mb::launch_bits_wgrad<8>(obs_bits, dy, accumulate_out, n_frames, st);
    return mb::launch_bits_wgrad<16>(obs_bits, dy, accumulate_out, n_frames, st);
}
