// The encoder's 3x3 / stride 1 / pad 1 convolutions in fp32, for sm_100a: weight (+ bias) gradient first,
// then the forward / data-gradient kernel and the max-pool backward further down.
//
// Reference: the encoder of model.py:56-137 (Residual_Block / ConvSequence) is differentiated by
// loss.backward() at libs/utils.py:334.  The encoder stays on PyTorch; this file replaces ONE cuDNN
// kernel family inside its backward pass.  For these shapes (16-32 channels, 2x2 .. 16x16 maps,
// ~2e4 frames) the weight gradient
//     dW[co, ci, ky, kx] = sum_{n, y, x} dY[n, co, y, x] * X[n, ci, y + ky - 1, x + kx - 1]
// is a reduction over millions of positions into a tiny output (2.3 k - 9.2 k numbers); cuDNN's
// fp32 kernels (wgrad_alg0 / alg1) run it at ~8 TFLOP/s and it was 13.6 ms of the 41 ms learner
// step at 16x16 (profiles/r01_encoder_layers.txt).  Same mathematics, fp32 FMA, only the summation
// order differs from cuDNN's (which is itself not deterministic; this kernel is).
//
// Design (FMA-bound, not HBM-bound: 72 flop per byte read):
//   * the whole dW lives in REGISTERS: a thread owns 4 output channels x 1 input channel x 9 taps =
//     36 accumulators; (COUT/4 x CIN) threads cover dW, and G such groups per 256-thread CTA work on
//     different frames and are summed through shared memory at the end;
//   * frames are staged into shared memory by cp.async (16-byte LDGSTS, zero-fill past the end),
//     double buffered, in the NCHW order they have in global memory; channel planes are padded so
//     that PLANE/4 is odd and the per-lane float4 reads of X (lanes = input channels) are
//     conflict free while the dY reads are warp broadcasts;
//   * rows slide through registers (up / mid / down), so per image row a thread issues 1 X-row load
//     and 4 dY-row loads for 36 * W FMAs; taps that would fall outside the image are skipped at
//     compile time, so there is no halo;
//   * persistent CTAs walk the frame chunks; every CTA writes its partial dW -- and the bias gradient,
//     summed from the dY planes it staged anyway -- to its own slot of a [grid, COUT*CIN*9 + COUT]
//     buffer, and the caller adds the slots up (fixed order: deterministic).
#include <atomic>

#include "common.cuh"

#include "../../include/mbeast.h"

namespace mb {

int sm_count(int device);

namespace {

constexpr int kWgThreads = 256;
constexpr int kWgTM = 4;          // output channels per thread

__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src, bool valid) {
    const uint32_t bytes = valid ? 16u : 0u;            // 0 source bytes: the 16 destination bytes are zero-filled
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <int CIN_, int COUT_, int H_, int CIS_, int G_, int FG_>
struct WgCfg {
    static constexpr int CIN = CIN_, COUT = COUT_, H = H_, W = H_, CIS = CIS_, G = G_, FG = FG_;
    static constexpr int HW = H * W;
    static constexpr int PLANE = ((HW / 4) % 2 == 1) ? HW : HW + 4;     // floats; PLANE/4 odd
    static constexpr int COG = COUT / kWgTM;                            // groups of 4 output channels
    static constexpr int P = COG * CIS;                                 // threads that cover dW once
    static constexpr int CH = G * FG;                                   // frames per chunk
    static constexpr int FRAME = (CIN + COUT) * PLANE;                  // floats per staged frame: X planes, dY planes
    static constexpr int PIECES = (CIN + COUT) * (HW / 4);              // 16-byte pieces per frame
    static constexpr int OUT = COUT * CIN * 9;
    static constexpr int SLOT = OUT + COUT;                             // a CTA's partial: dW, then the bias gradient
    static constexpr int BSL = kWgThreads / COUT;                       // threads per output channel for sum(dY)
    static constexpr int RED = G * P * kWgTM * 9 + kWgThreads;          // floats of the end-of-kernel reduction
    static constexpr int STAGE_FLOATS = 2 * CH * FRAME;
    static constexpr int SMEM_FLOATS = STAGE_FLOATS > RED ? STAGE_FLOATS : RED;
    static constexpr int SMEM = SMEM_FLOATS * 4;
    static_assert(P * G == kWgThreads, "thread layout must fill the CTA");
    static_assert(COUT % kWgTM == 0 && CIS >= CIN && HW % 4 == 0, "shape");
    static_assert(H >= 2, "a 1x1 map has a single tap; not handled here");
};

// One image row of W floats from shared memory.
template <int W>
__device__ __forceinline__ void load_row(float (&r)[W], const float* p) {
    if constexpr (W == 2) {
        const float2 v = *reinterpret_cast<const float2*>(p);
        r[0] = v.x;
        r[1] = v.y;
    } else {
#pragma unroll
        for (int i = 0; i < W / 4; ++i) {
            const float4 v = *reinterpret_cast<const float4*>(p + 4 * i);
            r[4 * i + 0] = v.x;
            r[4 * i + 1] = v.y;
            r[4 * i + 2] = v.z;
            r[4 * i + 3] = v.w;
        }
    }
}

// acc[m][ky*3+kx] += dY[m][x] * Xrow_ky[x + kx - 1] for one output row; rows outside the image are
// compiled out (UP / DN), and so are the columns (static range check on x + kx - 1).
template <int W, bool UP, bool DN>
__device__ __forceinline__ void row_fma(float (&acc)[kWgTM][9], const float (&up)[W], const float (&mid)[W],
                                        const float (&dn)[W], const float* ybase, int plane) {
    constexpr int XB = W < 8 ? W : 8;                    // dY columns held in registers at a time
#pragma unroll
    for (int x0 = 0; x0 < W; x0 += XB) {
        float d[kWgTM][XB];
#pragma unroll
        for (int m = 0; m < kWgTM; ++m) load_row<XB>(d[m], ybase + m * plane + x0);
#pragma unroll
        for (int xx = 0; xx < XB; ++xx) {
            const int x = x0 + xx;
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
                const int xi = x + kx - 1;
                if (xi < 0 || xi >= W) continue;
#pragma unroll
                for (int m = 0; m < kWgTM; ++m) {
                    if (UP) acc[m][0 + kx] = fmaf(d[m][xx], up[xi], acc[m][0 + kx]);
                    acc[m][3 + kx] = fmaf(d[m][xx], mid[xi], acc[m][3 + kx]);
                    if (DN) acc[m][6 + kx] = fmaf(d[m][xx], dn[xi], acc[m][6 + kx]);
                }
            }
        }
    }
}

// All rows of one frame for this thread's (4 output channels, 1 input channel).
template <int W>
__device__ __forceinline__ void relu_row(float (&r)[W]) {
#pragma unroll
    for (int i = 0; i < W; ++i) r[i] = fmaxf(r[i], 0.0f);
}

template <class C>
__device__ __forceinline__ void frame_fma(float (&acc)[kWgTM][9], const float* xplane, const float* ybase, bool relu_x) {
    constexpr int W = C::W, H = C::H;
    float up[W], mid[W], dn[W];
    load_row<W>(mid, xplane);
    load_row<W>(dn, xplane + W);
    if (relu_x) {
        relu_row<W>(mid);
        relu_row<W>(dn);
    }
    row_fma<W, false, true>(acc, mid, mid, dn, ybase, C::PLANE);                 // y = 0: no row above
    if constexpr (H > 2) {
        // small maps unroll completely; at 16x16 the 14 interior rows would be 8 k FMAs of straight-line
        // code (> the instruction cache), so they run as a loop unrolled by two
#pragma unroll(H <= 8 ? H : 2)
        for (int y = 1; y < H - 1; ++y) {
#pragma unroll
            for (int i = 0; i < W; ++i) {
                up[i] = mid[i];
                mid[i] = dn[i];
            }
            load_row<W>(dn, xplane + (y + 1) * W);
            if (relu_x) relu_row<W>(dn);
            row_fma<W, true, true>(acc, up, mid, dn, ybase + y * W, C::PLANE);
        }
    }
#pragma unroll
    for (int i = 0; i < W; ++i) {
        up[i] = mid[i];
        mid[i] = dn[i];
    }
    row_fma<W, true, false>(acc, up, mid, mid, ybase + (H - 1) * W, C::PLANE);   // y = H-1: no row below
}

template <class C>
__global__ void __launch_bounds__(kWgThreads) conv3x3_wgrad_kernel(const float* __restrict__ x,
                                                                    const float* __restrict__ dy,
                                                                    float* __restrict__ partials, long long n_frames,
                                                                    int relu_x, float* __restrict__ atomic_out) {
    extern __shared__ __align__(16) float wg_smem[];
    const int tid = threadIdx.x;
    const int g = tid / C::P;                       // frame group
    const int r = tid % C::P;
    const int cog = r / C::CIS;                     // output channels 4*cog .. 4*cog+3
    const int ci = r % C::CIS;                      // input channel (lanes past CIN idle along on channel CIN-1)
    const int ci_eff = ci < C::CIN ? ci : C::CIN - 1;
    const long long n_chunks = (n_frames + C::CH - 1) / C::CH;

    float acc[kWgTM][9];
#pragma unroll
    for (int m = 0; m < kWgTM; ++m)
#pragma unroll
        for (int t = 0; t < 9; ++t) acc[m][t] = 0.0f;
    const int b_co = tid % C::COUT, b_sl = tid / C::COUT;
    float bsum = 0.0f;

    auto stage = [&](long long chunk, int buf) {
        float* base = wg_smem + buf * (C::CH * C::FRAME);
        const long long f0 = chunk * C::CH;
        for (int i = tid; i < C::CH * C::PIECES; i += kWgThreads) {
            const int fs = i / C::PIECES, piece = i % C::PIECES;
            const int plane = piece / (C::HW / 4), within = piece % (C::HW / 4);
            const long long f = f0 + fs;
            const bool valid = f < n_frames;
            const float* src = x;
            if (valid) {
                src = (plane < C::CIN) ? x + (f * C::CIN + plane) * C::HW + within * 4
                                       : dy + (f * C::COUT + (plane - C::CIN)) * C::HW + within * 4;
            }
            cp_async16(base + fs * C::FRAME + plane * C::PLANE + within * 4, src, valid);
        }
        cp_async_commit();
    };

    long long chunk = blockIdx.x;
    int buf = 0;
    if (chunk < n_chunks) stage(chunk, 0);
    for (; chunk < n_chunks; chunk += gridDim.x) {
        const long long next = chunk + gridDim.x;
        if (next < n_chunks) {
            stage(next, buf ^ 1);           // its buffer was released by the barrier that ended the last iteration
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        const float* base = wg_smem + buf * (C::CH * C::FRAME);
#pragma unroll 1
        for (int j = 0; j < C::FG; ++j) {
            const float* fr = base + (g * C::FG + j) * C::FRAME;
            frame_fma<C>(acc, fr + ci_eff * C::PLANE, fr + (C::CIN + cog * kWgTM) * C::PLANE, relu_x != 0);
        }
        // bias gradient = sum of dY over frames and pixels: thread (b_co, b_sl) adds its share of the staged dY
        // planes of channel b_co (a handful of float4 per chunk next to thousands of FMAs)
#pragma unroll 1
        for (int fs = 0; fs < C::CH; ++fs) {
            const float* pl = base + fs * C::FRAME + (C::CIN + b_co) * C::PLANE;
            for (int q4 = b_sl; q4 < C::HW / 4; q4 += C::BSL) {
                const float4 v = *reinterpret_cast<const float4*>(pl + 4 * q4);
                bsum += (v.x + v.y) + (v.z + v.w);
            }
        }
        __syncthreads();
        buf ^= 1;
    }

    // fold the G frame groups and write this CTA's slot
    float* red = wg_smem;
#pragma unroll
    for (int m = 0; m < kWgTM; ++m)
#pragma unroll
        for (int t = 0; t < 9; ++t) red[tid * (kWgTM * 9) + m * 9 + t] = acc[m][t];
    float* bred = red + kWgThreads * (kWgTM * 9);
    bred[tid] = bsum;
    __syncthreads();
    // atomic_out: add this CTA's share straight into the [OUT + COUT] result (zeroed by the caller) instead of writing a
    // slot for a later fixed-order sum: one pass less per convolution, at the price of a run-to-run summation order
    float* out = atomic_out ? atomic_out : partials + (long long)blockIdx.x * C::SLOT;
    if (tid < C::COUT) {
        float s = 0.0f;
#pragma unroll
        for (int k = 0; k < C::BSL; ++k) s += bred[k * C::COUT + tid];
        if (atomic_out) atomicAdd(out + C::OUT + tid, s);
        else out[C::OUT + tid] = s;
    }
    for (int e = tid; e < C::OUT; e += kWgThreads) {
        const int co = e / (C::CIN * 9), rem = e % (C::CIN * 9);
        const int ci2 = rem / 9, tap = rem % 9;
        float s = 0.0f;
#pragma unroll
        for (int gg = 0; gg < C::G; ++gg)
            s += red[((gg * C::COG + co / kWgTM) * C::CIS + ci2) * (kWgTM * 9) + (co % kWgTM) * 9 + tap];
        if (atomic_out) atomicAdd(out + e, s);
        else out[e] = s;
    }
}

template <class C>
int wg_grid(int dev, long long n_frames, int* grid) {
    static int occ[64] = {};
    int& o = occ[(dev >= 0 && dev < 64) ? dev : 0];
    if (o == 0) {
        MB_CUDA(cudaFuncSetAttribute(conv3x3_wgrad_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
        int n = 0;
        MB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, conv3x3_wgrad_kernel<C>, kWgThreads, C::SMEM));
        o = n > 0 ? n : 1;
    }
    const long long chunks = (n_frames + C::CH - 1) / C::CH;
    const long long cap = (long long)sm_count(dev) * o;
    *grid = int(chunks < cap ? (chunks > 0 ? chunks : 1) : cap);
    return MB_OK;
}

template <class C>
int wg_launch(const float* x, const float* dy, float* partials, int n_blocks, long long n_frames, int relu_x,
              float* atomic_out, cudaStream_t st) {
    int dev = 0, grid = 0;
    MB_CUDA(cudaGetDevice(&dev));
    const int rc = wg_grid<C>(dev, n_frames, &grid);
    if (rc != MB_OK) return rc;
    if (n_blocks != grid) {
        set_error("mb_conv3x3_wgrad: n_blocks=%d, but mb_conv3x3_wgrad_blocks gives %d for this shape", n_blocks, grid);
        return MB_ERR_INVALID;
    }
    conv3x3_wgrad_kernel<C><<<grid, kWgThreads, C::SMEM, st>>>(x, dy, partials, n_frames, relu_x, atomic_out);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}

// shape -> configuration <CIN, COUT, H, ci slots, frame groups, frames per group and chunk>
#define MB_WG_SHAPES(X)                     \
    X(27, 16, 16, 32, 2, 1) \
    X(27, 16, 8, 32, 2, 2)   \
    X(16, 16, 8, 16, 4, 1)   \
    X(16, 16, 4, 16, 4, 2)   \
    X(16, 32, 16, 16, 2, 1) \
    X(16, 32, 8, 16, 2, 1)   \
    X(16, 32, 4, 16, 2, 2)   \
    X(32, 32, 8, 32, 1, 2)   \
    X(32, 32, 4, 32, 1, 4)   \
    X(32, 32, 2, 32, 1, 16)

}  // namespace

int conv3x3_wgrad_blocks(long long n_frames, int cin, int cout, int h, int w, int* grid) {
    *grid = 0;
    if (h != w) return MB_OK;
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
#define X(CI, CO, HH, CIS, GG, FG) \
    if (cin == CI && cout == CO && h == HH) return wg_grid<WgCfg<CI, CO, HH, CIS, GG, FG>>(dev, n_frames, grid);
    MB_WG_SHAPES(X)
#undef X
    return MB_OK;          // *grid == 0: shape not covered, the caller keeps its own path
}

int conv3x3_wgrad(const float* x, const float* dy, float* partials, int n_blocks, long long n_frames, int cin,
                  int cout, int h, int w, int relu_x, float* atomic_out, cudaStream_t st) {
    if (h == w) {
#define X(CI, CO, HH, CIS, GG, FG)              \
    if (cin == CI && cout == CO && h == HH)     \
        return wg_launch<WgCfg<CI, CO, HH, CIS, GG, FG>>(x, dy, partials, n_blocks, n_frames, relu_x, atomic_out, st);
        MB_WG_SHAPES(X)
#undef X
    }
    set_error("mb_conv3x3_wgrad: shape cin=%d cout=%d %dx%d is not covered (see mb_conv3x3_wgrad_blocks)", cin, cout, h, w);
    return MB_ERR_UNSUPPORTED;
}

// ---------------------------------------------------------------------------------------------
// Forward-type 3x3 / stride 1 / pad 1 convolution in fp32:
//     out[n, co, y, x] = bias[co] + sum_{ci, ky, kx} wk[ci][co][ky*3+kx] * in[n, ci, y+ky-1, x+kx-1]
// used for the encoder's forward convolutions (wk = W permuted to [cin][cout][9]) and, with the kernel
// flipped and the channel axes swapped, for their data gradients (wk = W.flip(2,3) viewed [cout][cin][9],
// no bias).  cuDNN runs these small-channel shapes at ~20 TFLOP/s (implicit_convolve_sgemm / winograd,
// 9 ms of a 23 ms step at 16x16).
//
// A thread owns 4 output channels x one row segment of up to 8 pixels (32 accumulators) of one frame and
// walks the input channels: per channel it reads its three input rows and the 36 weights of its channel
// group from shared memory (float4) and issues 36*8 FMAs; taps outside the image are compiled out.  Lanes
// of a warp that share a row read the same input addresses, lanes that share a channel group the same
// weights (broadcasts), so shared-memory traffic is ~15 wavefronts per 288 FMA issue slots.  Frames are
// staged by cp.async, double buffered, exactly as they lie in global memory; the weights stay resident.
// ---------------------------------------------------------------------------------------------
template <int CIN_, int COUT_, int H_, int FG_, int FPT_>
struct FwCfg {
    static constexpr int CIN = CIN_, COUT = COUT_, H = H_, W = H_, FG = FG_, FPT = FPT_;
    static constexpr int HW = H * W;
    static constexpr int XW = W < 8 ? W : 8;                // pixels of a row per thread
    static constexpr int SEG = W / XW;
    static constexpr int RPT = (H == 2) ? 2 : 1;            // rows per thread (a 2x2 map is one thread's)
    static constexpr int ROWS = H / RPT;
    static constexpr int COG = COUT / kWgTM;
    static constexpr int TPF = COG * ROWS * SEG;            // threads per frame
    static constexpr int FPP = kWgThreads / TPF;            // frames per pass
    static constexpr int CH = FPP * FG * FPT;               // frames per chunk; a thread takes FPT of them at a time
    // staged planes carry one zero row above and below (not for 2x2, where the row index is static), so every
    // lane of a warp runs the same three-row code whatever its row
    static constexpr int HALO = (H > 2) ? 1 : 0;
    static constexpr int PLANE = (H + 2 * HALO) * W;        // floats per staged plane
    static constexpr int XFRAME = CIN * HW;                 // floats per frame in global memory
    static constexpr int SFRAME = CIN * PLANE;              // floats per staged frame
    static constexpr int WK = CIN * COUT * 9;
    static constexpr int SMEM = (2 * CH * SFRAME + WK + COUT) * 4;
    // lanes of a quarter warp (the unit a 128-bit shared load is served in) cover YL consecutive rows -- 128 B,
    // every bank once -- times CL channel groups (which read the same row addresses: broadcasts)
    static constexpr int YL0 = (32 / W) < 1 ? 1 : (32 / W);
    static constexpr int YL = ROWS < YL0 ? ROWS : YL0;
    static constexpr int CL = 8 / YL;
    static constexpr int YH = ROWS / YL;
    static_assert(kWgThreads % TPF == 0 && FPP >= 1, "thread layout");
    static_assert(COG % CL == 0 && ROWS % YL == 0, "quarter-warp layout");
    static_assert(COUT % kWgTM == 0 && HW % 4 == 0 && (WK + COUT) % 4 == 0 && (W * HALO) % 4 == 0, "shape / alignment");
};

// acc[m][x] += w[m][ky*3+kx] * row_ky[X0 + x + kx - 1] over one input channel; columns outside the image are
// compiled out (X0, the segment's first column, is a template constant)
template <int W, int XW, int X0, bool UP, bool DN>
__device__ __forceinline__ void fwd_fma(float (&acc)[kWgTM][XW], const float (&wt)[kWgTM][9], const float (&up)[W],
                                        const float (&mid)[W], const float (&dn)[W]) {
#pragma unroll
    for (int x = 0; x < XW; ++x) {
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
            const int xi = X0 + x + kx - 1;
            if (xi < 0 || xi >= W) continue;
#pragma unroll
            for (int m = 0; m < kWgTM; ++m) {
                if (UP) acc[m][x] = fmaf(wt[m][0 + kx], up[xi], acc[m][x]);
                acc[m][x] = fmaf(wt[m][3 + kx], mid[xi], acc[m][x]);
                if (DN) acc[m][x] = fmaf(wt[m][6 + kx], dn[xi], acc[m][x]);
            }
        }
    }
}

// FPT frames x RPT rows of one thread over all input channels: the 36 weights of a channel are read once
// and used for every frame and row the thread holds.  `rowp[t]` points at the thread's (first) row in the
// staged plane of channel 0 of frame t.
template <class C, int X0>
__device__ __forceinline__ void fwd_rows(float (&acc)[C::FPT][C::RPT][kWgTM][C::XW], const float* const (&rowp)[C::FPT],
                                         const float* wcog, bool relu_in) {
    constexpr int W = C::W;
#pragma unroll 2
    for (int ci = 0; ci < C::CIN; ++ci) {
        float wt[kWgTM][9];
        const float* wp = wcog + ci * (C::COUT * 9);
#pragma unroll
        for (int q = 0; q < 9; ++q) {
            const float4 v = *reinterpret_cast<const float4*>(wp + 4 * q);
            (&wt[0][0])[4 * q + 0] = v.x;
            (&wt[0][0])[4 * q + 1] = v.y;
            (&wt[0][0])[4 * q + 2] = v.z;
            (&wt[0][0])[4 * q + 3] = v.w;
        }
#pragma unroll
        for (int t = 0; t < C::FPT; ++t) {
            const float* pr = rowp[t] + ci * C::PLANE;
            if constexpr (C::H == 2) {                     // the whole 2x2 plane: row 0 then row 1
                float r0[W], r1[W];
                load_row<W>(r0, pr);
                load_row<W>(r1, pr + W);
                if (relu_in) {
                    relu_row<W>(r0);
                    relu_row<W>(r1);
                }
                fwd_fma<W, C::XW, X0, false, true>(acc[t][0], wt, r0, r0, r1);
                fwd_fma<W, C::XW, X0, true, false>(acc[t][1], wt, r0, r1, r1);
            } else {                                       // halo rows make -W / +W valid for every row
                float up[W], mid[W], dn[W];
                load_row<W>(mid, pr);
                load_row<W>(up, pr - W);
                load_row<W>(dn, pr + W);
                if (relu_in) {                             // halo rows are zero: relu leaves them zero
                    relu_row<W>(mid);
                    relu_row<W>(up);
                    relu_row<W>(dn);
                }
                fwd_fma<W, C::XW, X0, true, true>(acc[t][0], wt, up, mid, dn);
            }
        }
    }
}

template <class C>
__global__ void __launch_bounds__(kWgThreads) conv3x3_fwd_kernel(const float* __restrict__ in, const float* __restrict__ wk,
                                                                  const float* __restrict__ bias, float* __restrict__ out,
                                                                  long long n_frames, int wmode, int relu_in,
                                                                  const float* __restrict__ mask_src,
                                                                  const float* __restrict__ add_src) {
    extern __shared__ __align__(16) float fw_smem[];
    float* wsm = fw_smem;                              // [CIN][COUT][9]
    float* bsm = wsm + C::WK;                          // [COUT]
    float* xsm = bsm + C::COUT;                        // 2 x [CH][CIN][H + halo][W]
    const int tid = threadIdx.x;
    // weights into [CIN][COUT][9]: wmode 0 = given so; 1 = a torch Conv2d weight [COUT][CIN][3][3] (forward);
    // 2 = the torch weight of the convolution whose DATA GRADIENT this is, [CIN][COUT][3][3] with the taps flipped
    for (int i = tid; i < C::WK; i += kWgThreads) {
        const int k = i % 9, co = (i / 9) % C::COUT, ci = i / (9 * C::COUT);
        wsm[i] = wmode == 0 ? wk[i] : wmode == 1 ? wk[(co * C::CIN + ci) * 9 + k] : wk[(ci * C::COUT + co) * 9 + (8 - k)];
    }
    for (int i = tid; i < C::COUT; i += kWgThreads) bsm[i] = bias ? bias[i] : 0.0f;
    for (int i = tid; i < 2 * C::CH * C::SFRAME; i += kWgThreads) xsm[i] = 0.0f;      // halo rows stay zero for good
    __syncthreads();                                   // before any cp.async lands on the zeroed interior

    // thread -> (frame slot, row segment, channel group, row); the segment is uniform per warp and the low lane
    // bits follow the quarter-warp layout of FwCfg (YL rows x CL channel groups)
    const int slot = tid / C::TPF;
    const int r = tid % C::TPF;
    const int seg = r / (C::COG * C::ROWS);
    const int r2 = r % (C::COG * C::ROWS);
    const int yl = r2 % C::YL, cl = (r2 / C::YL) % C::CL;
    const int yh = (r2 / (C::YL * C::CL)) % C::YH, chh = r2 / (C::YL * C::CL * C::YH);
    const int yy = yl + C::YL * yh;
    const int cog = cl + C::CL * chh;
    const float* wcog = wsm + cog * (kWgTM * 9);

    const long long n_chunks = (n_frames + C::CH - 1) / C::CH;
    auto stage = [&](long long chunk, int buf) {
        float* base = xsm + buf * (C::CH * C::SFRAME);
        const long long f0 = chunk * C::CH;
        constexpr int PIECES = C::CH * C::XFRAME / 4;
        for (int i = tid; i < PIECES; i += kWgThreads) {
            const int e = i * 4;
            const int fs = e / C::XFRAME, rem = e % C::XFRAME;
            const int plane = rem / C::HW, off = rem % C::HW;
            const bool valid = f0 + fs < n_frames;
            cp_async16(base + fs * C::SFRAME + plane * C::PLANE + C::HALO * C::W + off,
                       valid ? in + f0 * C::XFRAME + e : in, valid);
        }
        cp_async_commit();
    };

    long long chunk = blockIdx.x;
    int buf = 0;
    if (chunk < n_chunks) stage(chunk, 0);
    for (; chunk < n_chunks; chunk += gridDim.x) {
        const long long next = chunk + gridDim.x;
        if (next < n_chunks) {
            stage(next, buf ^ 1);
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        const float* base = xsm + buf * (C::CH * C::SFRAME);
#pragma unroll 1
        for (int j = 0; j < C::FG; ++j) {
            float acc[C::FPT][C::RPT][kWgTM][C::XW];
            const float* rowp[C::FPT];
#pragma unroll
            for (int t = 0; t < C::FPT; ++t) {
                const int fs = (j * C::FPT + t) * C::FPP + slot;
                rowp[t] = base + fs * C::SFRAME + (yy * C::RPT + C::HALO) * C::W;
#pragma unroll
                for (int rr = 0; rr < C::RPT; ++rr)
#pragma unroll
                    for (int m = 0; m < kWgTM; ++m)
#pragma unroll
                        for (int x = 0; x < C::XW; ++x) acc[t][rr][m][x] = bsm[cog * kWgTM + m];
            }
            if (C::SEG == 1 || seg == 0) {             // uniform per warp
                fwd_rows<C, 0>(acc, rowp, wcog, relu_in != 0);
            } else {
                fwd_rows<C, (C::SEG > 1 ? 8 : 0)>(acc, rowp, wcog, relu_in != 0);
            }
#pragma unroll
            for (int t = 0; t < C::FPT; ++t) {
                const long long f = chunk * C::CH + (j * C::FPT + t) * C::FPP + slot;
                if (f < n_frames) {
#pragma unroll
                    for (int rr = 0; rr < C::RPT; ++rr)
#pragma unroll
                        for (int m = 0; m < kWgTM; ++m) {
                            const long long oi = ((f * C::COUT + cog * kWgTM + m) * C::H + yy * C::RPT + rr) * C::W + seg * C::XW;
                            float* o = out + oi;
                            // fused element-wise tails: ReLU backward (zero where the reference's pre-activation was
                            // not positive) and / or a residual add, at the positions this thread writes
                            if (mask_src) {
#pragma unroll
                                for (int x = 0; x < C::XW; ++x) acc[t][rr][m][x] = __ldg(mask_src + oi + x) > 0.0f ? acc[t][rr][m][x] : 0.0f;
                            }
                            if (add_src) {
#pragma unroll
                                for (int x = 0; x < C::XW; ++x) acc[t][rr][m][x] += __ldg(add_src + oi + x);
                            }
                            if constexpr (C::XW == 2) {
                                *reinterpret_cast<float2*>(o) = make_float2(acc[t][rr][m][0], acc[t][rr][m][1]);
                            } else {
#pragma unroll
                                for (int q = 0; q < C::XW / 4; ++q)
                                    *reinterpret_cast<float4*>(o + 4 * q) =
                                        make_float4(acc[t][rr][m][4 * q], acc[t][rr][m][4 * q + 1],
                                                    acc[t][rr][m][4 * q + 2], acc[t][rr][m][4 * q + 3]);
                            }
                        }
                }
            }
        }
        __syncthreads();
        buf ^= 1;
    }
}

template <class C>
int fw_launch(const float* in, const float* wk, const float* bias, float* out, long long n_frames, int wmode,
              int relu_in, const float* mask_src, const float* add_src, cudaStream_t st) {
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    static int occ[64] = {};
    int& o = occ[(dev >= 0 && dev < 64) ? dev : 0];
    if (o == 0) {
        MB_CUDA(cudaFuncSetAttribute(conv3x3_fwd_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
        int n = 0;
        MB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, conv3x3_fwd_kernel<C>, kWgThreads, C::SMEM));
        o = n > 0 ? n : 1;
    }
    const long long chunks = (n_frames + C::CH - 1) / C::CH;
    const long long cap = (long long)sm_count(dev) * o;
    const int grid = int(chunks < cap ? (chunks > 0 ? chunks : 1) : cap);
    conv3x3_fwd_kernel<C><<<grid, kWgThreads, C::SMEM, st>>>(in, wk, bias, out, n_frames, wmode, relu_in, mask_src, add_src);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}

// (cin, cout, h, passes per chunk, frames per thread and pass): the encoder's forward shapes and their
// data-gradient mirrors
#define MB_FW_SHAPES(X) \
    X(27, 16, 16, 1, 1) X(27, 16, 8, 1, 1) X(16, 16, 16, 1, 1) X(16, 16, 8, 1, 1) X(16, 16, 4, 1, 1) \
    X(16, 32, 16, 1, 1) X(16, 32, 8, 1, 1) X(16, 32, 4, 1, 1) X(32, 16, 16, 1, 1) X(32, 16, 8, 1, 1) X(32, 16, 4, 1, 1) \
    X(32, 32, 8, 1, 1) X(32, 32, 4, 1, 1) X(32, 32, 2, 1, 2)

int conv3x3_fwd_supported(int cin, int cout, int h, int w) {
    if (h != w) return 0;
#define X(CI, CO, HH, FG, FPT) \
    if (cin == CI && cout == CO && h == HH) return 1;
    MB_FW_SHAPES(X)
#undef X
    return 0;
}

// csrc/convtc.cu: the same convolution on the tensor cores, for the 16- / 32-channel maps of 8x8 and smaller
int conv3x3_tc_supported(int cin, int cout, int h, int w);
int conv3x3_tc(const float* in, const float* wk, const float* bias, float* out, long long n_frames, int cin, int cout, int h,
               int w, int wmode, int relu_in, const float* mask_src, const float* add_src, cudaStream_t st);
static std::atomic<int> g_conv_tc{1};                       // mb_set_tuning("conv_tc", 0): keep every shape on the fp32 FMA kernel below
void conv_set_tuning(int tensor_cores) { g_conv_tc = tensor_cores; }

int conv3x3_fwd(const float* in, const float* wk, const float* bias, float* out, long long n_frames, int cin, int cout,
                int h, int w, int wmode, int relu_in, const float* mask_src, const float* add_src, cudaStream_t st) {
    if (g_conv_tc && conv3x3_tc_supported(cin, cout, h, w))
        return conv3x3_tc(in, wk, bias, out, n_frames, cin, cout, h, w, wmode, relu_in, mask_src, add_src, st);
    if (h == w) {
#define X(CI, CO, HH, FG, FPT)                  \
    if (cin == CI && cout == CO && h == HH)     \
        return fw_launch<FwCfg<CI, CO, HH, FG, FPT>>(in, wk, bias, out, n_frames, wmode, relu_in, mask_src, add_src, st);
        MB_FW_SHAPES(X)
#undef X
    }
    set_error("mb_conv3x3_fwd: shape cin=%d cout=%d %dx%d is not covered (see mb_conv3x3_fwd_supported)", cin, cout, h, w);
    return MB_ERR_UNSUPPORTED;
}

// ---------------------------------------------------------------------------------------------
// Backward of max_pool2d(kernel 3, stride 2, padding 1) (reference model.py:96) from the argmax
// indices the forward saved.  ATen's max_pool_backward_nchw took 2.0 ms of the step for ~0.9 GB of
// traffic; this is a gather: one thread per INPUT pixel looks at the (at most four) windows that
// cover it and adds the gradients of those whose argmax it is.  No atomics, deterministic, every
// dx element written exactly once.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) maxpool3s2_bwd_kernel(const float* __restrict__ dy,
                                                             const long long* __restrict__ idx,
                                                             float* __restrict__ dx, long long planes, int h, int w,
                                                             int oh, int ow) {
    const long long total = planes * h * w;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long pl = i / (h * w);
        const int pix = int(i - pl * (h * w));
        const int y = pix / w, x = pix - y * w;
        // window o covers input rows [2o-1, 2o+1]: y even -> o = y/2; y odd -> o = (y-1)/2 and (y+1)/2
        const int oy0 = y >> 1, oy1 = min((y + 1) >> 1, oh - 1);
        const int ox0 = x >> 1, ox1 = min((x + 1) >> 1, ow - 1);
        const float* g = dy + pl * (oh * ow);
        const long long* ix = idx + pl * (oh * ow);
        float s = 0.0f;
        for (int oy = oy0; oy <= oy1; ++oy)
            for (int ox = ox0; ox <= ox1; ++ox)
                if (__ldg(ix + oy * ow + ox) == pix) s += __ldg(g + oy * ow + ox);
        dx[i] = s;
    }
}

// Same gather, four input pixels of a row per thread (w % 4 == 0): one float4 store, the six windows that can
// cover them read once.  LW / LH = log2 of the (power of two) map size, so the index split is shifts only.
template <int LW, int LH>
__global__ void __launch_bounds__(256) maxpool3s2_bwd_vec_kernel(const float* __restrict__ dy,
                                                                 const long long* __restrict__ idx,
                                                                 float* __restrict__ dx, long long planes) {
    constexpr int W = 1 << LW, H = 1 << LH, OW = W / 2, OH = H / 2, QW = W / 4;
    const long long total = planes * (H * QW);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long pl = i / (H * QW);                       // compile-time power of two: a shift
        const int r = int(i - pl * (H * QW));
        const int y = r / QW, x0 = (r - y * QW) * 4;
        const int oy0 = y >> 1, oy1 = min((y + 1) >> 1, OH - 1);
        const int oxa = x0 >> 1;                                 // windows oxa, oxa+1, oxa+2 (if inside)
        const float* g = dy + pl * (OH * OW);
        const long long* ix = idx + pl * (OH * OW);
        float o[4] = {0.f, 0.f, 0.f, 0.f};
        const int base = y * W + x0;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int oy = k == 0 ? oy0 : oy1;
            if (k == 1 && oy1 == oy0) break;
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                const int ox = oxa + j;
                if (ox >= OW) break;
                const int d = int(__ldg(ix + oy * OW + ox)) - base;     // which of my four pixels, if any
                const float v = __ldg(g + oy * OW + ox);
#pragma unroll
                for (int e = 0; e < 4; ++e) o[e] += (d == e) ? v : 0.0f;
            }
        }
        *reinterpret_cast<float4*>(dx + pl * (H * W) + base) = make_float4(o[0], o[1], o[2], o[3]);
    }
}

// ---------------------------------------------------------------------------------------------
// max_pool2d(kernel 3, stride 2, padding 1) forward with a ONE-BYTE argmax (reference model.py:96).  ATen's
// max_pool_forward_nchw writes an int64 index per output (8 B next to the 4-B value) and took 0.45 ms of the step;
// here the winner is stored as its position inside the 3x3 window, code = ky * 3 + kx, first maximum in row-major
// scan order (ATen's tie rule), and the backward below reads that byte.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) maxpool3s2_fwd_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                             uint8_t* __restrict__ code, long long planes, int h, int w,
                                                             int oh, int ow) {
    const long long total = planes * oh * ow;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long pl = i / (oh * ow);
        const int o = int(i - pl * (oh * ow));
        const int oy = o / ow, ox = o - oy * ow;
        const float* p = x + pl * (h * w);
        float best = -INFINITY;
        int bc = 0;
        bool first = true;
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
            const int yy = 2 * oy - 1 + ky;
            if (yy < 0 || yy >= h) continue;
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
                const int xx = 2 * ox - 1 + kx;
                if (xx < 0 || xx >= w) continue;
                const float v = __ldg(p + yy * w + xx);
                if (first || v > best || v != v) {          // ATen: (val > maxval) || isnan(val)
                    best = v;
                    bc = ky * 3 + kx;
                    first = false;
                }
            }
        }
        y[i] = best;
        code[i] = uint8_t(bc);
    }
}

// gather-form backward from the one-byte codes: an input pixel looks at the <= 4 windows that cover it
__global__ void __launch_bounds__(256) maxpool3s2_bwd_code_kernel(const float* __restrict__ dy, const uint8_t* __restrict__ code,
                                                                  float* __restrict__ dx, long long planes, int h, int w,
                                                                  int oh, int ow) {
    const long long total = planes * h * w;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long pl = i / (h * w);
        const int pix = int(i - pl * (h * w));
        const int y = pix / w, x = pix - y * w;
        const int oy0 = y >> 1, oy1 = min((y + 1) >> 1, oh - 1);
        const int ox0 = x >> 1, ox1 = min((x + 1) >> 1, ow - 1);
        const float* g = dy + pl * (oh * ow);
        const uint8_t* cd = code + pl * (oh * ow);
        float s = 0.0f;
        for (int oy = oy0; oy <= oy1; ++oy)
            for (int ox = ox0; ox <= ox1; ++ox) {
                const int c = __ldg(cd + oy * ow + ox);
                const int wy = 2 * oy - 1 + c / 3, wx = 2 * ox - 1 + c % 3;
                if (wy == y && wx == x) s += __ldg(g + oy * ow + ox);
            }
        dx[i] = s;
    }
}

// Two neighbouring outputs of a row per thread (w % 4 == 0): per input row one 16-byte load (columns 2*ox .. 2*ox+3) and
// the column to their left; float2 / two-byte stores.  LW / LH = log2 of the map size: index split by shifts.
template <int LW, int LH>
__global__ void __launch_bounds__(256) maxpool3s2_fwd_vec_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                                 uint8_t* __restrict__ code, long long planes) {
    constexpr int W = 1 << LW, H = 1 << LH, OW = W / 2, OH = H / 2, PW = OW / 2;
    const long long total = planes * (OH * PW);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long pl = i / (OH * PW);
        const int r = int(i - pl * (OH * PW));
        const int oy = r / PW, ox = (r - oy * PW) * 2;
        const float* p = x + pl * (H * W);
        float best[2] = {-INFINITY, -INFINITY};
        int bc[2] = {0, 0};
        bool first[2] = {true, true};
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
            const int yy = 2 * oy - 1 + ky;
            if (yy < 0 || yy >= H) continue;
            const float4 v4 = __ldg(reinterpret_cast<const float4*>(p + yy * W + 2 * ox));
            const float left = ox > 0 ? __ldg(p + yy * W + 2 * ox - 1) : 0.0f;
            const float right = (2 * ox + 4 < W) ? __ldg(p + yy * W + 2 * ox + 4) : 0.0f;
            (void)right;
            // output ox: columns 2ox-1, 2ox, 2ox+1 = left, v4.x, v4.y;  output ox+1: columns 2ox+1, 2ox+2, 2ox+3 = v4.y, v4.z, v4.w
            const float c0[3] = {left, v4.x, v4.y};
            const float c1[3] = {v4.y, v4.z, v4.w};
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
                if (!(kx == 0 && ox == 0)) {                 // column -1 does not exist
                    const float v = c0[kx];
                    if (first[0] || v > best[0] || v != v) { best[0] = v; bc[0] = ky * 3 + kx; first[0] = false; }
                }
                {
                    const float v = c1[kx];
                    if (first[1] || v > best[1] || v != v) { best[1] = v; bc[1] = ky * 3 + kx; first[1] = false; }
                }
            }
        }
        const long long o = pl * (OH * OW) + oy * OW + ox;
        *reinterpret_cast<float2*>(y + o) = make_float2(best[0], best[1]);
        *reinterpret_cast<uchar2*>(code + o) = make_uchar2((unsigned char)bc[0], (unsigned char)bc[1]);
    }
}

// The gather of maxpool3s2_bwd_vec_kernel with one-byte window codes instead of int64 positions.
template <int LW, int LH>
__global__ void __launch_bounds__(256) maxpool3s2_bwd_code_vec_kernel(const float* __restrict__ dy, const uint8_t* __restrict__ code,
                                                                      float* __restrict__ dx, long long planes) {
    constexpr int W = 1 << LW, H = 1 << LH, OW = W / 2, OH = H / 2, QW = W / 4;
    const long long total = planes * (H * QW);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long pl = i / (H * QW);
        const int r = int(i - pl * (H * QW));
        const int y = r / QW, x0 = (r - y * QW) * 4;
        const int oy0 = y >> 1, oy1 = min((y + 1) >> 1, OH - 1);
        const int oxa = x0 >> 1;
        const float* g = dy + pl * (OH * OW);
        const uint8_t* cd = code + pl * (OH * OW);
        float o[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int oy = k == 0 ? oy0 : oy1;
            if (k == 1 && oy1 == oy0) break;
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                const int ox = oxa + j;
                if (ox >= OW) break;
                const int c = __ldg(cd + oy * OW + ox);
                const int ky = (c * 11) >> 5, kx = c - 3 * ky;              // c / 3, c % 3 for c in 0..8
                const bool row = (2 * oy - 1 + ky) == y;
                const int d = (2 * ox - 1 + kx) - x0;                       // which of my four pixels, if any
                const float v = __ldg(g + oy * OW + ox);
#pragma unroll
                for (int e = 0; e < 4; ++e) o[e] += (row && d == e) ? v : 0.0f;
            }
        }
        *reinterpret_cast<float4*>(dx + pl * (H * W) + y * W + x0) = make_float4(o[0], o[1], o[2], o[3]);
    }
}

// Scatter form of the same backward.  A pixel is covered by at most four windows -- (its row's earlier / later
// window) x (its column's earlier / later window) -- so each window drops its gradient into the slot
// [2 * (ky == 0) + (kx == 0)] of its arg-max pixel in a zeroed shared-memory tile, and every float4 of dx is the sum of
// the four slot planes in slot order: the same additions in the same order as the gather above (bit-identical),
// with ~45 instructions per thread instead of ~100 and every input read once, coalesced.
// 256 threads = 256 / (H * W / 4) planes per pass; thread = one window, then one float4 of dx.
template <int LW, int LH>
__global__ void __launch_bounds__(256) maxpool3s2_bwd_code_tile_kernel(const float* __restrict__ dy, const uint8_t* __restrict__ code,
                                                                       float* __restrict__ dx, long long planes) {
    constexpr int W = 1 << LW, H = 1 << LH, OW = W / 2, QW = W / 4;
    constexpr int QPP = H * QW;                       // float4 quads of dx per plane = windows per plane
    constexpr int PPB = 256 / QPP;                    // planes per block pass
    constexpr int PIX = PPB * H * W;                  // pixels per pass (1024 for every map size)
    __shared__ __align__(16) float slots[4][PIX];
    const int t = threadIdx.x, pl = t / QPP, r = t % QPP;
    const int oy = r / OW, ox = r % OW;               // as a window
    const int qy = r / QW, qx0 = (r % QW) * 4;        // as a quad of dx
    for (long long p0 = (long long)blockIdx.x * PPB; p0 < planes; p0 += (long long)gridDim.x * PPB) {
        const bool live = p0 + pl < planes;
        float v = 0.0f;
        int c = 4;
        if (live) {                                   // issue the loads before the zero fill
            v = __ldg(dy + (p0 + pl) * QPP + r);
            c = __ldg(code + (p0 + pl) * QPP + r);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) reinterpret_cast<float4*>(&slots[k][0])[t] = make_float4(0.f, 0.f, 0.f, 0.f);
        __syncthreads();
        if (live) {
            const int ky = (c * 11) >> 5, kx = c - 3 * ky;                      // c / 3, c % 3 for c in 0..8
            const int y = 2 * oy - 1 + ky, x = 2 * ox - 1 + kx;                 // the arg-max pixel: always inside the map
            slots[2 * (ky == 0) + (kx == 0)][(pl * H + y) * W + x] = v;
        }
        __syncthreads();
        if (live) {
            const int q = ((pl * H + qy) * W + qx0) / 4;
            const float4 s0 = reinterpret_cast<const float4*>(&slots[0][0])[q], s1 = reinterpret_cast<const float4*>(&slots[1][0])[q];
            const float4 s2 = reinterpret_cast<const float4*>(&slots[2][0])[q], s3 = reinterpret_cast<const float4*>(&slots[3][0])[q];
            float4 o;
            o.x = (((0.0f + s0.x) + s1.x) + s2.x) + s3.x;
            o.y = (((0.0f + s0.y) + s1.y) + s2.y) + s3.y;
            o.z = (((0.0f + s0.z) + s1.z) + s2.z) + s3.z;
            o.w = (((0.0f + s0.w) + s1.w) + s2.w) + s3.w;
            *reinterpret_cast<float4*>(dx + (p0 + pl) * (H * W) + qy * W + qx0) = o;
        }
        __syncthreads();
    }
}

int maxpool3s2_fwd(const float* x, float* y, uint8_t* code, long long planes, int h, int w, cudaStream_t st) {
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    const int oh = (h + 1) / 2, ow = (w + 1) / 2;
    if (h == w && (h == 4 || h == 8 || h == 16 || h == 32) && (reinterpret_cast<uintptr_t>(x) & 15) == 0 &&
        (reinterpret_cast<uintptr_t>(y) & 7) == 0 && (reinterpret_cast<uintptr_t>(code) & 1) == 0) {
        const long long tot2 = planes * oh * (ow / 2);
        const long long want2 = (tot2 + 255) / 256, cap2 = (long long)sm_count(dev) * 16;
        const int grid = int(want2 < cap2 ? want2 : cap2);
        if (h == 4) maxpool3s2_fwd_vec_kernel<2, 2><<<grid, 256, 0, st>>>(x, y, code, planes);
        if (h == 8) maxpool3s2_fwd_vec_kernel<3, 3><<<grid, 256, 0, st>>>(x, y, code, planes);
        if (h == 16) maxpool3s2_fwd_vec_kernel<4, 4><<<grid, 256, 0, st>>>(x, y, code, planes);
        if (h == 32) maxpool3s2_fwd_vec_kernel<5, 5><<<grid, 256, 0, st>>>(x, y, code, planes);
        MB_CUDA(cudaGetLastError());
        return MB_OK;
    }
    const long long total = planes * oh * ow;
    const long long want = (total + 255) / 256, cap = (long long)sm_count(dev) * 16;
    maxpool3s2_fwd_kernel<<<int(want < cap ? want : cap), 256, 0, st>>>(x, y, code, planes, h, w, oh, ow);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}

int maxpool3s2_bwd_code(const float* dy, const uint8_t* code, float* dx, long long planes, int h, int w, cudaStream_t st) {
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    const int oh = (h + 1) / 2, ow = (w + 1) / 2;
    if (h == w && (h == 4 || h == 8 || h == 16) && (reinterpret_cast<uintptr_t>(dx) & 15) == 0) {
        const int ppb = 256 / (h * (w / 4));
        const long long want = (planes + ppb - 1) / ppb, cap = (long long)sm_count(dev) * 8;
        const int grid = int(want < cap ? want : cap);
        if (h == 4) maxpool3s2_bwd_code_tile_kernel<2, 2><<<grid, 256, 0, st>>>(dy, code, dx, planes);
        if (h == 8) maxpool3s2_bwd_code_tile_kernel<3, 3><<<grid, 256, 0, st>>>(dy, code, dx, planes);
        if (h == 16) maxpool3s2_bwd_code_tile_kernel<4, 4><<<grid, 256, 0, st>>>(dy, code, dx, planes);
        MB_CUDA(cudaGetLastError());
        return MB_OK;
    }
    if (h == w && (h == 4 || h == 8 || h == 16 || h == 32) && (reinterpret_cast<uintptr_t>(dx) & 15) == 0) {
        const long long tot4 = planes * h * (w / 4);
        const long long want4 = (tot4 + 255) / 256, cap4 = (long long)sm_count(dev) * 16;
        const int grid = int(want4 < cap4 ? want4 : cap4);
        if (h == 4) maxpool3s2_bwd_code_vec_kernel<2, 2><<<grid, 256, 0, st>>>(dy, code, dx, planes);
        if (h == 8) maxpool3s2_bwd_code_vec_kernel<3, 3><<<grid, 256, 0, st>>>(dy, code, dx, planes);
        if (h == 16) maxpool3s2_bwd_code_vec_kernel<4, 4><<<grid, 256, 0, st>>>(dy, code, dx, planes);
        if (h == 32) maxpool3s2_bwd_code_vec_kernel<5, 5><<<grid, 256, 0, st>>>(dy, code, dx, planes);
        MB_CUDA(cudaGetLastError());
        return MB_OK;
    }
    const long long total = planes * h * w;
    const long long want = (total + 255) / 256, cap = (long long)sm_count(dev) * 16;
    maxpool3s2_bwd_code_kernel<<<int(want < cap ? want : cap), 256, 0, st>>>(dy, code, dx, planes, h, w, oh, ow);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}

int maxpool3s2_bwd(const float* dy, const long long* idx, float* dx, long long planes, int h, int w, cudaStream_t st) {
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    if (h == w && (h == 4 || h == 8 || h == 16 || h == 32) && (reinterpret_cast<uintptr_t>(dx) & 15) == 0) {
        const long long total = planes * h * (w / 4);
        const long long want = (total + 255) / 256, cap = (long long)sm_count(dev) * 16;
        const int grid = int(want < cap ? want : cap);
        if (h == 4) maxpool3s2_bwd_vec_kernel<2, 2><<<grid, 256, 0, st>>>(dy, idx, dx, planes);
        if (h == 8) maxpool3s2_bwd_vec_kernel<3, 3><<<grid, 256, 0, st>>>(dy, idx, dx, planes);
        if (h == 16) maxpool3s2_bwd_vec_kernel<4, 4><<<grid, 256, 0, st>>>(dy, idx, dx, planes);
        if (h == 32) maxpool3s2_bwd_vec_kernel<5, 5><<<grid, 256, 0, st>>>(dy, idx, dx, planes);
        MB_CUDA(cudaGetLastError());
        return MB_OK;
    }
    const int oh = (h + 1) / 2, ow = (w + 1) / 2;
    const long long total = planes * h * w;
    const long long want = (total + 255) / 256;
    const long long cap = (long long)sm_count(dev) * 16;
    maxpool3s2_bwd_kernel<<<int(want < cap ? want : cap), 256, 0, st>>>(dy, idx, dx, planes, h, w, oh, ow);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}

}  // namespace mb

extern "C" int mb_conv3x3_fwd_supported(int32_t cin, int32_t cout, int32_t h, int32_t w) {
    return mb::conv3x3_fwd_supported(cin, cout, h, w);
}

extern "C" int mb_conv3x3_fwd(const float* in, const float* wk, const float* bias, float* out, int64_t n_frames,
                              int32_t cin, int32_t cout, int32_t h, int32_t w, int32_t wmode, void* stream) {
    if (n_frames < 0 || wmode < 0 || wmode > 2) {
        mb::set_error("mb_conv3x3_fwd: n_frames=%lld wmode=%d out of range", (long long)n_frames, wmode);
        return MB_ERR_INVALID;
    }
    if (n_frames == 0) return MB_OK;
    if (!in || !wk || !out) {
        mb::set_error("mb_conv3x3_fwd: null pointer");
        return MB_ERR_INVALID;
    }
    if ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) {
        mb::set_error("mb_conv3x3_fwd: in and out must be 16-byte aligned");
        return MB_ERR_INVALID;
    }
    return mb::conv3x3_fwd(in, wk, bias, out, n_frames, cin, cout, h, w, wmode, 0, nullptr, nullptr,
                           static_cast<cudaStream_t>(stream));
}

extern "C" int mb_conv3x3_fwd_fused(const float* in, const float* wk, const float* bias, float* out, int64_t n_frames,
                                    int32_t cin, int32_t cout, int32_t h, int32_t w, int32_t wmode, int32_t relu_in,
                                    const float* mask_src, const float* add_src, void* stream) {
    if (n_frames < 0 || wmode < 0 || wmode > 2) {
        mb::set_error("mb_conv3x3_fwd_fused: n_frames=%lld wmode=%d out of range", (long long)n_frames, wmode);
        return MB_ERR_INVALID;
    }
    if (n_frames == 0) return MB_OK;
    if (!in || !wk || !out || ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15)) {
        mb::set_error("mb_conv3x3_fwd_fused: null pointer, or in / out not 16-byte aligned");
        return MB_ERR_INVALID;
    }
    return mb::conv3x3_fwd(in, wk, bias, out, n_frames, cin, cout, h, w, wmode, relu_in, mask_src, add_src,
                           static_cast<cudaStream_t>(stream));
}

extern "C" int mb_maxpool3s2_bwd(const float* dy, const int64_t* idx, float* dx, int64_t planes, int32_t h, int32_t w,
                                 void* stream) {
    if (planes < 0 || h <= 0 || w <= 0) {
        mb::set_error("mb_maxpool3s2_bwd: planes=%lld %dx%d out of range", (long long)planes, h, w);
        return MB_ERR_INVALID;
    }
    if (planes == 0) return MB_OK;
    if (!dy || !idx || !dx) {
        mb::set_error("mb_maxpool3s2_bwd: null pointer");
        return MB_ERR_INVALID;
    }
    return mb::maxpool3s2_bwd(dy, reinterpret_cast<const long long*>(idx), dx, planes, h, w,
                              static_cast<cudaStream_t>(stream));
}

extern "C" int mb_conv3x3_wgrad_blocks(int64_t n_frames, int32_t cin, int32_t cout, int32_t h, int32_t w) {
    if (n_frames < 0 || cin <= 0 || cout <= 0 || h <= 0 || w <= 0) {
        mb::set_error("mb_conv3x3_wgrad_blocks: n_frames=%lld cin=%d cout=%d %dx%d out of range", (long long)n_frames,
                      cin, cout, h, w);
        return -1;
    }
    int grid = 0;
    if (mb::conv3x3_wgrad_blocks(n_frames, cin, cout, h, w, &grid) != MB_OK) return -1;
    return grid;
}

extern "C" int mb_conv3x3_wgrad(const float* x, const float* dy, float* partials, int32_t n_blocks, int64_t n_frames,
                                int32_t cin, int32_t cout, int32_t h, int32_t w, void* stream) {
    if (n_frames < 0 || n_blocks <= 0) {
        mb::set_error("mb_conv3x3_wgrad: n_frames=%lld n_blocks=%d out of range", (long long)n_frames, n_blocks);
        return MB_ERR_INVALID;
    }
    if (!x || !dy || !partials) {
        mb::set_error("mb_conv3x3_wgrad: null pointer");
        return MB_ERR_INVALID;
    }
    if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dy)) & 15) {
        mb::set_error("mb_conv3x3_wgrad: x and dy must be 16-byte aligned");
        return MB_ERR_INVALID;
    }
    return mb::conv3x3_wgrad(x, dy, partials, n_blocks, n_frames, cin, cout, h, w, 0, nullptr, static_cast<cudaStream_t>(stream));
}

extern "C" int mb_conv3x3_wgrad_relu(const float* x, const float* dy, float* partials, int32_t n_blocks, int64_t n_frames,
                                     int32_t cin, int32_t cout, int32_t h, int32_t w, int32_t relu_x, float* accumulate_out,
                                     void* stream) {
    if (n_frames < 0 || n_blocks <= 0 || !x || !dy || (!partials && !accumulate_out) ||
        ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dy)) & 15)) {
        mb::set_error("mb_conv3x3_wgrad_relu: bad arguments (null / misaligned pointer, n_frames=%lld, n_blocks=%d)",
                      (long long)n_frames, n_blocks);
        return MB_ERR_INVALID;
    }
    return mb::conv3x3_wgrad(x, dy, partials, n_blocks, n_frames, cin, cout, h, w, relu_x, accumulate_out,
                             static_cast<cudaStream_t>(stream));
}

extern "C" int mb_maxpool3s2_fwd(const float* x, float* y, uint8_t* code, int64_t planes, int32_t h, int32_t w, void* stream) {
    if (planes < 0 || h <= 0 || w <= 0) {
        mb::set_error("mb_maxpool3s2_fwd: planes=%lld %dx%d out of range", (long long)planes, h, w);
        return MB_ERR_INVALID;
    }
    if (planes == 0) return MB_OK;
    if (!x || !y || !code) {
        mb::set_error("mb_maxpool3s2_fwd: null pointer");
        return MB_ERR_INVALID;
    }
    return mb::maxpool3s2_fwd(x, y, code, planes, h, w, static_cast<cudaStream_t>(stream));
}

extern "C" int mb_maxpool3s2_bwd_code(const float* dy, const uint8_t* code, float* dx, int64_t planes, int32_t h, int32_t w,
                                      void* stream) {
    if (planes < 0 || h <= 0 || w <= 0) {
        mb::set_error("mb_maxpool3s2_bwd_code: planes=%lld %dx%d out of range", (long long)planes, h, w);
        return MB_ERR_INVALID;
    }
    if (planes == 0) return MB_OK;
    if (!dy || !code || !dx) {
        mb::set_error("mb_maxpool3s2_bwd_code: null pointer");
        return MB_ERR_INVALID;
    }
    return mb::maxpool3s2_bwd_code(dy, code, dx, planes, h, w, static_cast<cudaStream_t>(stream));
}
