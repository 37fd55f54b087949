// The encoder's 3x3 convolutions on the tensor cores (tcgen05.mma kind::tf32, accumulators in TMEM), for sm_100a.
//
// Reference: model.py:56-109 -- ResidualBlock.conv0 / conv1 and ConvSequence.conv (Conv2d(c, c', 3, padding=1)) on
// 16- and 32-channel maps of 8x8, 4x4 and 2x2 pixels, and their data gradients in the backward pass (the same
// convolution with the transposed, flipped kernel).  Same contract as mb_conv3x3_fwd_fused (csrc/convgrad.cu), which
// dispatches here for the shapes listed at the bottom.
//
// Why not an im2col GEMM.  As a GEMM over pixels a 3x3 convolution has N = cout = 16 or 32, and tcgen05.mma does not
// go below ~44 clocks per M128 x K8 instruction however small N is (it reads the 4 KB A slab from shared memory every
// time): measured 199 / 399 TFLOP/s at N = 16 / 32 against 1117 at N >= 128 (profiles/r02_micro_umma_n.txt), which
// after the 3 passes of the split-TF32 product leaves less than the fp32 FMA pipe gives.  So the kernel makes N wide
// instead: ONE product per pixel tile against all nine taps at once,
//     Y[p, tap, co] = sum_ci X[p, ci] * W[co, ci, tap]            (M = 128 pixels, N = 9 * 16 = 144, K = cin)
// and the convolution is the shift-and-add of Y that follows:  out[q, co] = sum_tap Y[q + d(tap), tap, co]  over the
// taps whose source pixel q + d(tap) lies inside the frame (zero padding drops the rest).  No im2col, no halo rows in
// M, every pixel is read once.  A tile is 128 pixels = whole frames (2, 8 or 32 of them), so no shift leaves the tile.
//
// Precision: X and W are split into TF32 halves (x = hi + lo); the cross terms lo*hi and hi*lo are accumulated first
// and hi*hi last, all into one fp32 TMEM accumulator (K <= 32: at most four additions at full magnitude; the tensor
// core truncates on accumulation, csrc/gemm.cu).  Result: fp32-grade, ~1e-6 relative; tests hold it to 1e-5 vs fp64.
//
// CTA = 17 warps, persistent over tiles (one CTA per SM):
//   warps 12-15 stage A: the tile's NCHW floats arrive by cp.async, 2 (cin 32) or 5 (cin 16) tiles ahead -- one tile in
//              flight per SM is 8-16 KB, which latency holds to ~1.7 TB/s; then a thread per pixel applies the optional
//              ReLU, splits, and stores four channels per 16-byte store into the K-major, 128-byte-swizzled operand
//              tiles the MMA descriptors expect (conflict-free: eight consecutive pixels hit the eight chunk slots);
//   warp 16    one lane issues the MMAs: per 16-channel output group ("job") 3 passes x cin/8 k-steps of
//              M128 x N144 x K8 into one of three TMEM accumulator slots, then commits;
//   warps 0-11 three epilogue groups, one per accumulator slot (the epilogue is a long dependent chain per thread, so
//              it is the groups' number that sets the rate), thread = pixel (its TMEM lane): x- and y-shifts by warp
//              shuffles as the 144 columns are read (8x8 maps: the two frame rows that meet at a warp boundary go
//              through shared memory), then bias / ReLU-mask / residual add and coalesced NCHW stores.
// The weights (both halves, all taps) stay in shared memory for the CTA's lifetime.
#include <atomic>

#include "common.cuh"
#include "tc.cuh"

#include "../../include/mbeast.h"

namespace mb {

int sm_count(int device);

namespace {

constexpr int kTcThreads = 544;
constexpr int kTcStagers = 128;                  // staging threads (warps 12-15)
constexpr int kTcGroups = 3;                     // epilogue groups = accumulator slots
constexpr int kTcM = 128;                        // pixels per tile = TMEM lanes
constexpr int kTcJobN = 144;                     // 9 taps x 16 output channels per accumulator
constexpr int kTcSlotCols = 160;                 // TMEM columns from one accumulator slot to the next
constexpr int kTcSlots = 3;
constexpr int kTcStages = 2;
constexpr int kTcTileA = kTcM * 128;             // bytes of an A operand half: 128 rows x 32 fp32
constexpr int kTcXchPitch = 17;                  // floats per pixel in the warp-boundary exchange (16 + 1: no bank conflicts)
constexpr int kTcXchWarp = 2 * 8 * kTcXchPitch;    // floats a warp publishes: its last row's z0 and its first row's z2
constexpr int kTcXchBytes = kTcGroups * 2 * 4 * kTcXchWarp * 4;      // [group][job parity][quarter]
constexpr int kTcJobB = kTcJobN * 128;           // bytes of a job's weight rows (18 swizzle atoms)

template <int CIN_, int COUT_, int H_>
struct TcCfg {
    static constexpr int CIN = CIN_, COUT = COUT_, H = H_, W = H_, HW = H_ * H_;
    static constexpr int FPT = kTcM / HW;                          // frames per tile
    static constexpr int JOBS = COUT / 16;
    static constexpr int KSTEPS = CIN / 8;
    static constexpr int BBYTES = JOBS * kTcJobB;                  // one half (hi or lo) of the weights
    static constexpr int LOADS = 32 * CIN / kTcStagers;            // float4 pieces per staging thread and tile
    static constexpr int RAW_SLOTS = CIN == 32 ? 3 : 6;               // raw input tiles in the cp.async ring
    static constexpr int RAW_BYTES = RAW_SLOTS * (kTcM * CIN + kTcM) * 4;  // frames padded by HW floats each: + 128 floats per tile
    static constexpr int AHEAD = RAW_SLOTS - 1;                       // tiles whose loads are in flight
    static constexpr int SMEM = 1024 + 2 * BBYTES + kTcStages * 2 * kTcTileA + kTcXchBytes + RAW_BYTES;
    static_assert(kTcM % HW == 0 && HW >= 4 && (CIN == 16 || CIN == 32) && (COUT == 16 || COUT == 32), "shape");
};

// byte offset of element (row, k) in a K-major tile of 128-byte rows with the 128-byte swizzle (what TMA writes with
// CU_TENSOR_MAP_SWIZZLE_128B): 8-row atoms of 1024 B, the 16-byte chunk index XORed with the row's index in the atom
__device__ __forceinline__ uint32_t sw128_off(int row, int k) {
    return uint32_t((row >> 3) * 1024 + (row & 7) * 128 + ((((k >> 2) ^ (row & 7))) << 4) + (k & 3) * 4);
}

__device__ __forceinline__ void tc_cp_async16(void* dst_smem, const void* src, bool valid) {
    const uint32_t bytes = valid ? 16u : 0u;                 // 0 source bytes: the destination is zero-filled
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes) : "memory");
}

__device__ __forceinline__ void split_tf32_pair(float v, float& hi, float& lo) {
    uint32_t h, l;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(v));
    const float r = v - __uint_as_float(h);
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(l) : "f"(r));
    hi = __uint_as_float(h);
    lo = __uint_as_float(l);
}

template <class C>
__global__ void __launch_bounds__(kTcThreads, 1) conv3x3_tc_kernel(const float* __restrict__ in, const float* __restrict__ wk,
                                                                    const float* __restrict__ bias, float* __restrict__ out,
                                                                    long long n_frames, int wmode, int relu_in,
                                                                    const float* __restrict__ mask_src,
                                                                    const float* __restrict__ add_src, int debug) {
    constexpr int CIN = C::CIN, COUT = C::COUT, H = C::H, W = C::W, HW = C::HW;
    extern __shared__ uint8_t tc_smem_raw[];
    __shared__ __align__(8) uint64_t a_full[kTcStages], a_empty[kTcStages], acc_full[kTcSlots], acc_empty[kTcSlots];
    __shared__ uint32_t tmem_slot;
    __shared__ float s_bias[32];
    uint8_t* smem = tc_smem_raw + ((1024u - (smem_u32(tc_smem_raw) & 1023u)) & 1023u);
    uint8_t* b_hi = smem;
    uint8_t* b_lo = smem + C::BBYTES;
    uint8_t* a_base = smem + 2 * C::BBYTES;
    float* xch = reinterpret_cast<float*>(a_base + kTcStages * 2 * kTcTileA);
    uint8_t* raw_base = reinterpret_cast<uint8_t*>(xch) + kTcXchBytes;                 // 16-byte aligned (kTcXchBytes = 26112)

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < kTcStages; ++s) {
            mbar_init(&a_full[s], kTcStagers);               // every staging thread arrives
            mbar_init(&a_empty[s], 1);                       // tcgen05.commit
        }
        for (int s = 0; s < kTcSlots; ++s) {
            mbar_init(&acc_full[s], 1);                      // tcgen05.commit
            mbar_init(&acc_empty[s], 4);                     // one arrival per epilogue warp
        }
        fence_mbar_init();
    }
    if (warp == 16) tmem_alloc(&tmem_slot, 512);
    // weights -> B rows n = (job * 9 + tap) * 16 + c, K-major (k = input channel), split; wmode as mb_conv3x3_fwd:
    // 0 = [cin][cout][9], 1 = a torch Conv2d weight [cout][cin][3][3], 2 = the torch weight [cin][cout][3][3] of the
    // convolution whose data gradient this is, taps flipped
    // (all the loads first: one round trip to L2 instead of one per element)
    constexpr int kWPer = (9 * COUT * CIN + kTcThreads - 1) / kTcThreads;
    float wreg[kWPer];
#pragma unroll
    for (int r = 0; r < kWPer; ++r) {
        const int i = tid + r * kTcThreads, ci = i % CIN, n = i / CIN;
        const int c16 = n % 16, tap = (n / 16) % 9, co = (n / kTcJobN) * 16 + c16;
        wreg[r] = 0.0f;
        if (i < 9 * COUT * CIN)
            wreg[r] = __ldg(wk + (wmode == 0 ? (ci * COUT + co) * 9 + tap : wmode == 1 ? (co * CIN + ci) * 9 + tap : (ci * COUT + co) * 9 + (8 - tap)));
    }
#pragma unroll
    for (int r = 0; r < kWPer; ++r) {
        const int i = tid + r * kTcThreads, ci = i % CIN, n = i / CIN;
        if (i < 9 * COUT * CIN) {
            float hi, lo;
            split_tf32_pair(wreg[r], hi, lo);
            *reinterpret_cast<float*>(b_hi + sw128_off(n, ci)) = hi;
            *reinterpret_cast<float*>(b_lo + sw128_off(n, ci)) = lo;
        }
    }
    if (tid < COUT) s_bias[tid] = bias ? bias[tid] : 0.0f;
    fence_proxy_async_smem();                                // the MMAs read these generic-proxy writes
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;
    const long long n_tiles = (n_frames + C::FPT - 1) / C::FPT;

    if (warp >= 12 && warp < 16) {
        // ===================== staging: NCHW floats -> swizzled K-major TF32 halves =====================
        // thread = pixel of the tile.  The raw ring holds tiles as they lie in memory ([frame][channel][pixel]), each
        // frame followed by HW floats of padding so that the frames a warp straddles (4x4, 2x2 maps) fall on other banks.
        const int t = tid - 384;
        constexpr int kFrameF = CIN * HW + HW;                                    // floats per frame in the ring
        auto issue = [&](long long tile, int slot) {
            const float* src = in + tile * (long long)(kTcM * CIN);
            float* dst = reinterpret_cast<float*>(raw_base) + slot * (C::FPT * kFrameF);
#pragma unroll
            for (int r = 0; r < C::LOADS; ++r) {
                const int i = t + kTcStagers * r;                                 // 16-byte piece of the tile, in memory order
                const int fr = i / (CIN * HW / 4), rem = i % (CIN * HW / 4);
                const bool valid = tile * C::FPT + fr < n_frames;                 // frames past the end read as zeros
                tc_cp_async16(dst + fr * kFrameF + 4 * rem, valid ? src + 4 * i : in, valid);
            }
        };
#pragma unroll
        for (int d = 0; d < C::AHEAD; ++d) {
            const long long tile = blockIdx.x + (long long)d * gridDim.x;
            if (tile < n_tiles) issue(tile, d);
            asm volatile("cp.async.commit_group;" ::: "memory");
        }
        const int sfr = t / HW, spix = t % HW, r7 = t & 7;
        const uint32_t row_off = uint32_t((t >> 3) * 1024 + r7 * 128);
        uint32_t it = 0;
        for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
            const int s = it % kTcStages;
            asm volatile("cp.async.wait_group %0;" ::"n"(C::AHEAD - 1) : "memory");       // my pieces of this tile have landed
            asm volatile("bar.sync 4, 128;" ::: "memory");   // ... everyone's have, and everyone is done with the previous tile's slot
            const long long ahead = tile + (long long)C::AHEAD * gridDim.x;
            if (ahead < n_tiles) issue(ahead, (it + C::AHEAD) % C::RAW_SLOTS);
            asm volatile("cp.async.commit_group;" ::: "memory");
            mbar_wait(&a_empty[s], ((it / kTcStages) & 1) ^ 1, nullptr);          // passes at once the first time round
            if (debug & 2) {                                                      // timing experiment: no conversion
                mbar_arrive(&a_full[s]);
                continue;
            }
            uint8_t* ah = a_base + s * 2 * kTcTileA + row_off;
            const float* px = reinterpret_cast<const float*>(raw_base) + (it % C::RAW_SLOTS) * (C::FPT * kFrameF) + sfr * kFrameF + spix;
#pragma unroll
            for (int cg = 0; cg < CIN / 4; ++cg) {
                uint32_t hi[4], lo[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    float xv = px[(4 * cg + k) * HW];
                    if (relu_in) xv = fmaxf(xv, 0.0f);
                    // nearest TF32 (ties away from zero, as cvt.rna.tf32.f32), twice: x = hi + lo up to 2^-22 |x|
                    hi[k] = (__float_as_uint(xv) + 0x1000u) & 0xffffe000u;
                    lo[k] = (__float_as_uint(xv - __uint_as_float(hi[k])) + 0x1000u) & 0xffffe000u;
                }
                const uint32_t off = uint32_t((cg ^ r7) << 4);                   // the 128-byte swizzle: 16-byte chunk ^ row-in-atom
                *reinterpret_cast<uint4*>(ah + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                *reinterpret_cast<uint4*>(ah + kTcTileA + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            }
            fence_proxy_async_smem();
            mbar_arrive(&a_full[s]);
        }
    } else if (warp == 16) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            constexpr uint32_t idesc = tf32_idesc(kTcM, kTcJobN);
            uint32_t it = 0, job = 0;
            for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
                const int s = it % kTcStages;
                mbar_wait(&a_full[s], (it / kTcStages) & 1, nullptr);
                tc_fence_after();
                const uint32_t ah = smem_u32(a_base + s * 2 * kTcTileA);
                const uint64_t a_hi = make_kmajor_sw128_desc(ah), a_lo = make_kmajor_sw128_desc(ah + kTcTileA);
#pragma unroll 1
                for (int h = 0; h < C::JOBS; ++h, ++job) {
                    const int slot = job % kTcSlots;
                    mbar_wait(&acc_empty[slot], ((job / kTcSlots) & 1) ^ 1, nullptr);
                    tc_fence_after();
                    const uint32_t d = tmem_base + uint32_t(slot * kTcSlotCols);
                    const uint64_t bh = make_kmajor_sw128_desc(smem_u32(b_hi + h * kTcJobB));
                    const uint64_t bl = make_kmajor_sw128_desc(smem_u32(b_lo + h * kTcJobB));
                    if (debug & 4) {                         // timing experiment: no MMAs
                        tc_commit(&acc_full[slot]);
                        continue;
                    }
                    // the two small cross terms first, the hi * hi bulk last (see the header)
#pragma unroll
                    for (int k = 0; k < C::KSTEPS; ++k) mma_tf32(d, a_lo + uint64_t(k * 2), bh + uint64_t(k * 2), idesc, k != 0);
#pragma unroll
                    for (int k = 0; k < C::KSTEPS; ++k) mma_tf32(d, a_hi + uint64_t(k * 2), bl + uint64_t(k * 2), idesc, 1);
#pragma unroll
                    for (int k = 0; k < C::KSTEPS; ++k) mma_tf32(d, a_hi + uint64_t(k * 2), bh + uint64_t(k * 2), idesc, 1);
                    tc_commit(&acc_full[slot]);
                }
                tc_commit(&a_empty[s]);                      // the stage is free once every MMA above has read it
            }
        }
    } else {
        // ===================== epilogue: shift-and-add, fused tails, stores =====================
        const int group = warp >> 2, quarter = warp & 3;     // group g drains accumulator slot g: jobs g, g + 3, ...
        const int p = quarter * 32 + lane;                   // this thread's pixel of the tile = its TMEM lane
        const int x = p % W, y = (p / W) % H, fr = p / HW, pix = p % HW;
        // neighbours as 0 / 1 factors: left / right inside the pixel row; the rows above / below when they sit in this
        // warp (always, unless the map is 8x8: there a warp holds four rows of a frame and rows 3 | 4 meet across warps)
        const float mxl = x >= 1 ? 1.0f : 0.0f, mxr = x <= W - 2 ? 1.0f : 0.0f;
        const float myu = (y >= 1 && lane >= W) ? 1.0f : 0.0f, myd = (y <= H - 2 && lane < 32 - W) ? 1.0f : 0.0f;
        constexpr bool kExchange = H == 8;
        const uint32_t taddr = tmem_base + (uint32_t(quarter * 32) << 16) + uint32_t(group * kTcSlotCols);
        uint32_t job = 0, mine = 0;
        for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const long long f = tile * C::FPT + fr;
            const bool live = f < n_frames;
#pragma unroll 1
            for (int h = 0; h < C::JOBS; ++h, ++job) {
                if (int(job % kTcSlots) != group) continue;
                mbar_wait(&acc_full[group], mine & 1, nullptr);
                tc_fence_after();
                if (debug & 1) {                             // timing experiment: no epilogue work
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&acc_empty[group]);
                    ++mine;
                    continue;
                }
                float acc[16], z0[16], z2[16];
                // 12 units of (tap row ky, 4 channels): the TMEM reads of unit u + 1 are in flight while unit u is shifted
                uint32_t ybuf[2][12];
                auto unit_issue = [&](int u, uint32_t (&dst)[12]) {
                    const uint32_t col = taddr + uint32_t((u / 4) * 48 + (u % 4) * 4);          // tap (ky, 0), channels 4 * (u % 4) ..
                    tmem_ld4x3_issue(col, col + 16u, col + 32u, dst);
                };
                unit_issue(0, ybuf[0]);
#pragma unroll
                for (int u = 0; u < 12; ++u) {
                    uint32_t (&cur)[12] = ybuf[u & 1];
                    tmem_ld_wait12(cur);
                    if (u + 1 < 12) unit_issue(u + 1, ybuf[(u + 1) & 1]);
                    const int ky = u / 4, c0 = (u % 4) * 4;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        // Z[q][ky] = Y[q-1][ky,0] + Y[q][ky,1] + Y[q+1][ky,2] within the pixel row
                        const float l = __shfl_up_sync(0xffffffffu, __uint_as_float(cur[i]), 1);
                        const float r = __shfl_down_sync(0xffffffffu, __uint_as_float(cur[8 + i]), 1);
                        const float z = fmaf(mxl, l, fmaf(mxr, r, __uint_as_float(cur[4 + i])));
                        if (ky == 0) z0[c0 + i] = z;
                        if (ky == 1) acc[c0 + i] = z;
                        if (ky == 2) z2[c0 + i] = z;
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&acc_empty[group]);                    // the slot can take another job
                // out[q] = Z[q][1] + Z[q-W][0] + Z[q+W][2] within the frame
#pragma unroll
                for (int c = 0; c < 16; ++c) {
                    const float up = __shfl_up_sync(0xffffffffu, z0[c], W), dn = __shfl_down_sync(0xffffffffu, z2[c], W);
                    acc[c] = fmaf(myu, up, fmaf(myd, dn, acc[c]));
                }
                if constexpr (kExchange) {
                    float* xb = xch + ((group * 2 + int(mine & 1)) * 4 + quarter) * kTcXchWarp;
                    if (lane >= 24) {
#pragma unroll
                        for (int c = 0; c < 16; ++c) xb[(lane - 24) * kTcXchPitch + c] = z0[c];          // my last row, for the warp below
                    }
                    if (lane < 8) {
#pragma unroll
                        for (int c = 0; c < 16; ++c) xb[(8 + lane) * kTcXchPitch + c] = z2[c];           // my first row, for the warp above
                    }
                    asm volatile("bar.sync %0, 128;" ::"r"(group + 1) : "memory");
                    if (lane < 8 && (quarter & 1)) {                              // row 4 of a frame: row 3 is the previous warp's last
                        const float* src = xb - kTcXchWarp + lane * kTcXchPitch;
#pragma unroll
                        for (int c = 0; c < 16; ++c) acc[c] += src[c];
                    }
                    if (lane >= 24 && !(quarter & 1)) {                           // row 3 of a frame: row 4 is the next warp's first
                        const float* src = xb + kTcXchWarp + (8 + lane - 24) * kTcXchPitch;
#pragma unroll
                        for (int c = 0; c < 16; ++c) acc[c] += src[c];
                    }
                }
                ++mine;
                if (live) {
                    float* o = out + (f * COUT + h * 16) * HW + pix;
                    if (mask_src || add_src) {
                        const long long oi = (f * COUT + h * 16) * HW + pix;
#pragma unroll
                        for (int c = 0; c < 16; ++c) {
                            float v = acc[c] + s_bias[h * 16 + c];
                            if (mask_src) v = __ldg(mask_src + oi + c * HW) > 0.0f ? v : 0.0f;
                            if (add_src) v += __ldg(add_src + oi + c * HW);
                            o[c * HW] = v;
                        }
                    } else {
#pragma unroll
                        for (int c = 0; c < 16; ++c) o[c * HW] = acc[c] + s_bias[h * 16 + c];
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 16) tmem_dealloc(tmem_base, 512);
}

std::atomic<int> g_tc_debug{0};          // mb_set_tuning("conv_tc_debug", bits): timing experiments that skip a role's work (wrong results)

template <class C>
int tc_launch(const float* in, const float* wk, const float* bias, float* out, long long n_frames, int wmode, int relu_in,
              const float* mask_src, const float* add_src, cudaStream_t st) {
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    static bool attr_set[64] = {};
    bool& done = attr_set[(dev >= 0 && dev < 64) ? dev : 0];
    if (!done) {
        MB_CUDA(cudaFuncSetAttribute(conv3x3_tc_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
        done = true;
    }
    const long long tiles = (n_frames + C::FPT - 1) / C::FPT, cap = sm_count(dev);
    conv3x3_tc_kernel<C><<<int(tiles < cap ? tiles : cap), kTcThreads, C::SMEM, st>>>(in, wk, bias, out, n_frames, wmode, relu_in,
                                                                                    mask_src, add_src, g_tc_debug);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}

// (cin, cout, h): residual blocks and ConvSequence convolutions after the first pool, and their data gradients
#define MB_TC_SHAPES(X) \
    X(16, 16, 8) X(16, 16, 4) X(16, 32, 8) X(16, 32, 4) X(32, 16, 8) X(32, 16, 4) X(32, 32, 8) X(32, 32, 4) X(32, 32, 2)

}  // namespace

void conv_tc_set_debug(int bits) { g_tc_debug = bits; }

int conv3x3_tc_supported(int cin, int cout, int h, int w) {
    if (h != w) return 0;
#define X(CI, CO, HH) \
    if (cin == CI && cout == CO && h == HH) return 1;
    MB_TC_SHAPES(X)
#undef X
    return 0;
}

int conv3x3_tc(const float* in, const float* wk, const float* bias, float* out, long long n_frames, int cin, int cout, int h,
               int w, int wmode, int relu_in, const float* mask_src, const float* add_src, cudaStream_t st) {
    if (h == w) {
#define X(CI, CO, HH)                           \
    if (cin == CI && cout == CO && h == HH)     \
        return tc_launch<TcCfg<CI, CO, HH>>(in, wk, bias, out, n_frames, wmode, relu_in, mask_src, add_src, st);
        MB_TC_SHAPES(X)
#undef X
    }
    set_error("conv3x3_tc: shape cin=%d cout=%d %dx%d is not covered", cin, cout, h, w);
    return MB_ERR_UNSUPPORTED;
}

}  // namespace mb
