// Actor Linear + policy head as ONE tensor-core kernel, for sm_100a.
//
// Reference: model.py:167 (logits = actor(hidden)) followed by model.py:168-214 (split into 7*C masked
// categoricals, log_prob of the batch's actions, masked entropy, sums over the frame).  Run as two kernels
// the [N, 78*C] fp32 logits make a round trip through HBM (1.66 GB written and read back at 16x16, T=80,
// 256 rows) although the learner never looks at them (libs/utils.py:237-329 uses logprobs / entropy /
// baseline only).  Here the epilogue of the tcgen05 GEMM evaluates the head on the accumulator rows it
// already holds: logits exist only in tensor memory and registers (storing them stays possible: logits_out).
//
// Pieces:
//   pack_cells         mask u8 [N,78C] + action i64 [N,7C]  ->  cell records [C][N] of 16 B: 78 mask bits and
//                      seven 6-bit action indices.  Cell-major so that the 32 lanes of an epilogue warp (32
//                      consecutive frames, one cell) read one contiguous 512-B run.
//   pack_actor_weight  W [78C,K], b [78C] -> TF32 hi / lo copies with every cell's 78 rows padded to 80
//                      ([80C,K]; rows 78, 79 zero), so that an N tile of 240 columns is exactly three cells.
//   actor_head_fwd     C tile 128 frames x 240 columns = 3 cells; TMA -> 2-stage ring -> tcgen05.mma
//                      kind::tf32, three products of hi/lo-split operands (fp32-level accuracy, see gemm.cu)
//                      into two TMEM accumulators; 12 epilogue warps, one thread per (frame, cell): 80 + 80
//                      columns out of TMEM, bias, mask, the seven segment soft-maxes, log-prob of the taken
//                      actions and entropy -> cell_out [C][N] float2.
//   frame_sum          per frame, the C cell results added in cell order (deterministic) -> logp, ent [N].
#include <atomic>

#include "common.cuh"
#include "tc.cuh"

#include <cuda.h>

#include <mutex>

#include "../../include/mbeast.h"

namespace mb {

int sm_count(int device);

namespace {

constexpr int kCellPad = 80;                         // logit rows per cell in the padded weight copy

// ---- cell record -----------------------------------------------------------------------------
// x = mask bits 0..31, y = 32..63, z[0:14) = 64..77, z[14:32) = actions 0..2 (6 bits each), w[0:24) = actions 3..6
__device__ __forceinline__ bool meta_bit(const uint4& m, int i) {      // i is a compile-time constant at every call
    return i < 32 ? ((m.x >> i) & 1u) : i < 64 ? ((m.y >> (i - 32)) & 1u) : ((m.z >> (i - 64)) & 1u);
}
__device__ __forceinline__ int meta_action(const uint4& m, int k) {
    return k < 3 ? int((m.z >> (14 + 6 * k)) & 63u) : int((m.w >> (6 * (k - 3))) & 63u);
}
__device__ __forceinline__ bool meta_alive(const uint4& m) { return (m.x | m.y | (m.z & 0x3fffu)) != 0u; }

// One block: 32 frames x 8 cells.  The 624 mask bytes and 448 action bytes a frame contributes are contiguous and
// 16-B aligned when cells % 8 == 0 (78*8 = 16*39): they are staged with coalesced 16-B loads into padded rows,
// then thread (frame fi, cell ci) assembles its record and lanes write 32 consecutive frames of one cell.
constexpr int kPackFrames = 32, kPackCells = 8, kPackThreads = 256;
constexpr int kPackMaskPitch = 628, kPackActPitch = 456;

__global__ void __launch_bounds__(kPackThreads) pack_cells_kernel(const uint8_t* __restrict__ mask,
                                                                  const int64_t* __restrict__ action,
                                                                  uint4* __restrict__ meta, uint32_t* __restrict__ alive_bits,
                                                                  long long n_frames, int cells, int* status) {
    __shared__ __align__(16) uint8_t s_mask[kPackFrames * kPackMaskPitch];
    __shared__ __align__(16) uint8_t s_act[kPackFrames * kPackActPitch];
    const int cell_tiles = (cells + kPackCells - 1) / kPackCells;
    const long long f0 = (long long)(blockIdx.x / cell_tiles) * kPackFrames;
    const int c0 = int(blockIdx.x % cell_tiles) * kPackCells;
    const int nc = min(kPackCells, cells - c0);
    const int nf = int(min((long long)kPackFrames, n_frames - f0));
    const int mbytes = nc * kLogitsPerCell, abytes = nc * kCompsPerCell * 8;
    const bool vec = (cells % 8 == 0) && nc == kPackCells &&
                     ((reinterpret_cast<uintptr_t>(mask) | reinterpret_cast<uintptr_t>(action)) & 15) == 0;
    if (vec) {
        // 39 + 28 sixteen-byte pieces per frame; rows are padded by 4 / 8 bytes, so the pieces are stored as words.
        // Warp w stages frames 4w .. 4w+3, its lanes the pieces lane, lane+32, lane+64 of each (no integer division);
        // all of a thread's loads are issued before its first store (12 x 16 B in flight per thread).
        const int wq = threadIdx.x >> 5, ln = threadIdx.x & 31;
        uint4 v[12];
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int q = 0; q < 3; ++q) {
                const int fi = 4 * wq + j, r = ln + 32 * q;
                if (fi < nf && r < 67) {
                    const bool is_m = r < 39;
                    const uint8_t* g = is_m ? mask + ((f0 + fi) * cells + c0) * kLogitsPerCell
                                            : reinterpret_cast<const uint8_t*>(action + ((f0 + fi) * cells + c0) * kCompsPerCell);
                    v[3 * j + q] = __ldg(reinterpret_cast<const uint4*>(g + 16 * (is_m ? r : r - 39)));
                }
            }
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int q = 0; q < 3; ++q) {
                const int fi = 4 * wq + j, r = ln + 32 * q;
                if (fi < nf && r < 67) {
                    const bool is_m = r < 39;
                    uint32_t* d = reinterpret_cast<uint32_t*>((is_m ? s_mask + fi * kPackMaskPitch : s_act + fi * kPackActPitch) +
                                                              16 * (is_m ? r : r - 39));
                    d[0] = v[3 * j + q].x; d[1] = v[3 * j + q].y; d[2] = v[3 * j + q].z; d[3] = v[3 * j + q].w;
                }
            }
    } else {
        for (int fi = 0; fi < nf; ++fi) {
            const uint8_t* gm = mask + ((f0 + fi) * cells + c0) * kLogitsPerCell;
            const uint8_t* ga = reinterpret_cast<const uint8_t*>(action + ((f0 + fi) * cells + c0) * kCompsPerCell);
            for (int i = threadIdx.x; i < mbytes; i += kPackThreads) s_mask[fi * kPackMaskPitch + i] = gm[i];
            for (int i = threadIdx.x; i < abytes; i += kPackThreads) s_act[fi * kPackActPitch + i] = ga[i];
        }
    }
    __syncthreads();
    const int fi = threadIdx.x & 31, ci = threadIdx.x >> 5;
    if (ci >= nc) return;                               // whole warps leave together (ci is the warp index)
    const bool live = fi < nf;                          // lanes past the last frame only take part in the ballot
    const int fr = live ? fi : 0;
    // 78 mask bytes -> 78 bits, four bytes at a time: the cell starts on an even byte of a 4-byte aligned row, so its
    // bytes are read as 21 aligned words shifted by 0 or 16 bits; per word, vcmpne4 gives 0xff per non-zero byte and
    // a multiply gathers one bit of each byte into a nibble
    const uint32_t* mw = reinterpret_cast<const uint32_t*>(s_mask + fr * kPackMaskPitch + ((ci * kLogitsPerCell) & ~3));
    const int sh = ((ci * kLogitsPerCell) & 2) * 8;
    uint32_t w[3] = {0u, 0u, 0u};
    uint32_t prev = mw[0];
#pragma unroll
    for (int i = 0; i < 20; ++i) {                     // 20 words = 80 bytes; bytes 78, 79 are masked off below
        const uint32_t next = mw[i + 1];
        const uint32_t word = __funnelshift_r(prev, next, sh);
        prev = next;
        const uint32_t nz = __vcmpne4(word, 0u) & 0x08040201u;
        const uint32_t nib = (nz * 0x01010101u) >> 24;              // bit j = byte j of the word is non-zero
        w[(4 * i) >> 5] |= nib << ((4 * i) & 31);
    }
    w[2] &= 0x3fffu;
    const long long* a = reinterpret_cast<const long long*>(s_act + fr * kPackActPitch + ci * kCompsPerCell * 8);
    constexpr int LEN[7] = {6, 4, 4, 4, 4, 7, 49};
    uint32_t act[7];
    bool bad = false;
#pragma unroll
    for (int k = 0; k < 7; ++k) {
        const long long v = a[k];
        bad = bad || v < 0 || v >= LEN[k];
        act[k] = uint32_t(v < 0 ? 0 : (v >= LEN[k] ? LEN[k] - 1 : v));       // clamped like csrc/head.cu; flagged below
    }
    if (bad && live && status) atomicOr(status, kStatusBadAction);
    uint4 r;
    r.x = w[0];
    r.y = w[1];
    r.z = w[2] | (act[0] << 14) | (act[1] << 20) | (act[2] << 26);
    r.w = act[3] | (act[4] << 6) | (act[5] << 12) | (act[6] << 18);
    if (live) meta[(long long)(c0 + ci) * n_frames + f0 + fi] = r;
    if (alive_bits) {
        // one word per (cell, 32 frames): bit i = frame f0 + i has a valid action in this cell (the lists of mb_cells_bwd)
        const unsigned bal = __ballot_sync(0xffffffffu, live && (w[0] | w[1] | w[2]) != 0u);
        if (fi == 0) alive_bits[(long long)(c0 + ci) * ((n_frames + 31) / 32) + f0 / 32] = bal;
    }
}

// W [78C, K] -> hi / lo [80C, K] (TF32 split as in gemm.cu: x = hi + lo up to 2^-22 |x|), bias -> [80C]
__global__ void __launch_bounds__(256) pack_actor_weight_kernel(const float* __restrict__ w, const float* __restrict__ b,
                                                                float* __restrict__ hi, float* __restrict__ lo,
                                                                float* __restrict__ bp, int cells, int K) {
    const long long total = (long long)cells * kCellPad * K;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long row = i / K;
        const int k = int(i - row * K);
        const int cell = int(row / kCellPad), r = int(row % kCellPad);
        float v = 0.0f;
        if (r < kLogitsPerCell) v = w[((long long)cell * kLogitsPerCell + r) * K + k];
        uint32_t h, l;
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(v));
        const float rem = v - __uint_as_float(h);
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(l) : "f"(rem));
        hi[i] = __uint_as_float(h);
        lo[i] = __uint_as_float(l);
        if (k == 0) bp[row] = (r < kLogitsPerCell && b) ? b[cell * kLogitsPerCell + r] : 0.0f;
    }
}

// ---- head math on one cell held in registers ---------------------------------------------------
// Same formulas as csrc/head.cu (z' = mask ? z : -1e8; m = max z'; e = 2^((z'-m) log2 e); S = sum e; W = sum e (z'-m);
// lse = m + ln S; logp = z'_a - lse; H = ln S - W/S; a fully masked component gives exactly 0 / 0, quirk Q1).
__device__ __forceinline__ float fused_log(float x) { return lg2_approx(x) * kLn2; }

template <int OFF, int LEN>
__device__ __forceinline__ void seg_forward(const float (&z)[kCellPad], const uint4& meta, int a, float& lp, float& en) {
    float zz[LEN];
#pragma unroll
    for (int i = 0; i < LEN; ++i) zz[i] = meta_bit(meta, OFF + i) ? z[OFF + i] : kMaskFill;       // model.py:43
    float m4[4] = {zz[0], zz[0], zz[0], zz[0]};
#pragma unroll
    for (int i = 1; i < LEN; ++i) m4[i & 3] = fmaxf(m4[i & 3], zz[i]);
    const float m = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
    float S4[4] = {0.f, 0.f, 0.f, 0.f}, W4[4] = {0.f, 0.f, 0.f, 0.f};
    float za = kMaskFill;
#pragma unroll
    for (int i = 0; i < LEN; ++i) {
        const float d = zz[i] - m;
        const float e = ex2_approx(d * kLog2e);
        S4[i & 3] += e;
        W4[i & 3] = fmaf(e, d, W4[i & 3]);
        za = (i == a) ? zz[i] : za;                       // sentinel when the taken entry is masked (Q2)
    }
    const float S = (S4[0] + S4[1]) + (S4[2] + S4[3]), W = (W4[0] + W4[1]) + (W4[2] + W4[3]);
    const float logS = fused_log(S);
    const float lse = m + logS;
    const bool dead = m < -5e7f;
    const float h0 = dead ? (lse - m) : logS;             // fully masked: fl(-1e8 + ln n) = -1e8, so 0 (Q1)
    lp += za - lse;
    en += h0 - __fdividef(W, S);
}

__device__ __forceinline__ void cell_forward(const float (&z)[kCellPad], const uint4& meta, float& lp, float& en) {
    lp = 0.f;
    en = 0.f;
    seg_forward<0, 6>(z, meta, meta_action(meta, 0), lp, en);
    seg_forward<6, 4>(z, meta, meta_action(meta, 1), lp, en);
    seg_forward<10, 4>(z, meta, meta_action(meta, 2), lp, en);
    seg_forward<14, 4>(z, meta, meta_action(meta, 3), lp, en);
    seg_forward<18, 4>(z, meta, meta_action(meta, 4), lp, en);
    seg_forward<22, 7>(z, meta, meta_action(meta, 5), lp, en);
    seg_forward<29, 49>(z, meta, meta_action(meta, 6), lp, en);
}

// ---- the fused GEMM --------------------------------------------------------------------------
constexpr int kFuBM = 128, kFuBN = 3 * kCellPad, kFuBK = 32, kFuStages = 2;
constexpr int kFuEpiWarps = 12;                      // 4 TMEM lane quadrants x 3 cells
constexpr int kFuThreads = 32 * (2 + kFuEpiWarps);   // TMA warp, MMA warp, epilogue
constexpr int kFuTileA = kFuBM * kFuBK * 4;          // 16 KB
constexpr int kFuTileB = kFuBN * kFuBK * 4;          // 30 KB
constexpr int kFuStage = 2 * kFuTileA + 2 * kFuTileB;        // 92 KB
constexpr int kFuSmem = kFuStages * kFuStage + 1024;
constexpr uint32_t kFuTmemCols = 512;                // hi*hi sum in columns [0,240), cross terms in [256,496)
constexpr uint32_t kFuIdesc = tf32_idesc(kFuBM, kFuBN);
static_assert(kFuTileA % 1024 == 0 && kFuTileB % 1024 == 0, "swizzle atoms are 1024 B");

struct FuMaps {
    CUtensorMap a_hi, a_lo, b_hi, b_lo;
};

__global__ void __launch_bounds__(kFuThreads, 1) actor_head_fwd_kernel(const __grid_constant__ FuMaps maps,
                                                                       const float* __restrict__ bias_p,
                                                                       const uint4* __restrict__ meta,
                                                                       float2* __restrict__ cell_out,
                                                                       float* __restrict__ logits_out, int alive_only,
                                                                       int M, int cells, int K, int* status) {
    extern __shared__ uint8_t fu_smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[kFuStages];
    __shared__ __align__(8) uint64_t empty_bar[kFuStages];
    __shared__ __align__(8) uint64_t tfull_bar, tempty_bar;
    __shared__ uint32_t tmem_slot;
    uint8_t* smem = fu_smem_raw + ((1024u - (smem_u32(fu_smem_raw) & 1023u)) & 1023u);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kFuStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&tfull_bar, 1);
        mbar_init(&tempty_bar, kFuEpiWarps);
        fence_mbar_init();
        tma_prefetch_desc(&maps.a_hi);
        tma_prefetch_desc(&maps.a_lo);
        tma_prefetch_desc(&maps.b_hi);
        tma_prefetch_desc(&maps.b_lo);
    }
    if (warp == 1) tmem_alloc(&tmem_slot, kFuTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    const int tiles_m = (M + kFuBM - 1) / kFuBM, tiles_n = (cells + 2) / 3;
    const int n_tiles = tiles_m * tiles_n;
    const int kblocks = K / kFuBK;
    // tiles of one frame block are consecutive: the A tile (128 frames) is fetched from L2 by neighbouring CTAs at
    // about the same time, the B tiles (the whole weight, 40 MB with hi + lo at 16x16) stream past every frame block

    if (warp == 0) {
        if (lane == 0) {
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                const int m0 = (tile / tiles_n) * kFuBM, n0 = (tile % tiles_n) * kFuBN;
                for (int kb = 0; kb < kblocks; ++kb, ++it) {
                    const int s = it % kFuStages;
                    mbar_wait(&empty_bar[s], ((it / kFuStages) & 1) ^ 1, status);
                    uint8_t* st = smem + s * kFuStage;
                    mbar_arrive_expect_tx(&full_bar[s], uint32_t(kFuStage));
                    tma_load_2d(st, &maps.a_hi, kb * kFuBK, m0, &full_bar[s]);
                    tma_load_2d(st + kFuTileA, &maps.a_lo, kb * kFuBK, m0, &full_bar[s]);
                    tma_load_2d(st + 2 * kFuTileA, &maps.b_hi, kb * kFuBK, n0, &full_bar[s]);
                    tma_load_2d(st + 2 * kFuTileA + kFuTileB, &maps.b_lo, kb * kFuBK, n0, &full_bar[s]);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            uint32_t it = 0, tcount = 0;
            const uint32_t d_main = tmem_base, d_lo = tmem_base + 256u;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++tcount) {
                mbar_wait(&tempty_bar, (tcount & 1) ^ 1, status);
                tc_fence_after();
                for (int kb = 0; kb < kblocks; ++kb, ++it) {
                    const int s = it % kFuStages;
                    mbar_wait(&full_bar[s], (it / kFuStages) & 1, status);
                    tc_fence_after();
                    const uint32_t st = smem_u32(smem + s * kFuStage);
                    const uint64_t a_hi = make_kmajor_sw128_desc(st), a_lo = make_kmajor_sw128_desc(st + kFuTileA);
                    const uint64_t b_hi = make_kmajor_sw128_desc(st + 2 * kFuTileA);
                    const uint64_t b_lo = make_kmajor_sw128_desc(st + 2 * kFuTileA + kFuTileB);
#pragma unroll
                    for (int k = 0; k < kFuBK / 8; ++k) {
                        const uint64_t adv = uint64_t((k * 8 * 4) >> 4);
                        mma_tf32(d_main, a_hi + adv, b_hi + adv, kFuIdesc, (kb | k) != 0);
                        mma_tf32(d_lo, a_lo + adv, b_hi + adv, kFuIdesc, (kb | k) != 0);
                        mma_tf32(d_lo, a_hi + adv, b_lo + adv, kFuIdesc, 1);
                    }
                    tc_commit(&empty_bar[s]);
                }
                tc_commit(&tfull_bar);
            }
        }
    } else {
        // ===================== epilogue: one thread per (frame, cell) =====================
        const int q = warp & 3;                      // TMEM lane quadrant this warp may read (hardware rule)
        const int j = (warp - 2) >> 2;               // which of the tile's three cells
        uint32_t tcount = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++tcount) {
            const int m0 = (tile / tiles_n) * kFuBM;
            const int cell = (tile % tiles_n) * 3 + j;
            mbar_wait(&tfull_bar, tcount & 1, status);
            tc_fence_after();
            const uint32_t lane_base = uint32_t(q * 32) << 16;
            float z[kCellPad];
#pragma unroll
            for (int c = 0; c < kCellPad / 16; ++c) {
                float v[16], w[16];
                tmem_ld16(tmem_base + lane_base + uint32_t(j * kCellPad + c * 16), v);
                tmem_ld16(tmem_base + lane_base + 256u + uint32_t(j * kCellPad + c * 16), w);
#pragma unroll
                for (int i = 0; i < 16; ++i) z[c * 16 + i] = v[i] + w[i];
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty_bar);       // the MMA warp may start the next tile
            const int row = m0 + q * 32 + lane;
            if (row < M && cell < cells) {
                const float4* bp = reinterpret_cast<const float4*>(bias_p + cell * kCellPad);
#pragma unroll
                for (int i = 0; i < kCellPad / 4; ++i) {
                    const float4 b = __ldg(bp + i);
                    z[4 * i] += b.x; z[4 * i + 1] += b.y; z[4 * i + 2] += b.z; z[4 * i + 3] += b.w;
                }
                const uint4 mt = __ldg(meta + (long long)cell * M + row);
                const bool alive = meta_alive(mt);
                // model.py:208 policy_logits, on request: every cell, or only the cells that have a valid action (all
                // the backward pass needs: the gradient of a cell without one is exactly zero, quirk Q1)
                if (logits_out && (alive || !alive_only)) {
                    float2* o = reinterpret_cast<float2*>(logits_out + ((long long)row * cells + cell) * kLogitsPerCell);
#pragma unroll
                    for (int i = 0; i < kLogitsPerCell / 2; ++i) o[i] = make_float2(z[2 * i], z[2 * i + 1]);
                }
                float lp = 0.f, en = 0.f;
                if (alive) cell_forward(z, mt, lp, en);       // a cell without a valid action adds exactly 0
                cell_out[(long long)cell * M + row] = make_float2(lp, en);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 1) tmem_dealloc(tmem_base, kFuTmemCols);
}

// ---- the same forward over the LIVE (frame, cell) pairs only -------------------------------------------------
// The learner never reads the logits (libs/utils.py:237-329) and a cell without a valid action contributes exactly 0 to
// log-prob and entropy (quirk Q1), so only the pairs with a valid action -- 6 % at microRTS densities -- need their 78
// logits at all.  Tiles are 128 entries of ONE cell out of the cell-ordered lists of csrc/cells.cu (build_cell_lists,
// chunk = 128): M = 128 gathered rows of `hidden`, N = 80 (that cell's padded weight rows), K = 256.
//   warps 8-11  gather: thread = entry; the row's hi / lo halves arrive by 16-byte cp.async straight into the K-major,
//               128-byte-swizzled operand stage (the halves are pre-split in global memory, so no conversion), three
//               k-blocks deep (a fourth stage changes nothing); thread 0 also issues the two TMA boxes of the cell's weight rows;
//   warp 12     one lane issues the MMAs (same three products, same accumulator split and order as the dense kernel:
//               the logits are bit-identical);
//   warps 0-7   two epilogue groups over two TMEM slots: bias, the head of csrc/head.cu on the row's 80 columns,
//               `cell_out` scattered to [cell][frame] (pre-zeroed: dead pairs stay exactly 0), logits of the live
//               cells on request (all the backward pass needs).
constexpr int kLfM = 128, kLfStages = 3, kLfThreads = 32 * 13;
constexpr int kLfTileA = kLfM * kFuBK * 4;           // 16 KB
constexpr int kLfTileB = kCellPad * kFuBK * 4;       // 10 KB = ten swizzle atoms
constexpr int kLfStage = 2 * kLfTileA + 2 * kLfTileB;
constexpr int kLfSmem = kLfStages * kLfStage + 1024;
constexpr uint32_t kLfIdesc = tf32_idesc(kLfM, kCellPad);
static_assert(kLfTileB % 1024 == 0, "swizzle atoms are 1024 B");

struct LfMaps {
    CUtensorMap b_hi, b_lo;
};

__device__ __forceinline__ void lf_cp_async16(void* dst_smem, const void* src, bool valid) {
    const uint32_t bytes = valid ? 16u : 0u;         // 0 source bytes: the destination is zero-filled
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes) : "memory");
}

__global__ void __launch_bounds__(kLfThreads, 1) actor_head_list_fwd_kernel(const __grid_constant__ LfMaps maps,
                                                                            const float* __restrict__ a_hi, const float* __restrict__ a_lo,
                                                                            const float* __restrict__ bias_p,
                                                                            const uint4* __restrict__ meta,
                                                                            const int* __restrict__ entries,
                                                                            const int* __restrict__ chunk_cell,
                                                                            const int* __restrict__ header,
                                                                            float2* __restrict__ cell_out,
                                                                            float* __restrict__ logits_out, int M, int cells, int K,
                                                                            int* status, int debug) {
    extern __shared__ uint8_t lf_smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[kLfStages], empty_bar[kLfStages], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_slot;
    uint8_t* smem = lf_smem_raw + ((1024u - (smem_u32(lf_smem_raw) & 1023u)) & 1023u);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kLfStages; ++s) {
            mbar_init(&full_bar[s], 129);            // 128 gathering threads + the expect_tx arrival of the weight boxes
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&acc_full[s], 1);
            mbar_init(&acc_empty[s], 4);
        }
        fence_mbar_init();
        tma_prefetch_desc(&maps.b_hi);
        tma_prefetch_desc(&maps.b_lo);
    }
    if (warp == 12) tmem_alloc(&tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;
    const int n_tiles = __ldg(header);               // chunks of 128 entries, each of one cell
    const int kblocks = K / kFuBK;

    if (warp >= 8 && warp < 12) {
        // ===================== gather =====================
        // eight lanes fetch the eight 16-byte pieces of one row's 128-byte k-block, so a warp instruction is four
        // whole-line requests instead of 32 half-used sectors (16-byte pieces from 32 different rows left the kernel
        // bound by L2 request rate: 0.34 ms against 0.22 ms); a thread serves rows t/8 + 16 i, i < 8
        const int t = threadIdx.x - 256;
        const int c8 = t & 7, r0 = t >> 3;
        uint32_t issued = 0, arrived = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const int cell = __ldg(chunk_cell + tile);
            const float* rh[8];
            const float* rl[8];
            bool valid[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int f = __ldg(entries + (long long)tile * kLfM + r0 + 16 * i);       // -1: padding of the cell's last chunk
                valid[i] = f >= 0;
                rh[i] = a_hi + (long long)(valid[i] ? f : 0) * K + 4 * c8;
                rl[i] = a_lo + (long long)(valid[i] ? f : 0) * K + 4 * c8;
            }
            for (int kb = 0; kb < kblocks; ++kb) {
                const int s = issued % kLfStages;
                mbar_wait(&empty_bar[s], ((issued / kLfStages) & 1) ^ 1, status);
                uint8_t* st = smem + s * kLfStage;
                if (t == 0) {
                    mbar_arrive_expect_tx(&full_bar[s], uint32_t(2 * kLfTileB));
                    tma_load_2d(st + 2 * kLfTileA, &maps.b_hi, kb * kFuBK, cell * kCellPad, &full_bar[s]);
                    tma_load_2d(st + 2 * kLfTileA + kLfTileB, &maps.b_lo, kb * kFuBK, cell * kCellPad, &full_bar[s]);
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int r = r0 + 16 * i;
                    const uint32_t off = uint32_t(r * 128) + ((uint32_t(c8) ^ uint32_t(r & 7)) << 4);
                    if (debug & 2) continue;                 // timing experiment: no gather
                    lf_cp_async16(st + off, rh[i] + kb * kFuBK, valid[i]);
                    lf_cp_async16(st + kLfTileA + off, rl[i] + kb * kFuBK, valid[i]);
                }
                asm volatile("cp.async.commit_group;" ::: "memory");
                ++issued;
                if (issued >= uint32_t(kLfStages)) {          // all but the last kLfStages - 1 groups have landed
                    asm volatile("cp.async.wait_group %0;" ::"n"(kLfStages - 1) : "memory");
                    fence_proxy_async_smem();
                    while (arrived + (kLfStages - 1) < issued) {
                        mbar_arrive(&full_bar[arrived % kLfStages]);
                        ++arrived;
                    }
                }
            }
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        fence_proxy_async_smem();
        while (arrived < issued) {
            mbar_arrive(&full_bar[arrived % kLfStages]);
            ++arrived;
        }
    } else if (warp == 12) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            uint32_t it = 0, job = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++job) {
                const uint32_t slot = job & 1u;
                mbar_wait(&acc_empty[slot], ((job >> 1) & 1u) ^ 1u, status);
                tc_fence_after();
                const uint32_t d_main = tmem_base + slot * 256u, d_lo = d_main + 128u;
                for (int kb = 0; kb < kblocks; ++kb, ++it) {
                    const int s = it % kLfStages;
                    mbar_wait(&full_bar[s], (it / kLfStages) & 1, status);
                    tc_fence_after();
                    const uint32_t st = smem_u32(smem + s * kLfStage);
                    const uint64_t ah = make_kmajor_sw128_desc(st), al = make_kmajor_sw128_desc(st + kLfTileA);
                    const uint64_t bh = make_kmajor_sw128_desc(st + 2 * kLfTileA);
                    const uint64_t bl = make_kmajor_sw128_desc(st + 2 * kLfTileA + kLfTileB);
#pragma unroll
                    for (int k = 0; k < kFuBK / 8; ++k) {
                        const uint64_t adv = uint64_t((k * 8 * 4) >> 4);
                        if (debug & 4) continue;             // timing experiment: no MMAs
                        mma_tf32(d_main, ah + adv, bh + adv, kLfIdesc, (kb | k) != 0);
                        mma_tf32(d_lo, al + adv, bh + adv, kLfIdesc, (kb | k) != 0);
                        mma_tf32(d_lo, ah + adv, bl + adv, kLfIdesc, 1);
                    }
                    tc_commit(&empty_bar[s]);
                }
                tc_commit(&acc_full[slot]);
            }
        }
    } else {
        // ===================== epilogue: one thread per list entry =====================
        const int group = warp >> 2, q = warp & 3;
        uint32_t job = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++job) {
            if (int(job & 1u) != group) continue;
            const uint32_t slot = job & 1u;
            mbar_wait(&acc_full[slot], (job >> 1) & 1u, status);
            tc_fence_after();
            const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + slot * 256u;
            if (debug & 1) {                                 // timing experiment: no epilogue work
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&acc_empty[slot]);
                continue;
            }
            float z[kCellPad];
#pragma unroll
            for (int c = 0; c < kCellPad / 16; ++c) {
                float v[16], w[16];
                tmem_ld16(taddr + uint32_t(c * 16), v);
                tmem_ld16(taddr + 128u + uint32_t(c * 16), w);
#pragma unroll
                for (int i = 0; i < 16; ++i) z[c * 16 + i] = v[i] + w[i];
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[slot]);
            const int f = __ldg(entries + (long long)tile * kLfM + q * 32 + lane);
            if (f >= 0) {
                const int cell = __ldg(chunk_cell + tile);
                const float4* bp = reinterpret_cast<const float4*>(bias_p + cell * kCellPad);
#pragma unroll
                for (int i = 0; i < kCellPad / 4; ++i) {
                    const float4 b = __ldg(bp + i);
                    z[4 * i] += b.x; z[4 * i + 1] += b.y; z[4 * i + 2] += b.z; z[4 * i + 3] += b.w;
                }
                const uint4 mt = __ldg(meta + (long long)cell * M + f);
                if (logits_out) {
                    // a cell's 78 logits are 312 contiguous bytes starting on an 8-byte boundary: 16-byte stores wherever
                    // the alignment allows (19 + one 8-byte store instead of 39 8-byte stores)
                    float* o = logits_out + ((long long)f * cells + cell) * kLogitsPerCell;
                    if ((reinterpret_cast<uintptr_t>(o) & 15) == 0) {
#pragma unroll
                        for (int i = 0; i < 19; ++i)
                            reinterpret_cast<float4*>(o)[i] = make_float4(z[4 * i], z[4 * i + 1], z[4 * i + 2], z[4 * i + 3]);
                        *reinterpret_cast<float2*>(o + 76) = make_float2(z[76], z[77]);
                    } else {
                        *reinterpret_cast<float2*>(o) = make_float2(z[0], z[1]);
#pragma unroll
                        for (int i = 0; i < 19; ++i)
                            reinterpret_cast<float4*>(o + 2)[i] = make_float4(z[4 * i + 2], z[4 * i + 3], z[4 * i + 4], z[4 * i + 5]);
                    }
                }
                float lp = 0.f, en = 0.f;
                cell_forward(z, mt, lp, en);
                cell_out[(long long)cell * M + f] = make_float2(lp, en);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 12) tmem_dealloc(tmem_base, 512);
}

// per frame: the C cell results in cell order -> logp, ent
__global__ void __launch_bounds__(64) frame_sum_kernel(const float2* __restrict__ cell_out, float* __restrict__ logp,
                                                        float* __restrict__ ent, long long M, int cells) {
    const long long f = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (f >= M) return;
    float lp = 0.f, en = 0.f;
    int c = 0;
    for (; c + 16 <= cells; c += 16) {                 // 16 loads in flight, added in cell order
        float2 v[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = __ldg(cell_out + (long long)(c + i) * M + f);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            lp += v[i].x;
            en += v[i].y;
        }
    }
    for (; c < cells; ++c) {
        const float2 v = __ldg(cell_out + (long long)c * M + f);
        lp += v.x;
        en += v.y;
    }
    logp[f] = lp;
    ent[f] = en;
}

// ---- host ------------------------------------------------------------------------------------
typedef CUresult (*FuEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                               const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                               CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

FuEncodeFn fu_encode_fn() {
    static FuEncodeFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<FuEncodeFn>(ptr);
    });
    return fn;
}

// Tensor maps are re-used across calls with the same operands (the learner calls with the same buffers every
// step): a small direct-mapped cache keyed by (pointer, rows, K, box rows).
struct FuMapKey {
    const void* base;
    long long rows;
    int K, box_rows;
    bool operator==(const FuMapKey& o) const { return base == o.base && rows == o.rows && K == o.K && box_rows == o.box_rows; }
};
struct FuMapSlot {
    FuMapKey key;
    CUtensorMap map;
    bool valid;
};
FuMapSlot g_fu_maps[64];
std::mutex g_fu_mutex;

bool fu_map(FuEncodeFn enc, CUtensorMap* m, const float* base, long long rows, int K, int box_rows) {
    const FuMapKey key{base, rows, K, box_rows};
    size_t h = (reinterpret_cast<uintptr_t>(base) >> 8) ^ (size_t(rows) * 0x9E3779B97F4A7C15ull) ^ (size_t(box_rows) << 5);
    FuMapSlot& slot = g_fu_maps[(h ^ (h >> 23)) & 63];
    std::lock_guard<std::mutex> lock(g_fu_mutex);
    if (slot.valid && slot.key == key) {
        *m = slot.map;
        return true;
    }
    const cuuint64_t gdim[2] = {cuuint64_t(K), cuuint64_t(rows)};
    const cuuint64_t gstride[1] = {cuuint64_t(K) * 4u};
    const cuuint32_t box[2] = {cuuint32_t(kFuBK), cuuint32_t(box_rows)};
    const cuuint32_t estr[2] = {1, 1};
    if (enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstride, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return false;
    slot.key = key;
    slot.map = *m;
    slot.valid = true;
    return true;
}

}  // namespace

int pack_cells(const uint8_t* mask, const int64_t* action, void* meta, void* alive_bits, long long n_frames, int cells,
               int* status, cudaStream_t st) {
    const long long blocks = ((n_frames + kPackFrames - 1) / kPackFrames) * ((cells + kPackCells - 1) / kPackCells);
    if (blocks > 0x7fffffffLL) {
        set_error("mb_pack_cells: problem too large for one launch");
        return MB_ERR_INVALID;
    }
    pack_cells_kernel<<<unsigned(blocks), kPackThreads, 0, st>>>(mask, action, reinterpret_cast<uint4*>(meta),
                                                                 reinterpret_cast<uint32_t*>(alive_bits), n_frames, cells, status);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}

int pack_actor_weight(const float* w, const float* b, float* hi, float* lo, float* bp, int cells, int K, cudaStream_t st) {
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    const long long total = (long long)cells * kCellPad * K;
    const long long want = (total + 255) / 256, cap = (long long)sm_count(dev) * 8;
    pack_actor_weight_kernel<<<int(want < cap ? want : cap), 256, 0, st>>>(w, b, hi, lo, bp, cells, K);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}

int actor_head_fwd(const float* a_hi, const float* a_lo, const float* wp_hi, const float* wp_lo, const float* bias_p,
                   const void* meta, float* cell_out, float* logits_out, int alive_only, float* logp, float* ent,
                   long long M, int cells, int K, int* status, cudaStream_t st) {
    FuEncodeFn enc = fu_encode_fn();
    if (!enc) {
        set_error("mb_actor_head_fwd: cuTensorMapEncodeTiled is not available from this driver");
        return MB_ERR_UNSUPPORTED;
    }
    FuMaps maps;
    const long long wrows = (long long)cells * kCellPad;
    if (!fu_map(enc, &maps.a_hi, a_hi, M, K, kFuBM) || !fu_map(enc, &maps.a_lo, a_lo, M, K, kFuBM) ||
        !fu_map(enc, &maps.b_hi, wp_hi, wrows, K, kFuBN) || !fu_map(enc, &maps.b_lo, wp_lo, wrows, K, kFuBN)) {
        set_error("mb_actor_head_fwd: cuTensorMapEncodeTiled rejected M=%lld cells=%d K=%d", M, cells, K);
        return MB_ERR_INVALID;
    }
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    static bool attr_set[64] = {};
    bool& done = attr_set[(dev >= 0 && dev < 64) ? dev : 0];
    if (!done) {
        MB_CUDA(cudaFuncSetAttribute(actor_head_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFuSmem));
        done = true;
    }
    const long long tiles = ((M + kFuBM - 1) / kFuBM) * ((cells + 2) / 3);
    const int sms = sm_count(dev);
    const int grid = int(tiles < sms ? tiles : sms);
    actor_head_fwd_kernel<<<grid, kFuThreads, kFuSmem, st>>>(maps, bias_p, reinterpret_cast<const uint4*>(meta),
                                                            reinterpret_cast<float2*>(cell_out), logits_out, alive_only,
                                                            int(M), cells, K, status);
    MB_CUDA(cudaGetLastError());
    if (logp && ent) {
        frame_sum_kernel<<<unsigned((M + 63) / 64), 64, 0, st>>>(reinterpret_cast<const float2*>(cell_out), logp, ent, M, cells);
        MB_CUDA(cudaGetLastError());
    }
    return MB_OK;
}

// csrc/cells.cu
size_t cell_lists_ws_ints(long long n_frames, int cells, int chunk);
int build_cell_lists(const uint32_t* alive_bits, int* ws, long long n_frames, int cells, int chunk, cudaStream_t st,
                     int** chunk_cell_out, int** header_out, int** entries_out);

std::atomic<int> g_lf_debug{0};      // mb_set_tuning("head_list_debug", bits): skip a role's work (timing experiments, wrong results)
void head_list_set_debug(int bits) { g_lf_debug = bits; }

int actor_head_list_fwd(const float* a_hi, const float* a_lo, const float* wp_hi, const float* wp_lo, const float* bias_p,
                        const void* meta, const void* alive_bits, float* cell_out, float* logits_out, float* logp, float* ent,
                        void* workspace, size_t workspace_bytes, long long M, int cells, int K, int* status, cudaStream_t st) {
    FuEncodeFn enc = fu_encode_fn();
    if (!enc) {
        set_error("mb_actor_head_list_fwd: cuTensorMapEncodeTiled is not available from this driver");
        return MB_ERR_UNSUPPORTED;
    }
    if (workspace_bytes < cell_lists_ws_ints(M, cells, kLfM) * sizeof(int)) {
        set_error("mb_actor_head_list_fwd: workspace smaller than mb_actor_head_list_workspace_bytes(n_frames, cells)");
        return MB_ERR_INVALID;
    }
    LfMaps maps;
    const long long wrows = (long long)cells * kCellPad;
    if (!fu_map(enc, &maps.b_hi, wp_hi, wrows, K, kCellPad) || !fu_map(enc, &maps.b_lo, wp_lo, wrows, K, kCellPad)) {
        set_error("mb_actor_head_list_fwd: cuTensorMapEncodeTiled rejected cells=%d K=%d", cells, K);
        return MB_ERR_INVALID;
    }
    int *chunk_cell = nullptr, *header = nullptr, *entries = nullptr;
    const int rc = build_cell_lists(reinterpret_cast<const uint32_t*>(alive_bits), reinterpret_cast<int*>(workspace), M, cells, kLfM,
                                    st, &chunk_cell, &header, &entries);
    if (rc != MB_OK) return rc;
    MB_CUDA(cudaMemsetAsync(cell_out, 0, size_t(cells) * size_t(M) * sizeof(float2), st));      // dead pairs: exactly 0 (Q1)
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    static bool attr_set[64] = {};
    bool& done = attr_set[(dev >= 0 && dev < 64) ? dev : 0];
    if (!done) {
        MB_CUDA(cudaFuncSetAttribute(actor_head_list_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kLfSmem));
        done = true;
    }
    actor_head_list_fwd_kernel<<<sm_count(dev), kLfThreads, kLfSmem, st>>>(maps, a_hi, a_lo, bias_p, reinterpret_cast<const uint4*>(meta),
                                                                          entries, chunk_cell, header,
                                                                          reinterpret_cast<float2*>(cell_out), logits_out, int(M), cells,
                                                                          K, status, g_lf_debug.load());
    MB_CUDA(cudaGetLastError());
    if (logp && ent) {
        frame_sum_kernel<<<unsigned((M + 63) / 64), 64, 0, st>>>(reinterpret_cast<const float2*>(cell_out), logp, ent, M, cells);
        MB_CUDA(cudaGetLastError());
    }
    return MB_OK;
}

}  // namespace mb

extern "C" size_t mb_actor_head_list_workspace_bytes(int64_t n_frames, int32_t cells) {
    if (n_frames <= 0 || cells <= 0) return 0;
    return (mb::cell_lists_ws_ints(n_frames, cells, mb::kLfM) * sizeof(int) + 255) & ~size_t(255);
}

extern "C" int mb_actor_head_list_fwd(const float* a_hi, const float* a_lo, const float* w_hi, const float* w_lo,
                                      const float* bias_p, const void* cell_meta, const void* alive_bits, float* cell_out,
                                      float* logits_out, float* logp_out, float* ent_out, void* workspace, size_t workspace_bytes,
                                      int64_t n_frames, int32_t cells, int32_t k, int32_t* status, void* stream) {
    if (n_frames < 0 || cells <= 0 || cells > 1024 || k <= 0 || k % 32 != 0 || n_frames > 0x7fffffff) {
        mb::set_error("mb_actor_head_list_fwd: n_frames=%lld cells=%d k=%d: need 0 < cells <= 1024 and k %% 32 == 0", (long long)n_frames,
                      cells, k);
        return MB_ERR_INVALID;
    }
    if (n_frames == 0) return MB_OK;
    if (!a_hi || !a_lo || !w_hi || !w_lo || !bias_p || !cell_meta || !alive_bits || !cell_out || !workspace ||
        ((logp_out == nullptr) != (ent_out == nullptr))) {
        mb::set_error("mb_actor_head_list_fwd: null pointer (logp_out / ent_out: both or neither)");
        return MB_ERR_INVALID;
    }
    const uintptr_t all = reinterpret_cast<uintptr_t>(a_hi) | reinterpret_cast<uintptr_t>(a_lo) | reinterpret_cast<uintptr_t>(w_hi) |
                          reinterpret_cast<uintptr_t>(w_lo) | reinterpret_cast<uintptr_t>(bias_p) |
                          reinterpret_cast<uintptr_t>(cell_meta) | reinterpret_cast<uintptr_t>(cell_out) |
                          reinterpret_cast<uintptr_t>(logits_out) | reinterpret_cast<uintptr_t>(workspace);
    if (all & 15) {
        mb::set_error("mb_actor_head_list_fwd: operands must be 16-byte aligned");
        return MB_ERR_INVALID;
    }
    return mb::actor_head_list_fwd(a_hi, a_lo, w_hi, w_lo, bias_p, cell_meta, alive_bits, cell_out, logits_out, logp_out, ent_out,
                                   workspace, workspace_bytes, n_frames, cells, k, status, static_cast<cudaStream_t>(stream));
}

extern "C" int mb_pack_cells(const uint8_t* mask, const int64_t* action, void* cell_meta, void* alive_bits, int64_t n_frames,
                             int32_t cells, int32_t* status, void* stream) {
    if (n_frames < 0 || cells <= 0) {
        mb::set_error("mb_pack_cells: n_frames=%lld cells=%d out of range", (long long)n_frames, cells);
        return MB_ERR_INVALID;
    }
    if (n_frames == 0) return MB_OK;
    if (!mask || !action || !cell_meta || (reinterpret_cast<uintptr_t>(cell_meta) & 15) ||
        (reinterpret_cast<uintptr_t>(action) & 7)) {
        mb::set_error("mb_pack_cells: null or misaligned pointer (cell_meta 16 B, action 8 B)");
        return MB_ERR_INVALID;
    }
    return mb::pack_cells(mask, action, cell_meta, alive_bits, n_frames, cells, status, static_cast<cudaStream_t>(stream));
}

extern "C" int mb_pack_actor_weight(const float* weight, const float* bias, float* w_hi, float* w_lo, float* bias_p,
                                    int32_t cells, int32_t k, void* stream) {
    if (cells <= 0 || k <= 0 || !weight || !w_hi || !w_lo || !bias_p) {
        mb::set_error("mb_pack_actor_weight: bad arguments (cells=%d k=%d)", cells, k);
        return MB_ERR_INVALID;
    }
    return mb::pack_actor_weight(weight, bias, w_hi, w_lo, bias_p, cells, k, static_cast<cudaStream_t>(stream));
}

extern "C" int mb_actor_head_fwd(const float* a_hi, const float* a_lo, const float* w_hi, const float* w_lo,
                                 const float* bias_p, const void* cell_meta, float* cell_out, float* logits_out,
                                 int32_t logits_alive_only, float* logp_out, float* ent_out, int64_t n_frames,
                                 int32_t cells, int32_t k, int32_t* status, void* stream) {
    if (n_frames < 0 || cells <= 0 || k <= 0 || k % 32 != 0 || n_frames > 0x7fffffff) {
        mb::set_error("mb_actor_head_fwd: n_frames=%lld cells=%d k=%d: need k %% 32 == 0", (long long)n_frames, cells, k);
        return MB_ERR_INVALID;
    }
    if (n_frames == 0) return MB_OK;
    if (!a_hi || !a_lo || !w_hi || !w_lo || !bias_p || !cell_meta || !cell_out || ((logp_out == nullptr) != (ent_out == nullptr))) {
        mb::set_error("mb_actor_head_fwd: null pointer (logp_out / ent_out: both or neither)");
        return MB_ERR_INVALID;
    }
    const uintptr_t all = reinterpret_cast<uintptr_t>(a_hi) | reinterpret_cast<uintptr_t>(a_lo) | reinterpret_cast<uintptr_t>(w_hi) |
                          reinterpret_cast<uintptr_t>(w_lo) | reinterpret_cast<uintptr_t>(bias_p) |
                          reinterpret_cast<uintptr_t>(cell_meta) | reinterpret_cast<uintptr_t>(cell_out) |
                          reinterpret_cast<uintptr_t>(logits_out);
    if (all & 15) {
        mb::set_error("mb_actor_head_fwd: operands must be 16-byte aligned");
        return MB_ERR_INVALID;
    }
    return mb::actor_head_fwd(a_hi, a_lo, w_hi, w_lo, bias_p, cell_meta, cell_out, logits_out, logits_alive_only, logp_out,
                              ent_out, n_frames, cells, k, status, static_cast<cudaStream_t>(stream));
}
