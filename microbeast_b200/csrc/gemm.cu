// Actor Linear forward on the 5th-generation tensor cores (tcgen05, TMEM accumulators, TMA operands)
// at fp32 accuracy, for sm_100a.
//
// Reference: model.py:136,167 -- logits = actor(hidden), a Linear 256 -> 78*C evaluated for every
// frame of the learner batch.  In fp32 on the CUDA cores (cuBLAS SIMT SGEMM) this product was 3.7 ms of a
// 17 ms learner step at 16x16, T=80, 256 rows.  A single TF32 tensor-core product would be 8x faster but
// carries 10 mantissa bits, far outside the 1e-5 parity budget.  This kernel keeps fp32 accuracy by
// splitting both operands into a TF32 "hi" part and a TF32 "lo" remainder (x = hi + lo exactly, up to
// 2^-22) and accumulating three tensor-core products in the fp32 TMEM accumulator:
//     A*B ~= A_hi*B_hi + A_lo*B_hi + A_hi*B_lo          (the dropped A_lo*B_lo term is <= 2^-22 |A||B|)
//
// C[M,N] = A[M,K] * B[N,K]^T + bias[N];  A = hidden (K-major), B = the Linear's weight [N,K] (K-major).
//
// Layout of one CTA (320 threads, one per SM, persistent over 128 x 256 output tiles):
//   warp 0, one lane: TMA producer -- per 32-wide K block four boxes (A_hi, A_lo 128x32; B_hi, B_lo
//                     256x32 fp32, 128-byte swizzle) into a 2-stage shared-memory ring (96 KB a stage);
//   warp 1, one lane: issues tcgen05.mma.kind::tf32 (M=128, N=256, K=8): 4 K-steps x 3 products per
//                     K block; the hi*hi products accumulate in TMEM columns [0,256), the two small cross
//                     terms in columns [256,512) (see the comment at the issue loop: it is what brings the
//                     error down to cuBLAS fp32 level); tcgen05.commit releases the stage to the producer
//                     and, after the last K block, hands both accumulators over;
//   warps 2-9:        epilogue -- tcgen05.ld both accumulators (a warp reads its own 32-lane quadrant, a
//                     thread 128 columns of one output row), sum them in registers, release the
//                     accumulators, then add the bias and store while warp 1 already runs the next tile.
#include <atomic>

#include "common.cuh"
#include "tc.cuh"

#include <cuda.h>

#include <mutex>

#include "../../include/mbeast.h"

namespace mb {

int sm_count(int device);

namespace {

constexpr int kGemmBM = 128, kGemmBN = 256, kGemmBK = 32, kGemmStages = 2;
constexpr int kGemmThreads = 320;                   // TMA warp, MMA warp, 8 epilogue warps
constexpr int kTileA = kGemmBM * kGemmBK * 4;        // 16 KB
constexpr int kTileB = kGemmBN * kGemmBK * 4;        // 32 KB
constexpr int kStageBytes = 2 * kTileA + 2 * kTileB; // 96 KB
constexpr int kGemmSmem = kGemmStages * kStageBytes + 1024;
constexpr uint32_t kTmemCols = 512;                  // 256 columns for the hi*hi sum + 256 for the lo terms

struct GemmMaps {
    CUtensorMap a_hi, a_lo, b_hi, b_lo;
};

// Instruction descriptor of tcgen05.mma.kind::tf32: [4,6) D format 1 = F32; [7,10) A format 2 = TF32;
// [10,13) B format 2 = TF32; bits 15 / 16 = 0: A and B are K-major; [17,23) N >> 3; [24,29) M >> 4.
constexpr uint32_t kIdescTf32 = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(kGemmBN >> 3) << 17) |
                                (uint32_t(kGemmBM >> 4) << 24);

__global__ void __launch_bounds__(256) split_tf32_kernel(const float* __restrict__ x, float* __restrict__ hi,
                                                         float* __restrict__ lo, long long n) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float v = x[i];
        uint32_t h, l;
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(v));                    // nearest TF32: low 13 mantissa bits zero
        const float r = v - __uint_as_float(h);                                // exact
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(l) : "f"(r));
        hi[i] = __uint_as_float(h);
        lo[i] = __uint_as_float(l);
    }
}

__global__ void __launch_bounds__(kGemmThreads, 1) linear_tf32x3_kernel(const __grid_constant__ GemmMaps maps,
                                                                        const float* __restrict__ bias,
                                                                        float* __restrict__ C, int M, int N, int K,
                                                                        int* status, int group_m, int stream_store) {
    extern __shared__ uint8_t gemm_smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[kGemmStages];      // TMA boxes of a stage have landed
    __shared__ __align__(8) uint64_t empty_bar[kGemmStages];     // the MMAs reading a stage have completed
    __shared__ __align__(8) uint64_t tfull_bar[2];               // an accumulator is complete
    __shared__ __align__(8) uint64_t tempty_bar[2];              // the epilogue has drained an accumulator
    __shared__ uint32_t tmem_slot;
    uint8_t* smem = gemm_smem_raw + ((1024u - (smem_u32(gemm_smem_raw) & 1023u)) & 1023u);     // swizzle atoms: 1024 B

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kGemmStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&tfull_bar[b], 1);
            mbar_init(&tempty_bar[b], 8);                        // one arrival per epilogue warp
        }
        fence_mbar_init();
        tma_prefetch_desc(&maps.a_hi);
        tma_prefetch_desc(&maps.a_lo);
        tma_prefetch_desc(&maps.b_hi);
        tma_prefetch_desc(&maps.b_lo);
    }
    if (warp == 1) tmem_alloc(&tmem_slot, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    const int tiles_m = (M + kGemmBM - 1) / kGemmBM, tiles_n = (N + kGemmBN - 1) / kGemmBN;
    const int n_tiles = tiles_m * tiles_n;
    // tile order: row-major (group_m = 0), or bands of group_m tile rows walked column by column so that the
    // CTAs running at the same time share a few A tiles and a few B tiles
    auto tile_mn = [&](int tile, int& tm, int& tn) {
        if (group_m <= 0) {
            tm = tile / tiles_n;
            tn = tile % tiles_n;
        } else {
            const int band = tile / (group_m * tiles_n), first = band * group_m;
            const int rows = min(group_m, tiles_m - first), in_band = tile - band * group_m * tiles_n;
            tm = first + in_band % rows;
            tn = in_band / rows;
        }
    };
    const int kblocks = K / kGemmBK;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                int tm, tn;
                tile_mn(tile, tm, tn);
                const int m0 = tm * kGemmBM, n0 = tn * kGemmBN;
                for (int kb = 0; kb < kblocks; ++kb, ++it) {
                    const int s = it % kGemmStages;
                    const uint32_t ph = (it / kGemmStages) & 1;
                    mbar_wait(&empty_bar[s], ph ^ 1, status);     // passes at once the first time round
                    uint8_t* st = smem + s * kStageBytes;
                    mbar_arrive_expect_tx(&full_bar[s], uint32_t(kStageBytes));
                    tma_load_2d(st, &maps.a_hi, kb * kGemmBK, m0, &full_bar[s]);
                    tma_load_2d(st + kTileA, &maps.a_lo, kb * kGemmBK, m0, &full_bar[s]);
                    tma_load_2d(st + 2 * kTileA, &maps.b_hi, kb * kGemmBK, n0, &full_bar[s]);
                    tma_load_2d(st + 2 * kTileA + kTileB, &maps.b_lo, kb * kGemmBK, n0, &full_bar[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        // The tensor core truncates when it adds a product into the fp32 accumulator, so every MMA into a LARGE
        // accumulator costs a fraction of an ulp of the result.  The hi*hi products (the bulk of the value) go to
        // columns [0,256), the two small cross terms to columns [256,512): their accumulator is 2^-11 the size, so
        // are its truncation errors, and the big one sees 32 additions instead of 96 at K = 256.
        if (lane == 0) {
            uint32_t it = 0, tcount = 0;
            const uint32_t d_main = tmem_base, d_lo = tmem_base + uint32_t(kGemmBN);
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++tcount) {
                mbar_wait(&tempty_bar[0], (tcount & 1) ^ 1, status);      // the epilogue has read the last tile out
                tc_fence_after();
                for (int kb = 0; kb < kblocks; ++kb, ++it) {
                    const int s = it % kGemmStages;
                    const uint32_t ph = (it / kGemmStages) & 1;
                    mbar_wait(&full_bar[s], ph, status);
                    tc_fence_after();
                    const uint32_t st = smem_u32(smem + s * kStageBytes);
                    const uint64_t a_hi = make_kmajor_sw128_desc(st), a_lo = make_kmajor_sw128_desc(st + kTileA);
                    const uint64_t b_hi = make_kmajor_sw128_desc(st + 2 * kTileA);
                    const uint64_t b_lo = make_kmajor_sw128_desc(st + 2 * kTileA + kTileB);
#pragma unroll
                    for (int k = 0; k < kGemmBK / 8; ++k) {
                        const uint64_t adv = uint64_t((k * 8 * 4) >> 4);      // 8 TF32 = 32 B along K inside the swizzle atom
                        mma_tf32(d_main, a_hi + adv, b_hi + adv, kIdescTf32, (kb | k) != 0);
                        mma_tf32(d_lo, a_lo + adv, b_hi + adv, kIdescTf32, (kb | k) != 0);
                        mma_tf32(d_lo, a_hi + adv, b_lo + adv, kIdescTf32, 1);
                    }
                    tc_commit(&empty_bar[s]);                     // the stage is free once these MMAs have read it
                }
                tc_commit(&tfull_bar[0]);                         // both accumulators are complete
            }
        }
    } else {
        // ===================== epilogue (warps 2..9) =====================
        // warp w reads TMEM lane quadrant w % 4 (a hardware rule); warps 2-5 take columns [0,128), warps 6-9
        // [128,256).  A thread pulls its 128 sums (main + lo) into registers, releases the accumulators to the MMA
        // warp at once, and only then does the slow part: bias and the stores.
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;
        uint32_t tcount = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++tcount) {
            int tm, tn;
            tile_mn(tile, tm, tn);
            const int m0 = tm * kGemmBM, n0 = tn * kGemmBN + half * 128;
            mbar_wait(&tfull_bar[0], tcount & 1, status);
            tc_fence_after();
            const uint32_t lane_base = uint32_t(q * 32) << 16;
            float acc[128];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                float v[32], w[32];
                tmem_ld32(tmem_base + lane_base + uint32_t(half * 128 + c * 32), v);
                tmem_ld32(tmem_base + lane_base + uint32_t(kGemmBN + half * 128 + c * 32), w);
#pragma unroll
                for (int j = 0; j < 32; ++j) acc[c * 32 + j] = v[j] + w[j];
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty_bar[0]);
            const int row = m0 + q * 32 + lane;
            if (row < M) {
                float* crow = C + (long long)row * N;
#pragma unroll
                for (int j = 0; j < 128; j += 4) {
                    const int col = n0 + j;
                    if (col < N) {                                // N % 4 == 0 (checked by the host)
                        const float4 bb = bias ? __ldg(reinterpret_cast<const float4*>(bias + col))
                                               : make_float4(0.f, 0.f, 0.f, 0.f);
                        const float4 o = make_float4(acc[j] + bb.x, acc[j + 1] + bb.y, acc[j + 2] + bb.z, acc[j + 3] + bb.w);
                        if (stream_store) __stcs(reinterpret_cast<float4*>(crow + col), o);   // written once, read later:
                        else *reinterpret_cast<float4*>(crow + col) = o;                     // keep the operands in L2 instead
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

// ---- host ------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn gemm_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(ptr);
    });
    return fn;
}

// [rows, K] fp32 row-major -> boxes of box_rows x 32 floats (128 B), 128-byte swizzle, zero fill past the edges.
// The learner calls with the same buffers every step, so encoded maps are kept in a small direct-mapped cache keyed
// by (pointer, rows, K, box rows); a map describes addresses and shapes only, never contents.
struct GemmMapKey {
    const void* base;
    long long rows;
    int K, box_rows;
    bool operator==(const GemmMapKey& o) const { return base == o.base && rows == o.rows && K == o.K && box_rows == o.box_rows; }
};
struct GemmMapSlot {
    GemmMapKey key;
    CUtensorMap map;
    bool valid;
};
GemmMapSlot g_gemm_maps[64];
std::mutex g_gemm_map_mutex;

bool gemm_map(EncodeTiledFn enc, CUtensorMap* m, const float* base, long long rows, int K, int box_rows) {
    const GemmMapKey key{base, rows, K, box_rows};
    size_t h = (reinterpret_cast<uintptr_t>(base) >> 8) ^ (size_t(rows) * 0x9E3779B97F4A7C15ull) ^ (size_t(box_rows) << 5) ^ size_t(K);
    GemmMapSlot& slot = g_gemm_maps[(h ^ (h >> 23)) & 63];
    std::lock_guard<std::mutex> lock(g_gemm_map_mutex);
    if (slot.valid && slot.key == key) {
        *m = slot.map;
        return true;
    }
    const cuuint64_t gdim[2] = {cuuint64_t(K), cuuint64_t(rows)};
    const cuuint64_t gstride[1] = {cuuint64_t(K) * 4u};
    const cuuint32_t box[2] = {cuuint32_t(kGemmBK), cuuint32_t(box_rows)};
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

// process-wide experiment knobs (mb_set_tuning): atomics, because autograd launches backward kernels from its own threads
static std::atomic<int> g_gemm_group_m{0}, g_gemm_stream_store{0};
void gemm_set_tuning(int group_m, int stream_store) {
    g_gemm_group_m = group_m;
    g_gemm_stream_store = stream_store;
}

int split_tf32(const float* x, float* hi, float* lo, long long n, cudaStream_t st) {
    if (n == 0) return MB_OK;
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    const long long want = (n + 255) / 256, cap = (long long)sm_count(dev) * 8;
    split_tf32_kernel<<<int(want < cap ? want : cap), 256, 0, st>>>(x, hi, lo, n);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}

int linear_tf32x3(const float* a_hi, const float* a_lo, const float* b_hi, const float* b_lo, const float* bias, float* c,
                  long long M, long long N, int K, int* status, cudaStream_t st) {
    EncodeTiledFn enc = gemm_encode_fn();
    if (!enc) {
        set_error("mb_linear_tf32x3: cuTensorMapEncodeTiled is not available from this driver");
        return MB_ERR_UNSUPPORTED;
    }
    GemmMaps maps;
    if (!gemm_map(enc, &maps.a_hi, a_hi, M, K, kGemmBM) || !gemm_map(enc, &maps.a_lo, a_lo, M, K, kGemmBM) ||
        !gemm_map(enc, &maps.b_hi, b_hi, N, K, kGemmBN) || !gemm_map(enc, &maps.b_lo, b_lo, N, K, kGemmBN)) {
        set_error("mb_linear_tf32x3: cuTensorMapEncodeTiled rejected M=%lld N=%lld K=%d", M, N, K);
        return MB_ERR_INVALID;
    }
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    static bool attr_set[64] = {};
    bool& done = attr_set[(dev >= 0 && dev < 64) ? dev : 0];
    if (!done) {
        MB_CUDA(cudaFuncSetAttribute(linear_tf32x3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmem));
        done = true;
    }
    const long long tiles = ((M + kGemmBM - 1) / kGemmBM) * ((N + kGemmBN - 1) / kGemmBN);
    const int sms = sm_count(dev);
    const int grid = int(tiles < sms ? tiles : sms);
    linear_tf32x3_kernel<<<grid, kGemmThreads, kGemmSmem, st>>>(maps, bias, c, int(M), int(N), K, status, g_gemm_group_m,
                                                                g_gemm_stream_store);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}

}  // namespace mb

extern "C" int mb_split_tf32(const float* x, float* hi, float* lo, int64_t n, void* stream) {
    if (n < 0 || (n > 0 && (!x || !hi || !lo))) {
        mb::set_error("mb_split_tf32: bad arguments (n=%lld)", (long long)n);
        return MB_ERR_INVALID;
    }
    return mb::split_tf32(x, hi, lo, n, static_cast<cudaStream_t>(stream));
}

extern "C" int mb_linear_tf32x3(const float* a_hi, const float* a_lo, const float* b_hi, const float* b_lo,
                                const float* bias, float* c, int64_t m, int64_t n, int32_t k, int32_t* status,
                                void* stream) {
    if (m < 0 || n < 0 || k <= 0 || k % 32 != 0 || n % 4 != 0 || m > 0x7fffffff || n > 0x7fffffff) {
        mb::set_error("mb_linear_tf32x3: m=%lld n=%lld k=%d: need k %% 32 == 0 and n %% 4 == 0", (long long)m,
                      (long long)n, k);
        return MB_ERR_INVALID;
    }
    if (m == 0 || n == 0) return MB_OK;
    if (!a_hi || !a_lo || !b_hi || !b_lo || !c) {
        mb::set_error("mb_linear_tf32x3: null pointer");
        return MB_ERR_INVALID;
    }
    const uintptr_t all = reinterpret_cast<uintptr_t>(a_hi) | reinterpret_cast<uintptr_t>(a_lo) |
                          reinterpret_cast<uintptr_t>(b_hi) | reinterpret_cast<uintptr_t>(b_lo) |
                          reinterpret_cast<uintptr_t>(c) | reinterpret_cast<uintptr_t>(bias);
    if (all & 15) {
        mb::set_error("mb_linear_tf32x3: operands must be 16-byte aligned");
        return MB_ERR_INVALID;
    }
    return mb::linear_tf32x3(a_hi, a_lo, b_hi, b_lo, bias, c, m, n, k, status, static_cast<cudaStream_t>(stream));
}
