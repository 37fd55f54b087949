// Backward of actor Linear + policy head over the cells that can have a gradient, for sm_100a.
//
// Reference: loss.backward() (libs/utils.py:334) through model.py:167-214.  d loss / d logits is exactly zero on
// every cell without a valid action (quirk Q1) -- 94 % of the (frame, cell) pairs at the synthetic density, more in
// real games -- so the two backward products of the actor Linear (dW = dlogits^T hidden, dhidden = dlogits W) only
// have to visit the pairs that are alive.  Round 1 did that with torch index / bmm / index_add_ kernels and two
// host synchronisations (2.5 ms of a 12.6 ms step); here the whole chain runs on the device:
//
//   cell_count / cell_scan / cell_fill   per cell, the alive frames in ascending order (deterministic), padded to
//                      chunks of 32 entries; every chunk belongs to one cell; all sizes stay on the device
//                      (a header word holds the chunk count), so nothing synchronises with the host;
//   cell_dz            one thread per entry: logits of the pair (kept by mb_actor_head_fwd), its 16-byte cell record
//                      and the frame's upstream gradients -> dlogits of the pair (the analytic formula of
//                      csrc/head.cu), stored [entry][80] and, per chunk, transposed [80][32];
//   cell_grad          persistent CTAs over contiguous ranges of chunks; per chunk the 32 hidden rows arrive by
//                      1 KB bulk copies (cp.async.bulk), the cell's weight block W_c [78 x 256] sits in shared
//                      memory while consecutive chunks share the cell, and two register-tiled fp32 FMA products
//                      give dhidden (added to the frames' rows with vector atomics) and dW_c / dbias_c
//                      (accumulated in registers, added to the output when the cell changes).
// fp32 FMA throughout (the arithmetic of the reference's CPU matmuls; only the summation order differs).
#include "common.cuh"

#include "../../include/mbeast.h"

namespace mb {

int sm_count(int device);

namespace {

constexpr int kChunk = 32;                 // list entries per chunk
constexpr int kPadJ = 80;                  // logits per cell, padded
constexpr int kHid = 256;                  // width of `hidden` (model.py:130)

__device__ __forceinline__ bool rec_bit(const uint4& m, int i) {
    return i < 32 ? ((m.x >> i) & 1u) : i < 64 ? ((m.y >> (i - 32)) & 1u) : ((m.z >> (i - 64)) & 1u);
}
__device__ __forceinline__ int rec_action(const uint4& m, int k) {
    return k < 3 ? int((m.z >> (14 + 6 * k)) & 63u) : int((m.w >> (6 * (k - 3))) & 63u);
}

// ---- lists -----------------------------------------------------------------------------------
// alive_bits [cells][words]: bit i of word w = frame 32*w + i has a valid action in the cell (written by mb_pack_cells)
__global__ void __launch_bounds__(256) cell_count_kernel(const uint32_t* __restrict__ alive_bits, int* __restrict__ counts, int words) {
    __shared__ int part[8];
    const int c = blockIdx.x;
    const uint32_t* row = alive_bits + (long long)c * words;
    int n = 0;
    for (int w = threadIdx.x; w < words; w += 256) n += __popc(__ldg(row + w));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = n;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < 8; ++w) t += part[w];
        counts[c] = t;
    }
}

// one block: chunk offsets per cell, the cell of every chunk, header = {n_chunks, n_entries_alive}
__global__ void __launch_bounds__(1024) cell_scan_kernel(const int* __restrict__ counts, int* __restrict__ chunk_off,
                                                         int* __restrict__ chunk_cell, int* __restrict__ header, int cells,
                                                         int chunk) {
    __shared__ int s_off[1025];
    __shared__ int s_cnt[1024];
    for (int c = threadIdx.x; c < cells; c += blockDim.x) s_cnt[c] = counts[c];      // one round trip, not one per cell
    __syncthreads();
    if (threadIdx.x == 0) {
        int run = 0, alive = 0;
        for (int c = 0; c < cells; ++c) {            // cells <= 1024: a serial scan of a few hundred adds
            s_off[c] = run;
            run += (s_cnt[c] + chunk - 1) / chunk;
            alive += s_cnt[c];
        }
        s_off[cells] = run;
        header[0] = run;
        header[1] = alive;
    }
    __syncthreads();
    for (int c = threadIdx.x; c < cells; c += blockDim.x) {
        chunk_off[c] = s_off[c];
        for (int k = s_off[c]; k < s_off[c + 1]; ++k) chunk_cell[k] = c;
    }
    if (threadIdx.x == 0) chunk_off[cells] = s_off[cells];
}

// per cell: alive frames in ascending order into entries[chunk_off[c] * 32 ...], tail of the last chunk = -1
__global__ void __launch_bounds__(256) cell_fill_kernel(const uint32_t* __restrict__ alive_bits, const int* __restrict__ chunk_off,
                                                        int* __restrict__ entries, int words, int chunk) {
    __shared__ int warp_tot[8];
    __shared__ int base_s;
    const int c = blockIdx.x;
    const uint32_t* row = alive_bits + (long long)c * words;
    int* out = entries + (long long)chunk_off[c] * chunk;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) base_s = 0;
    __syncthreads();
    for (int w0 = 0; w0 < words; w0 += 256) {
        const int w = w0 + threadIdx.x;
        uint32_t bits = w < words ? __ldg(row + w) : 0u;
        const int n = __popc(bits);
        int incl = n;                                       // inclusive prefix over the warp's 32 words
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) warp_tot[warp] = incl;
        __syncthreads();
        int pos = base_s + incl - n;
        for (int k = 0; k < warp; ++k) pos += warp_tot[k];
        while (bits) {                                      // the word's frames, ascending
            const int i = __ffs(bits) - 1;
            bits &= bits - 1;
            out[pos++] = 32 * w + i;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            int t = 0;
            for (int k = 0; k < 8; ++k) t += warp_tot[k];
            base_s += t;
        }
        __syncthreads();
    }
    const int total = base_s;
    const int padded = (total + chunk - 1) / chunk * chunk;
    for (int i = total + threadIdx.x; i < padded; i += 256) out[i] = -1;
}

// ---- dlogits of every list entry ---------------------------------------------------------------
// d/dz_i [g_lp * logp + g_h * H] = p_i * (-g_lp - g_h * (l_i + H)) + g_lp * [i == a]; masked entries have p_i = 0 and
// a fully masked component gets exactly 0 (same formula and SFU approximations as csrc/head.cu).
template <int OFF, int LEN>
__device__ __forceinline__ void seg_backward(float (&z)[kPadJ], const uint4& meta, int a, float g_lp, float g_h) {
    float zz[LEN], e[LEN];
    float m = -3.4e38f;
#pragma unroll
    for (int i = 0; i < LEN; ++i) {
        zz[i] = rec_bit(meta, OFF + i) ? z[OFF + i] : kMaskFill;
        m = fmaxf(m, zz[i]);
    }
    float S4[4] = {0.f, 0.f, 0.f, 0.f}, W4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < LEN; ++i) {
        const float d = zz[i] - m;
        e[i] = ex2_approx(d * kLog2e);
        S4[i & 3] += e[i];
        W4[i & 3] = fmaf(e[i], d, W4[i & 3]);
    }
    const float S = (S4[0] + S4[1]) + (S4[2] + S4[3]), W = (W4[0] + W4[1]) + (W4[2] + W4[3]);
    const float logS = lg2_approx(S) * kLn2;
    const float lse = m + logS;
    const bool dead = m < -5e7f;
    const float ent = (dead ? (lse - m) : logS) - __fdividef(W, S);
    const float invS = dead ? 0.f : __fdividef(1.f, S);
    const float c0 = -g_lp - g_h * ent;
#pragma unroll
    for (int i = 0; i < LEN; ++i) {
        const float l = zz[i] - lse;
        float v = (e[i] * invS) * fmaf(-g_h, l, c0);
        v += (i == a && rec_bit(meta, OFF + i)) ? g_lp : 0.f;
        z[OFF + i] = v;
    }
}

// dlogits of one list entry into shared memory, row-major [32][80] and transposed [80][32] (one lane per entry)
__device__ __forceinline__ void entry_dz(const float* __restrict__ logits, const uint4* __restrict__ meta,
                                         const float* __restrict__ g_logp, const float* __restrict__ g_ent, int f, int c,
                                         int n_frames, int cells, float* Dz, float* DzT, int slot) {
    float z[kPadJ];
#pragma unroll
    for (int i = 0; i < kPadJ; ++i) z[i] = 0.f;
    if (f >= 0) {
        const float2* src = reinterpret_cast<const float2*>(logits + ((long long)f * cells + c) * kLogitsPerCell);
#pragma unroll
        for (int i = 0; i < kLogitsPerCell / 2; ++i) {
            const float2 v = __ldg(src + i);
            z[2 * i] = v.x;
            z[2 * i + 1] = v.y;
        }
        const uint4 mt = __ldg(meta + (long long)c * n_frames + f);
        const float gl = __ldg(g_logp + f), gh = __ldg(g_ent + f);
        seg_backward<0, 6>(z, mt, rec_action(mt, 0), gl, gh);
        seg_backward<6, 4>(z, mt, rec_action(mt, 1), gl, gh);
        seg_backward<10, 4>(z, mt, rec_action(mt, 2), gl, gh);
        seg_backward<14, 4>(z, mt, rec_action(mt, 3), gl, gh);
        seg_backward<18, 4>(z, mt, rec_action(mt, 4), gl, gh);
        seg_backward<22, 7>(z, mt, rec_action(mt, 5), gl, gh);
        seg_backward<29, 49>(z, mt, rec_action(mt, 6), gl, gh);
        z[78] = 0.f;
        z[79] = 0.f;
    }
    float4* row = reinterpret_cast<float4*>(Dz + slot * kPadJ);
#pragma unroll
    for (int i = 0; i < kPadJ / 4; ++i) row[i] = make_float4(z[4 * i], z[4 * i + 1], z[4 * i + 2], z[4 * i + 3]);
#pragma unroll
    for (int j = 0; j < kPadJ; ++j) DzT[j * kChunk + slot] = z[j];
}

// ---- the two products ------------------------------------------------------------------------
constexpr int kGradThreads = 288;               // 8 product warps (warp 0 also issues the row copies) + 1 dlogits warp
constexpr int kWBytes = kPadJ * kHid * 4;                       // 80 KB: W_c, rows 78 and 79 zero
constexpr int kHBytes = kChunk * kHid * 4;                      // 32 KB
constexpr int kDzBytes = kChunk * kPadJ * 4;                    // 10 KB, twice (row-major and transposed)
constexpr int kGradStage = kHBytes + 2 * kDzBytes;
constexpr int kGradSmem = kWBytes + 2 * kGradStage + 128;

__global__ void __launch_bounds__(kGradThreads, 1) cell_grad_kernel(const float* __restrict__ hidden, const float* __restrict__ weight,
                                                                    const float* __restrict__ logits, const uint4* __restrict__ meta,
                                                                    const float* __restrict__ g_logp, const float* __restrict__ g_ent,
                                                                    const int* __restrict__ entries, const int* __restrict__ chunk_cell,
                                                                    const int* __restrict__ header, float* __restrict__ d_hidden,
                                                                    float* __restrict__ d_weight, float* __restrict__ d_bias,
                                                                    int n_frames, int cells) {
    extern __shared__ __align__(128) uint8_t cg_smem[];
    __shared__ __align__(8) uint64_t full[2];
    __shared__ __align__(8) uint64_t wbar;
    __shared__ int s_frame[2][kChunk];
    float* Wc = reinterpret_cast<float*>(cg_smem);
    uint8_t* stages = cg_smem + kWBytes;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n_chunks = header[0];
    const int first = int((long long)n_chunks * blockIdx.x / gridDim.x);
    const int last = int((long long)n_chunks * (blockIdx.x + 1) / gridDim.x);
    if (tid == 0) {
        mbar_init(&full[0], 1);
        mbar_init(&full[1], 1);
        mbar_init(&wbar, 1);
        fence_mbar_init();
    }
    for (int i = tid; i < 2 * kHid; i += kGradThreads) Wc[kLogitsPerCell * kHid + i] = 0.f;      // rows 78, 79
    __syncthreads();
    if (first >= last) return;

    auto issue_chunk = [&](int chunk, int s) {           // warp 0: the 32 hidden rows of the chunk, 1 KB each
        uint8_t* st = stages + s * kGradStage;
        const int f = entries[(long long)chunk * kChunk + lane];
        s_frame[s][lane] = f;
        if (lane == 0) mbar_arrive_expect_tx(&full[s], uint32_t(kHBytes));
        __syncwarp();
        // a padded entry has dz == 0; its hidden row is still read (frame 0: finite values, 0 * x stays 0)
        bulk_g2s(st + lane * (kHid * 4), hidden + (long long)(f < 0 ? 0 : f) * kHid, kHid * 4, &full[s]);
    };
    auto make_dz = [&](int chunk, int s) {               // warp 8: dlogits of the chunk's 32 entries, one lane each
        uint8_t* st = stages + s * kGradStage;
        entry_dz(logits, meta, g_logp, g_ent, entries[(long long)chunk * kChunk + lane], chunk_cell[chunk], n_frames, cells,
                 reinterpret_cast<float*>(st + kHBytes), reinterpret_cast<float*>(st + kHBytes + kDzBytes), lane);
    };

    // this thread's 8 hidden columns: 4*lane .. 4*lane+3 and 128 + 4*lane .. +3 (a warp's 128-bit loads and vector
    // atomics then cover one contiguous 512-B run each: no bank conflicts, full sectors)
    const int c0 = 4 * lane, c1 = kHid / 2 + 4 * lane;
    float accW[10][8];                   // dW_c rows 10*warp .. 10*warp+9, this thread's 8 columns
    float accB[10];
#pragma unroll
    for (int a = 0; a < 10; ++a) {
        accB[a] = 0.f;
#pragma unroll
        for (int b = 0; b < 8; ++b) accW[a][b] = 0.f;
    }
    int cur_cell = -1;
    uint32_t wphase = 0;

    auto flush = [&]() {                 // add this CTA's share of dW_c / dbias_c to the outputs
        if (cur_cell < 0) return;
#pragma unroll
        for (int a = 0; a < 10; ++a) {
            const int j = 10 * warp + a;
            if (j < kLogitsPerCell) {
                float* dst = d_weight + ((long long)cur_cell * kLogitsPerCell + j) * kHid;
                atomicAdd(reinterpret_cast<float4*>(dst + c0), make_float4(accW[a][0], accW[a][1], accW[a][2], accW[a][3]));
                atomicAdd(reinterpret_cast<float4*>(dst + c1), make_float4(accW[a][4], accW[a][5], accW[a][6], accW[a][7]));
                if (lane == 0) atomicAdd(d_bias + cur_cell * kLogitsPerCell + j, accB[a]);
            }
            accB[a] = 0.f;
#pragma unroll
            for (int b = 0; b < 8; ++b) accW[a][b] = 0.f;
        }
    };

    if (warp == 0) issue_chunk(first, 0);
    if (warp == 8) make_dz(first, 0);
    __syncthreads();
    for (int chunk = first; chunk < last; ++chunk) {
        const int s = (chunk - first) & 1;
        const int cell = chunk_cell[chunk];
        if (cell != cur_cell) {                          // rare: a CTA's range spans two or three cells
            flush();
            __syncthreads();                             // nobody still reads the old W_c
            if (tid == 0) {
                mbar_arrive_expect_tx(&wbar, uint32_t(kLogitsPerCell * kHid * 4));
                bulk_g2s(Wc, weight + (long long)cell * kLogitsPerCell * kHid, kLogitsPerCell * kHid * 4, &wbar);
            }
            mbar_wait(&wbar, wphase, nullptr);
            wphase ^= 1;
            cur_cell = cell;
        }
        if (warp == 0 && chunk + 1 < last) issue_chunk(chunk + 1, s ^ 1);
        if (warp == 8) {                                 // the next chunk's dlogits while the others multiply this one
            if (chunk + 1 < last) make_dz(chunk + 1, s ^ 1);
            __syncthreads();
            continue;
        }
        mbar_wait(&full[s], uint32_t(((chunk - first) >> 1) & 1), nullptr);
        const uint8_t* st = stages + s * kGradStage;
        const float* H = reinterpret_cast<const float*>(st);
        const float* Dz = reinterpret_cast<const float*>(st + kHBytes);              // [32][80]
        const float* DzT = reinterpret_cast<const float*>(st + kHBytes + kDzBytes);  // [80][32]

        // dhidden rows of entries 4*warp .. 4*warp+3: sum_j dz[e][j] * W_c[j][k]
        {
            float acc[4][8];
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 8; ++b) acc[a][b] = 0.f;
#pragma unroll 3
            for (int j = 0; j < kLogitsPerCell; ++j) {
                const float4 d4 = *reinterpret_cast<const float4*>(DzT + j * kChunk + 4 * warp);
                const float4 w0 = *reinterpret_cast<const float4*>(Wc + j * kHid + c0);
                const float4 w1 = *reinterpret_cast<const float4*>(Wc + j * kHid + c1);
                const float d[4] = {d4.x, d4.y, d4.z, d4.w};
                const float w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int b = 0; b < 8; ++b) acc[a][b] = fmaf(d[a], w[b], acc[a][b]);
            }
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                const int f = s_frame[s][4 * warp + a];
                if (f >= 0) {
                    float* dst = d_hidden + (long long)f * kHid;
                    atomicAdd(reinterpret_cast<float4*>(dst + c0), make_float4(acc[a][0], acc[a][1], acc[a][2], acc[a][3]));
                    atomicAdd(reinterpret_cast<float4*>(dst + c1), make_float4(acc[a][4], acc[a][5], acc[a][6], acc[a][7]));
                }
            }
        }
        // dW_c rows 10*warp .. 10*warp+9: sum_e dz[e][j] * hidden[f_e][k]
#pragma unroll 2
        for (int e = 0; e < kChunk; ++e) {
            const float2* dp = reinterpret_cast<const float2*>(Dz + e * kPadJ + 10 * warp);
            const float4 h0 = *reinterpret_cast<const float4*>(H + e * kHid + c0);
            const float4 h1 = *reinterpret_cast<const float4*>(H + e * kHid + c1);
            const float h[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
            float d[10];
#pragma unroll
            for (int a = 0; a < 5; ++a) {
                const float2 v = dp[a];
                d[2 * a] = v.x;
                d[2 * a + 1] = v.y;
            }
#pragma unroll
            for (int a = 0; a < 10; ++a) {
                accB[a] += d[a];
#pragma unroll
                for (int b = 0; b < 8; ++b) accW[a][b] = fmaf(d[a], h[b], accW[a][b]);
            }
        }
        __syncthreads();                                 // the stage may be refilled, s_frame rewritten
    }
    if (warp < 8) flush();
}

}  // namespace

size_t cell_list_capacity_chunks(long long n_frames, int cells) {
    return size_t(cells) * size_t((n_frames + kChunk - 1) / kChunk);
}

}  // namespace mb

namespace mb {

static size_t cells_ws_ints(long long n_frames, int cells) {
    const size_t cap = cell_list_capacity_chunks(n_frames, cells);
    return size_t(cells) + size_t(cells) + 1 + 4 + cap + cap * kChunk;
}

}  // namespace mb

namespace mb {
// Cell-ordered lists of the frames with a valid action, `chunk` entries per chunk (padding = -1), for other kernels of
// the library (csrc/fused.cu: the list form of the actor head's forward).  Layout in `ws` (ints): counts [cells],
// chunk_off [cells + 1], header [4] = {n_chunks, n_alive, ..}, chunk_cell [cap], entries [cap * chunk] with
// cap = n_frames * cells / chunk + cells.
size_t cell_lists_ws_ints(long long n_frames, int cells, int chunk) {
    const size_t cap = size_t(n_frames) * cells / chunk + cells;
    return size_t(cells) + cells + 1 + 4 + cap + cap * chunk;
}
int build_cell_lists(const uint32_t* alive_bits, int* ws, long long n_frames, int cells, int chunk, cudaStream_t st,
                     int** chunk_cell_out, int** header_out, int** entries_out) {
    const size_t cap = size_t(n_frames) * cells / chunk + cells;
    int* counts = ws;
    int* chunk_off = counts + cells;
    int* header = chunk_off + cells + 1;
    int* chunk_cell = header + 4;
    int* entries = chunk_cell + cap;
    const int words = int((n_frames + 31) / 32);
    cell_count_kernel<<<cells, 256, 0, st>>>(alive_bits, counts, words);
    cell_scan_kernel<<<1, 1024, 0, st>>>(counts, chunk_off, chunk_cell, header, cells, chunk);
    cell_fill_kernel<<<cells, 256, 0, st>>>(alive_bits, chunk_off, entries, words, chunk);
    MB_CUDA(cudaGetLastError());
    *chunk_cell_out = chunk_cell;
    *header_out = header;
    *entries_out = entries;
    return MB_OK;
}
}  // namespace mb

extern "C" size_t mb_cells_bwd_workspace_bytes(int64_t n_frames, int32_t cells) {
    if (n_frames <= 0 || cells <= 0) return 0;
    return (mb::cells_ws_ints(n_frames, cells) * sizeof(int) + 255) & ~size_t(255);
}

extern "C" int mb_cells_bwd(const float* hidden, const float* weight, const float* logits, const void* cell_meta,
                            const void* alive_bits, const float* g_logp, const float* g_ent, float* d_hidden, float* d_weight,
                            float* d_bias,
                            void* workspace, size_t workspace_bytes, int64_t n_frames, int32_t cells, int32_t k, void* stream) {
    using namespace mb;
    if (n_frames < 0 || cells <= 0 || cells > 1024 || k != kHid || n_frames > 0x7fffffff) {
        set_error("mb_cells_bwd: n_frames=%lld cells=%d k=%d: need 0 < cells <= 1024 and k == 256", (long long)n_frames, cells, k);
        return MB_ERR_INVALID;
    }
    if (!d_hidden || !d_weight || !d_bias) {
        set_error("mb_cells_bwd: null output pointer");
        return MB_ERR_INVALID;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    MB_CUDA(cudaMemsetAsync(d_weight, 0, size_t(cells) * kLogitsPerCell * kHid * sizeof(float), st));
    MB_CUDA(cudaMemsetAsync(d_bias, 0, size_t(cells) * kLogitsPerCell * sizeof(float), st));
    if (n_frames == 0) return MB_OK;
    MB_CUDA(cudaMemsetAsync(d_hidden, 0, size_t(n_frames) * kHid * sizeof(float), st));
    if (!hidden || !weight || !logits || !cell_meta || !alive_bits || !g_logp || !g_ent || !workspace) {
        set_error("mb_cells_bwd: null input pointer");
        return MB_ERR_INVALID;
    }
    const uintptr_t all = reinterpret_cast<uintptr_t>(hidden) | reinterpret_cast<uintptr_t>(weight) |
                          reinterpret_cast<uintptr_t>(cell_meta) | reinterpret_cast<uintptr_t>(d_hidden) |
                          reinterpret_cast<uintptr_t>(d_weight) | reinterpret_cast<uintptr_t>(workspace);
    if ((all & 15) || (reinterpret_cast<uintptr_t>(logits) & 7)) {
        set_error("mb_cells_bwd: hidden / weight / cell_meta / d_hidden / d_weight / workspace must be 16-byte aligned");
        return MB_ERR_INVALID;
    }
    if (workspace_bytes < mb_cells_bwd_workspace_bytes(n_frames, cells)) {
        set_error("mb_cells_bwd: workspace smaller than mb_cells_bwd_workspace_bytes(n_frames, cells)=%zu",
                  mb_cells_bwd_workspace_bytes(n_frames, cells));
        return MB_ERR_INVALID;
    }
    const size_t cap = cell_list_capacity_chunks(n_frames, cells);
    int* counts = reinterpret_cast<int*>(workspace);
    int* chunk_off = counts + cells;
    int* header = chunk_off + cells + 1;
    int* chunk_cell = header + 4;
    int* entries = chunk_cell + cap;
    const uint4* meta = reinterpret_cast<const uint4*>(cell_meta);
    const uint32_t* bits = reinterpret_cast<const uint32_t*>(alive_bits);
    const int words = int((n_frames + 31) / 32);
    cell_count_kernel<<<cells, 256, 0, st>>>(bits, counts, words);
    cell_scan_kernel<<<1, 1024, 0, st>>>(counts, chunk_off, chunk_cell, header, cells, kChunk);
    cell_fill_kernel<<<cells, 256, 0, st>>>(bits, chunk_off, entries, words, kChunk);
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    static bool attr_set[64] = {};
    bool& done = attr_set[(dev >= 0 && dev < 64) ? dev : 0];
    if (!done) {
        MB_CUDA(cudaFuncSetAttribute(cell_grad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kGradSmem));
        done = true;
    }
    cell_grad_kernel<<<sm_count(dev), kGradThreads, kGradSmem, st>>>(hidden, weight, logits, meta, g_logp, g_ent, entries, chunk_cell,
                                                                    header, d_hidden, d_weight, d_bias, int(n_frames), cells);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}
