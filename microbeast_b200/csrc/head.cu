// GridNet masked multi-discrete policy head for sm_100a.
//
// Replaces the 7*C-object Python loop of reference model.py:165-216 (Agent.get_action) and the
// per-component math of model.py:33-52 (CategoricalMasked: where(mask, z, -1e8) -> Categorical
// normalisation -> log_prob / masked entropy / sample), plus its autograd.
//
// Data layout.  logits [N, 78*C] f32, mask [N, 78*C] u8, action [N, 7*C] i64 and uniforms
// [N, 7*C] f32 are all contiguous, i.e. arrays of *cells*: 312 B + 78 B + 56 B (+ 28 B) per cell.
// A CTA streams tiles of TC=64 consecutive cells of one frame through a 3-stage shared-memory
// ring: one elected producer thread issues 1-D bulk copies (cp.async.bulk -> SASS UBLKCP) that
// complete on an mbarrier; six consumer warps read the tile out of shared memory with
// thread-per-cell (29 small-component logits) and lane-pair-per-cell (the 49-way component)
// mappings, so every global byte is fetched exactly once, fully coalesced, with no register
// staging.  Backward / sampling outputs are written in place into the stage and leave through
// bulk stores.  Per-frame sums are reduced in a fixed order (lane shuffle tree -> per-warp
// partial -> producer thread), so results are run-to-run deterministic and need no atomics.
//
// Roofline: HBM-bound.  Algorithmic bytes per cell: eval fwd 446, eval bwd 758, sample 474
// (+8 / +8 / +4..8 per frame) -- SURVEY.md section 8(d).
#include "common.cuh"

#include <math.h>

namespace mb {

enum HeadMode : int { kEvalFwd = 0, kEvalBwd = 1, kSample = 2 };

struct HeadParams {
    const float* logits;       // [N, 78C]
    const uint8_t* mask;       // [N, 78C]   non-zero = valid (model.py:40)
    const int64_t* action_in;  // [N, 7C]    eval modes
    const float* uniforms;     // [N, 7C]    sample mode
    const float* g_logp;       // [N]        bwd
    const float* g_ent;        // [N]        bwd
    float* logp_out;           // [N]
    float* ent_out;            // [N] (may be null in sample mode)
    float* dlogits;            // [N, 78C]   bwd
    int64_t* action_out;       // [N, 7C]    sample
    int* status;               // optional device status word
    long long n_frames;
    int cells;
    int tiles_per_frame;
};

template <int TC, int MODE>
struct HeadCfg {
    static constexpr int kGroups = TC / 32;            // 32-cell groups per tile
    static constexpr int kConsWarps = 3 * kGroups;     // A, B0, B1 per group
    static constexpr int kThreads = 32 * (1 + kConsWarps);
    static constexpr int kLogitBytes = TC * kLogitsPerCell * 4;
    static constexpr int kMaskBytes = TC * kLogitsPerCell;
    static constexpr int kActBytes = TC * kCompsPerCell * 8;
    static constexpr int kUniBytes = (MODE == kSample) ? TC * kCompsPerCell * 4 : 0;
    static constexpr int kOffMask = kLogitBytes;
    static constexpr int kOffAct = kOffMask + kMaskBytes;
    static constexpr int kOffUni = kOffAct + kActBytes;
    static constexpr int kStageBytes = kOffUni + kUniBytes;
    static_assert(kStageBytes % 128 == 0, "stage must stay 128-B aligned");
    static_assert(TC % 32 == 0 && TC % 8 == 0, "tile granularity");
};

template <int TC, int MODE, int NSTAGE>
constexpr size_t head_smem_bytes() {
    using Cfg = HeadCfg<TC, MODE>;
    return size_t(NSTAGE) * Cfg::kStageBytes + size_t(NSTAGE) * Cfg::kConsWarps * 2 * sizeof(float) +
           2 * NSTAGE * sizeof(uint64_t) + 128;
}

// ---------------------------------------------------------------------------------------------
// fp64 restatement of the sampler definition (oracle/head.py: sample_invcdf), used only when the
// fp32 fast path cannot decide (u within the guard band of a CDF edge).
//   p = softmax(l), l = z' - logsumexp(z')  in fp64;  cdf sequential;  a = #{k<n-1 : u >= cdf_k};
//   if the component has a valid entry and mask[a]==0, a = last valid index.
// ---------------------------------------------------------------------------------------------
__device__ __noinline__ int sample_component_fp64(const float* z, const uint8_t* mk, int len, float u) {
    double m1 = -1e300;
    for (int i = 0; i < len; ++i) {
        double zi = mk[i] ? double(z[i]) : -1e8;
        m1 = fmax(m1, zi);
    }
    double s1 = 0.0;
    for (int i = 0; i < len; ++i) {
        double zi = mk[i] ? double(z[i]) : -1e8;
        s1 += exp(zi - m1);
    }
    const double lse = m1 + log(s1);
    // second normalisation (Categorical.probs = softmax(normalised logits)); max(l) = m1 - lse
    const double m2 = m1 - lse;
    double s2 = 0.0;
    for (int i = 0; i < len; ++i) {
        double zi = mk[i] ? double(z[i]) : -1e8;
        s2 += exp((zi - lse) - m2);
    }
    const double ud = double(u);
    double cdf = 0.0;
    int a = 0, last_valid = -1;
    for (int i = 0; i < len; ++i) {
        double zi = mk[i] ? double(z[i]) : -1e8;
        if (mk[i]) last_valid = i;
        if (i < len - 1) {
            cdf += exp((zi - lse) - m2) / s2;
            a += (ud >= cdf) ? 1 : 0;
        }
    }
    if (last_valid >= 0 && mk[a] == 0) a = last_valid;
    return a;
}

// Statistics of one component held in registers: max, sum of exp, sum of exp * (z - max).
template <int LEN>
__device__ __forceinline__ void seg_stats(const float* zp, float* e, float& m, float& S, float& W) {
    m = zp[0];
#pragma unroll
    for (int i = 1; i < LEN; ++i) m = fmaxf(m, zp[i]);
    S = 0.f;
    W = 0.f;
#pragma unroll
    for (int i = 0; i < LEN; ++i) {
        const float d = zp[i] - m;
        const float ei = ex2_approx(d * kLog2e);
        e[i] = ei;
        S += ei;
        W = fmaf(ei, d, W);
    }
}

struct SegScalars {
    float lse;   // logsumexp of the masked logits (torch: max + log(sum exp(z - max)))
    float ent;   // masked entropy (model.py:46-52)
    bool dead;   // every entry masked (quirk Q1): log_prob 0, entropy 0, gradient 0, uniform sample
};

__device__ __forceinline__ SegScalars seg_finish(float m, float S, float W) {
    SegScalars r;
    const float logS = logf(S);
    r.lse = m + logS;
    r.dead = (m < -5e7f);
    // fully masked: lse = fl(-1e8 + ln n) = -1e8 in fp32, so lse - m == 0 and W == 0  ->  0 (Q1)
    const float h0 = r.dead ? (r.lse - m) : logS;
    r.ent = h0 - W / S;
    return r;
}

__device__ __forceinline__ int check_action(long long a, int len, int* status) {
    if (a < 0 || a >= len) {
        if (status) atomicOr(status, kStatusBadAction);
        a = a < 0 ? 0 : len - 1;
    }
    return int(a);
}

// Sequential inverse-CDF count over e[0..CNT) starting from running sum c; returns the number of
// k with t >= c_k and flags `near` when t is inside the guard band of any edge.
template <int CNT>
__device__ __forceinline__ int cdf_count(const float* e, float& c, float t, float guard, bool& near) {
    int a = 0;
#pragma unroll
    for (int i = 0; i < CNT; ++i) {
        c += e[i];
        a += (t >= c) ? 1 : 0;
        near = near || (fabsf(t - c) <= guard);
    }
    return a;
}

__device__ __forceinline__ float sample_guard(float S, float W, int len) {
    // fp32 error budget of the running CDF against the fp64 definition, with a 4x margin:
    // ex2.approx (2 ulp) + argument rounding (|d| * 2^-24 each) + len sequential adds + u*S.
    return S * (9.5367431640625e-07f + float(len + 1) * 2.384185791015625e-07f) + 4.76837158203125e-07f * fabsf(W);
}

// ---------------------------------------------------------------------------------------------
// A role: one thread per cell, the six small components (logit offsets 0..28).
// ---------------------------------------------------------------------------------------------
template <int MODE, int OFF, int LEN, int K>
__device__ __forceinline__ void a_role_component(float* z, float* zc, const uint8_t* mc, const int64_t* ain,
                                                 int64_t* aout, const float* uni, float g_lp, float g_h,
                                                 int* status, float& lp_acc, float& en_acc, int& taken,
                                                 bool& taken_valid) {
    float e[LEN];
    float m, S, W;
    seg_stats<LEN>(z + OFF, e, m, S, W);
    const SegScalars sc = seg_finish(m, S, W);
    int a;
    if (MODE == kSample) {
        const float u = uni[K];
        const float t = u * S;
        float c = 0.f;
        bool near = false;
        a = cdf_count<LEN - 1>(e, c, t, sample_guard(S, W, LEN), near);
        if (near) a = sample_component_fp64(zc + OFF, mc + OFF, LEN, u);
        aout[K] = a;
    } else {
        a = check_action(ain[K], LEN, status);
    }
    // z'_a re-read from shared memory (dynamic index); sentinel when the taken entry is masked (Q2)
    taken = OFF + a;
    taken_valid = mc[OFF + a] != 0;
    const float za = taken_valid ? zc[OFF + a] : kMaskFill;
    lp_acc += za - sc.lse;
    en_acc += sc.ent;
    if (MODE == kEvalBwd) {
        // d/dz_i [g_lp*logp + g_h*H] = p_i*(-g_lp - g_h*(l_i + H)) + g_lp*[i==a]; masked entries have
        // p_i == 0; a fully masked component gets exactly 0 (verified vs autograd)
        const float invS = sc.dead ? 0.f : 1.f / S;
        const float c0 = -g_lp - g_h * sc.ent;
#pragma unroll
        for (int i = 0; i < LEN; ++i) {
            const float l = z[OFF + i] - sc.lse;
            z[OFF + i] = (e[i] * invS) * fmaf(-g_h, l, c0);
        }
    }
}

template <int MODE>
__device__ __forceinline__ void a_role_cell(float* zc, const uint8_t* mc, const int64_t* ain, int64_t* aout,
                                            const float* uni, float g_lp, float g_h, int* status, float& lp,
                                            float& en) {
    float z[29];
#pragma unroll
    for (int j = 0; j < 14; ++j) {
        const float2 v = *reinterpret_cast<const float2*>(zc + 2 * j);
        const uint32_t h = *reinterpret_cast<const uint16_t*>(mc + 2 * j);
        z[2 * j] = (h & 0xffu) ? v.x : kMaskFill;
        z[2 * j + 1] = (h > 0xffu) ? v.y : kMaskFill;
    }
    z[28] = mc[28] ? zc[28] : kMaskFill;

    int a_k[6];
    bool v_k[6];
    a_role_component<MODE, 0, 6, 0>(z, zc, mc, ain, aout, uni, g_lp, g_h, status, lp, en, a_k[0], v_k[0]);
    a_role_component<MODE, 6, 4, 1>(z, zc, mc, ain, aout, uni, g_lp, g_h, status, lp, en, a_k[1], v_k[1]);
    a_role_component<MODE, 10, 4, 2>(z, zc, mc, ain, aout, uni, g_lp, g_h, status, lp, en, a_k[2], v_k[2]);
    a_role_component<MODE, 14, 4, 3>(z, zc, mc, ain, aout, uni, g_lp, g_h, status, lp, en, a_k[3], v_k[3]);
    a_role_component<MODE, 18, 4, 4>(z, zc, mc, ain, aout, uni, g_lp, g_h, status, lp, en, a_k[4], v_k[4]);
    a_role_component<MODE, 22, 7, 5>(z, zc, mc, ain, aout, uni, g_lp, g_h, status, lp, en, a_k[5], v_k[5]);

    if (MODE == kEvalBwd) {
        // z[] now holds dlogits; write it over the logits in the stage, then add the one-hot term
#pragma unroll
        for (int j = 0; j < 14; ++j) *reinterpret_cast<float2*>(zc + 2 * j) = make_float2(z[2 * j], z[2 * j + 1]);
        zc[28] = z[28];
#pragma unroll
        for (int k = 0; k < 6; ++k)
            if (v_k[k]) zc[a_k[k]] += g_lp;
    }
}

// ---------------------------------------------------------------------------------------------
// B role: a lane pair per cell, the 49-way component (logit offsets 29..77).
// even lane: element 29 + pairs 30..53 (25 entries); odd lane: pairs 54..77 (24 entries).
// ---------------------------------------------------------------------------------------------
template <int MODE>
__device__ __forceinline__ void b_role_cell(bool active, int half, float* zc, const uint8_t* mc,
                                            const int64_t* ain, int64_t* aout, const float* uni, float g_lp,
                                            float g_h, int* status, float& lp, float& en) {
    constexpr int OFF = 29, LEN = 49;
    const int pb = 30 + 24 * half;
    const bool has_x = (half == 0);
    float z[24], e[24];
    float zx = kMaskFill, ex = 0.f;
    if (active) {
#pragma unroll
        for (int j = 0; j < 12; ++j) {
            const float2 v = *reinterpret_cast<const float2*>(zc + pb + 2 * j);
            const uint32_t h = *reinterpret_cast<const uint16_t*>(mc + pb + 2 * j);
            z[2 * j] = (h & 0xffu) ? v.x : kMaskFill;
            z[2 * j + 1] = (h > 0xffu) ? v.y : kMaskFill;
        }
        if (has_x) zx = mc[OFF] ? zc[OFF] : kMaskFill;
    } else {
#pragma unroll
        for (int j = 0; j < 24; ++j) z[j] = kMaskFill;
    }
    float m = z[0];
#pragma unroll
    for (int i = 1; i < 24; ++i) m = fmaxf(m, z[i]);
    if (has_x) m = fmaxf(m, zx);
    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
    float S = 0.f, W = 0.f;
    if (has_x) {
        const float d = zx - m;
        ex = ex2_approx(d * kLog2e);
        S = ex;
        W = ex * d;
    }
#pragma unroll
    for (int i = 0; i < 24; ++i) {
        const float d = z[i] - m;
        const float ei = ex2_approx(d * kLog2e);
        e[i] = ei;
        S += ei;
        W = fmaf(ei, d, W);
    }
    const float S_own = S;
    const float S_other = __shfl_xor_sync(0xffffffffu, S, 1);
    S = has_x ? (S_own + S_other) : (S_other + S_own);     // same operand order on both lanes
    const float W_other = __shfl_xor_sync(0xffffffffu, W, 1);
    W = has_x ? (W + W_other) : (W_other + W);
    const SegScalars sc = seg_finish(m, S, W);

    int a = 0;
    if (MODE == kSample) {
        const float u = active ? uni[6] : 0.5f;
        const float t = u * S;
        const float guard = sample_guard(S, W, LEN);
        bool near = false;
        int cnt;
        if (has_x) {
            float c = ex;
            cnt = (t >= c) ? 1 : 0;
            near = fabsf(t - c) <= guard;
            cnt += cdf_count<24>(e, c, t, guard, near);
        } else {
            float c = S_other;                         // sum of the even lane's 25 entries
            cnt = cdf_count<23>(e, c, t, guard, near);  // entry 48 (k = n-1) is not counted
        }
        cnt += __shfl_xor_sync(0xffffffffu, cnt, 1);
        near = __shfl_xor_sync(0xffffffffu, int(near), 1) || near;
        a = cnt;
        if (near && has_x && active) a = sample_component_fp64(zc + OFF, mc + OFF, LEN, u);
        a = __shfl_sync(0xffffffffu, a, (threadIdx.x & 31) & ~1);
        if (has_x && active) aout[6] = a;
    } else if (active) {
        a = check_action(ain[6], LEN, has_x ? status : nullptr);
    }
    if (active && has_x) {
        const float za = mc[OFF + a] ? zc[OFF + a] : kMaskFill;
        lp += za - sc.lse;
        en += sc.ent;
    }
    if (MODE == kEvalBwd) {
        const bool taken_valid = active && (mc[OFF + a] != 0);
        __syncwarp();       // the even lane has read z'_a (possibly in the odd lane's range) above
        if (active) {
            const float invS = sc.dead ? 0.f : 1.f / S;
            const float c0 = -g_lp - g_h * sc.ent;
#pragma unroll
            for (int j = 0; j < 12; ++j) {
                const float l0 = z[2 * j] - sc.lse, l1 = z[2 * j + 1] - sc.lse;
                *reinterpret_cast<float2*>(zc + pb + 2 * j) =
                    make_float2((e[2 * j] * invS) * fmaf(-g_h, l0, c0), (e[2 * j + 1] * invS) * fmaf(-g_h, l1, c0));
            }
            if (has_x) zc[OFF] = (ex * invS) * fmaf(-g_h, zx - sc.lse, c0);
        }
        __syncwarp();
        if (has_x && taken_valid) zc[OFF + a] += g_lp;
    }
}

// ---------------------------------------------------------------------------------------------
// The kernel.
// ---------------------------------------------------------------------------------------------
template <int TC, int MODE, int NSTAGE>
__global__ void __launch_bounds__(HeadCfg<TC, MODE>::kThreads) head_kernel(const HeadParams p) {
    using Cfg = HeadCfg<TC, MODE>;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* stages = smem;
    float* part = reinterpret_cast<float*>(smem + size_t(NSTAGE) * Cfg::kStageBytes);   // [NSTAGE][kConsWarps][2]
    uint64_t* full = reinterpret_cast<uint64_t*>(part + NSTAGE * Cfg::kConsWarps * 2);
    uint64_t* done = full + NSTAGE;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < NSTAGE; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&done[s], Cfg::kConsWarps);
        }
        fence_mbar_init();
    }
    __syncthreads();

    const long long first = blockIdx.x;
    if (first >= p.n_frames) return;
    const long long nf_cta = (p.n_frames - first + gridDim.x - 1) / gridDim.x;
    const long long n_tiles = nf_cta * p.tiles_per_frame;
    const int tpf = p.tiles_per_frame;

    auto tile_frame = [&](long long j) -> long long { return first + (j / tpf) * (long long)gridDim.x; };
    auto tile_cell0 = [&](long long j) -> int { return int(j % tpf) * TC; };

    if (warp == 0) {
        // ===================== producer / retire warp =====================
        float acc_lp = 0.f, acc_en = 0.f;
        bool last_store_bulk = true;

        auto load_tile = [&](long long j) {
            const int s = int(j % NSTAGE);
            uint8_t* st = stages + size_t(s) * Cfg::kStageBytes;
            const long long frame = tile_frame(j);
            const int c0 = tile_cell0(j);
            const int nc = min(TC, p.cells - c0);
            const long long cell = frame * p.cells + c0;
            const float* gl = p.logits + cell * kLogitsPerCell;
            const uint8_t* gm = p.mask + cell * kLogitsPerCell;
            const int64_t* ga = (MODE == kSample) ? nullptr : p.action_in + cell * kCompsPerCell;
            const float* gu = (MODE == kSample) ? p.uniforms + cell * kCompsPerCell : nullptr;
            const uint32_t bl = uint32_t(nc) * kLogitsPerCell * 4, bm = uint32_t(nc) * kLogitsPerCell;
            const uint32_t ba = uint32_t(nc) * kCompsPerCell * 8, bu = uint32_t(nc) * kCompsPerCell * 4;
            bool bulk = ((reinterpret_cast<uintptr_t>(gl) | reinterpret_cast<uintptr_t>(gm) | bl | bm) & 15) == 0;
            if (MODE == kSample)
                bulk = bulk && ((reinterpret_cast<uintptr_t>(gu) | bu) & 15) == 0;
            else
                bulk = bulk && ((reinterpret_cast<uintptr_t>(ga) | ba) & 15) == 0;
            if (bulk) {
                if (lane == 0) {
                    const uint32_t total = bl + bm + (MODE == kSample ? bu : ba);
                    mbar_arrive_expect_tx(&full[s], total);
                    bulk_g2s(st, gl, bl, &full[s]);
                    bulk_g2s(st + Cfg::kOffMask, gm, bm, &full[s]);
                    if (MODE == kSample)
                        bulk_g2s(st + Cfg::kOffUni, gu, bu, &full[s]);
                    else
                        bulk_g2s(st + Cfg::kOffAct, ga, ba, &full[s]);
                }
            } else {
                // ragged / unaligned tile: plain cooperative copy by the producer warp
                float* sl = reinterpret_cast<float*>(st);
                for (int i = lane; i < nc * kLogitsPerCell; i += 32) sl[i] = gl[i];
                uint8_t* sm = st + Cfg::kOffMask;
                for (int i = lane; i < nc * kLogitsPerCell; i += 32) sm[i] = gm[i];
                if (MODE == kSample) {
                    float* su = reinterpret_cast<float*>(st + Cfg::kOffUni);
                    for (int i = lane; i < nc * kCompsPerCell; i += 32) su[i] = gu[i];
                } else {
                    int64_t* sa = reinterpret_cast<int64_t*>(st + Cfg::kOffAct);
                    for (int i = lane; i < nc * kCompsPerCell; i += 32) sa[i] = ga[i];
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&full[s]);
            }
        };

        // waits for the consumers of tile j, folds their partial sums (fixed order) and, on the
        // last tile of a frame, writes the per-frame outputs
        auto retire_tile = [&](long long j) {
            const int s = int(j % NSTAGE);
            mbar_wait(&done[s], uint32_t((j / NSTAGE) & 1), p.status);
            if (MODE != kEvalBwd && lane == 0) {
                const float* ps = part + s * Cfg::kConsWarps * 2;
#pragma unroll
                for (int w = 0; w < Cfg::kConsWarps; ++w) {
                    acc_lp += ps[2 * w];
                    acc_en += ps[2 * w + 1];
                }
                if ((j % tpf) == tpf - 1) {
                    const long long frame = tile_frame(j);
                    p.logp_out[frame] = acc_lp;
                    if (p.ent_out) p.ent_out[frame] = acc_en;
                    acc_lp = 0.f;
                    acc_en = 0.f;
                }
            }
        };

        auto store_tile = [&](long long j) {
            const int s = int(j % NSTAGE);
            uint8_t* st = stages + size_t(s) * Cfg::kStageBytes;
            const long long frame = tile_frame(j);
            const int c0 = tile_cell0(j);
            const int nc = min(TC, p.cells - c0);
            const long long cell = frame * p.cells + c0;
            if (MODE == kEvalBwd) {
                float* gd = p.dlogits + cell * kLogitsPerCell;
                const uint32_t bytes = uint32_t(nc) * kLogitsPerCell * 4;
                last_store_bulk = ((reinterpret_cast<uintptr_t>(gd) | bytes) & 15) == 0;
                if (last_store_bulk) {
                    if (lane == 0) {
                        bulk_s2g(gd, st, bytes);
                        bulk_commit();
                    }
                } else {
                    const float* sl = reinterpret_cast<const float*>(st);
                    for (int i = lane; i < nc * kLogitsPerCell; i += 32) gd[i] = sl[i];
                    __syncwarp();
                }
            } else {
                int64_t* ga = p.action_out + cell * kCompsPerCell;
                const uint32_t bytes = uint32_t(nc) * kCompsPerCell * 8;
                last_store_bulk = ((reinterpret_cast<uintptr_t>(ga) | bytes) & 15) == 0;
                if (last_store_bulk) {
                    if (lane == 0) {
                        bulk_s2g(ga, st + Cfg::kOffAct, bytes);
                        bulk_commit();
                    }
                } else {
                    const int64_t* sa = reinterpret_cast<const int64_t*>(st + Cfg::kOffAct);
                    for (int i = lane; i < nc * kCompsPerCell; i += 32) ga[i] = sa[i];
                    __syncwarp();
                }
            }
        };

        if (MODE == kEvalFwd) {
            for (long long j = 0; j < n_tiles + NSTAGE; ++j) {
                if (j >= NSTAGE) retire_tile(j - NSTAGE);
                if (j < n_tiles) load_tile(j);
            }
        } else {
            for (long long j = 0; j < n_tiles && j < NSTAGE; ++j) load_tile(j);
            for (long long j = 0; j < n_tiles; ++j) {
                retire_tile(j);
                store_tile(j);
                if (j >= 1 && j - 1 + NSTAGE < n_tiles) {
                    // the stage of tile j-1 may be refilled once its store has read shared memory
                    if (lane == 0) {
                        if (last_store_bulk)
                            bulk_wait_read<1>();
                        else
                            bulk_wait_read<0>();
                    }
                    __syncwarp();
                    load_tile(j - 1 + NSTAGE);
                }
            }
            if (lane == 0) bulk_wait_all<0>();
        }
    } else {
        // ===================== consumer warps =====================
        const int cw = warp - 1;             // 0 .. kConsWarps-1
        const int group = cw / 3;
        const int role = cw % 3;             // 0: A, 1: B cells 0..15, 2: B cells 16..31
        const int cell_in_tile = (role == 0) ? group * 32 + lane : group * 32 + (role - 1) * 16 + (lane >> 1);
        const int half = lane & 1;

        for (long long j = 0; j < n_tiles; ++j) {
            const int s = int(j % NSTAGE);
            uint8_t* st = stages + size_t(s) * Cfg::kStageBytes;
            const int nc = min(TC, p.cells - tile_cell0(j));
            float g_lp = 0.f, g_h = 0.f;
            if (MODE == kEvalBwd) {
                const long long frame = tile_frame(j);
                g_lp = __ldg(p.g_logp + frame);
                g_h = __ldg(p.g_ent + frame);
            }
            mbar_wait(&full[s], uint32_t((j / NSTAGE) & 1), p.status);

            const bool active = cell_in_tile < nc;
            float* zc = reinterpret_cast<float*>(st) + cell_in_tile * kLogitsPerCell;
            const uint8_t* mc = st + Cfg::kOffMask + cell_in_tile * kLogitsPerCell;
            int64_t* ac = reinterpret_cast<int64_t*>(st + Cfg::kOffAct) + cell_in_tile * kCompsPerCell;
            const float* uc = reinterpret_cast<const float*>(st + Cfg::kOffUni) + cell_in_tile * kCompsPerCell;
            float lp = 0.f, en = 0.f;
            if (role == 0) {
                if (active) a_role_cell<MODE>(zc, mc, ac, ac, uc, g_lp, g_h, p.status, lp, en);
            } else {
                b_role_cell<MODE>(active, half, zc, mc, ac, ac, uc, g_lp, g_h, p.status, lp, en);
            }
            if (MODE != kEvalBwd) {
                lp = warp_sum(lp);
                en = warp_sum(en);
                if (lane == 0) {
                    part[(s * Cfg::kConsWarps + cw) * 2] = lp;
                    part[(s * Cfg::kConsWarps + cw) * 2 + 1] = en;
                }
            }
            if (MODE != kEvalFwd) fence_proxy_async_smem();   // in-place outputs -> async-proxy bulk store
            __syncwarp();
            if (lane == 0) mbar_arrive(&done[s]);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Host launchers
// ---------------------------------------------------------------------------------------------
static int g_head_ctas_per_sm = 2;
static int g_head_grid_override = 0;
void head_set_tuning(int ctas_per_sm, int grid_override) {
    if (ctas_per_sm > 0) g_head_ctas_per_sm = ctas_per_sm;
    g_head_grid_override = grid_override;
}

template <int TC, int MODE, int NSTAGE>
static int launch_head(const HeadParams& p, cudaStream_t stream) {
    using Cfg = HeadCfg<TC, MODE>;
    constexpr size_t smem = head_smem_bytes<TC, MODE, NSTAGE>();
    static bool attr_set[64] = {};
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    if (dev < 64 && !attr_set[dev]) {
        MB_CUDA(cudaFuncSetAttribute(head_kernel<TC, MODE, NSTAGE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     int(smem)));
        attr_set[dev] = true;
    }
    long long grid = (long long)sm_count(dev) * g_head_ctas_per_sm;
    if (g_head_grid_override > 0) grid = g_head_grid_override;
    if (grid > p.n_frames) grid = p.n_frames;
    head_kernel<TC, MODE, NSTAGE><<<dim3((unsigned)grid), dim3(Cfg::kThreads), smem, stream>>>(p);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}

static int check_head_common(const void* logits, const void* mask, long long n_frames, int cells) {
    if (n_frames < 0 || cells <= 0) {
        set_error("head: n_frames=%lld cells=%d out of range", n_frames, cells);
        return MB_ERR_INVALID;
    }
    if (n_frames > 0 && (!logits || !mask)) {
        set_error("head: null logits/mask pointer");
        return MB_ERR_INVALID;
    }
    if ((reinterpret_cast<uintptr_t>(logits) & 3) != 0) {
        set_error("head: logits pointer must be 4-byte aligned");
        return MB_ERR_INVALID;
    }
    return MB_OK;
}

int head_eval_fwd(const float* logits, const uint8_t* mask, const int64_t* action, float* logp_out, float* ent_out,
                  long long n_frames, int cells, int* status, cudaStream_t stream) {
    if (int rc = check_head_common(logits, mask, n_frames, cells)) return rc;
    if (n_frames == 0) return MB_OK;
    if (!action || !logp_out || !ent_out) {
        set_error("head_eval_fwd: null action/logp_out/ent_out pointer");
        return MB_ERR_INVALID;
    }
    HeadParams p{};
    p.logits = logits; p.mask = mask; p.action_in = action; p.logp_out = logp_out; p.ent_out = ent_out;
    p.status = status; p.n_frames = n_frames; p.cells = cells; p.tiles_per_frame = (cells + 63) / 64;
    return launch_head<64, kEvalFwd, 3>(p, stream);
}

int head_eval_bwd(const float* logits, const uint8_t* mask, const int64_t* action, const float* g_logp,
                  const float* g_ent, float* dlogits, long long n_frames, int cells, int* status,
                  cudaStream_t stream) {
    if (int rc = check_head_common(logits, mask, n_frames, cells)) return rc;
    if (n_frames == 0) return MB_OK;
    if (!action || !g_logp || !g_ent || !dlogits) {
        set_error("head_eval_bwd: null action/g_logp/g_ent/dlogits pointer");
        return MB_ERR_INVALID;
    }
    HeadParams p{};
    p.logits = logits; p.mask = mask; p.action_in = action; p.g_logp = g_logp; p.g_ent = g_ent;
    p.dlogits = dlogits; p.status = status; p.n_frames = n_frames; p.cells = cells;
    p.tiles_per_frame = (cells + 63) / 64;
    return launch_head<64, kEvalBwd, 3>(p, stream);
}

int head_sample(const float* logits, const uint8_t* mask, const float* uniforms, int64_t* action_out,
                float* logp_out, float* ent_out, long long n_frames, int cells, int* status, cudaStream_t stream) {
    if (int rc = check_head_common(logits, mask, n_frames, cells)) return rc;
    if (n_frames == 0) return MB_OK;
    if (!uniforms || !action_out || !logp_out) {
        set_error("head_sample: null uniforms/action_out/logp_out pointer");
        return MB_ERR_INVALID;
    }
    HeadParams p{};
    p.logits = logits; p.mask = mask; p.uniforms = uniforms; p.action_out = action_out; p.logp_out = logp_out;
    p.ent_out = ent_out; p.status = status; p.n_frames = n_frames; p.cells = cells;
    p.tiles_per_frame = (cells + 63) / 64;
    return launch_head<64, kSample, 3>(p, stream);
}

}  // namespace mb
