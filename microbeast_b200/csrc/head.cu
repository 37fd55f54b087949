// GridNet masked multi-discrete policy head for sm_100a.
//
// Replaces the 7*C-object Python loop of reference model.py:165-216 (Agent.get_action) and the
// per-component math of model.py:33-52 (CategoricalMasked: where(mask, z, -1e8) -> Categorical
// normalisation -> log_prob / masked entropy / sample), plus its autograd.
//
// Data layout.  logits [N, 78*C] f32, mask [N, 78*C] u8, action [N, 7*C] i64 and uniforms
// [N, 7*C] f32 are all contiguous, i.e. arrays of *cells*: 312 B + 78 B + 56 B (+ 28 B) per cell.
// A CTA streams tiles of TC (32/64/128) consecutive cells of one frame through an NSTAGE-deep
// shared-memory ring: one elected producer thread issues 1-D bulk copies (cp.async.bulk -> SASS
// UBLKCP) that complete on an mbarrier; three consumer warps per 32 cells read the tile out of
// shared memory with thread-per-cell (29 small-component logits) and lane-pair-per-cell (the
// 49-way component) mappings, so every global byte is fetched exactly once, fully coalesced, with
// no register staging.  The backward writes dlogits in place into the stage and ships it with a
// bulk store; sampled actions go straight to global memory.  Per-frame sums are reduced in a fixed
// order (lane shuffle tree -> per-warp partial -> producer thread), so results are run-to-run
// deterministic and need no atomics and no scratch memory.
//
// Roofline: HBM-bound.  Algorithmic bytes per cell: eval fwd 446, eval bwd 758, sample 474
// (+8 / +8 / +4..8 per frame) -- SURVEY.md section 8(d).
#include <atomic>

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
    float* cell_out;           // per-cell mode (eval fwd): [4][n_frames*cells] = logp of the six small
                               // components, their entropy, logp of the 49-way component, its entropy
    int per_cell;              // 1: outputs / upstream gradients are per CELL, not per frame
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
    static constexpr int kActBytes = (MODE == kSample) ? 0 : TC * kCompsPerCell * 8;   // sample writes actions to global
    static constexpr int kUniBytes = (MODE == kSample) ? TC * kCompsPerCell * 4 : 0;
    static constexpr int kOffMask = kLogitBytes;
    static constexpr int kOffAct = kOffMask + kMaskBytes;
    static constexpr int kOffUni = kOffAct + kActBytes;
    static constexpr int kStageBytes = (kOffUni + kUniBytes + 127) / 128 * 128;
    static_assert(kOffMask % 16 == 0 && kOffAct % 16 == 0 && kOffUni % 16 == 0, "bulk copies need 16-B alignment");
    static_assert(TC % 32 == 0 && TC % 8 == 0, "tile granularity");
};

constexpr int kUcdfEntries = 64;     // exact uniform-CDF tables for n = 4 (3 edges), 6 (5), 7 (6), 49 (48)

template <int TC, int MODE, int NSTAGE>
constexpr size_t head_smem_bytes() {
    using Cfg = HeadCfg<TC, MODE>;
    return size_t(NSTAGE) * Cfg::kStageBytes + size_t(NSTAGE) * Cfg::kConsWarps * 2 * sizeof(float) +
           2 * NSTAGE * sizeof(uint64_t) + kUcdfEntries * sizeof(double) + 128;
}

// ---------------------------------------------------------------------------------------------
// Exact (fp64) sampler decisions, as the oracle defines them (oracle/head.py: sample_invcdf):
//   p = CategoricalMasked(logits.double(), mask).probs;  cdf sequential in fp64;
//   a = #{k < n-1 : double(u) >= cdf_k};  landing on a masked entry -> last valid index.
// Two places need them: (1) fully masked components (the common case, p == 1/n exactly) use a
// per-CTA table of the sequential fp64 sums of fl(1/n); (2) components whose fp32 running CDF
// puts u*S inside the guard band of an edge are recomputed by the WHOLE warp (the lanes would
// idle through a divergent single-thread fallback anyway).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ constexpr int ucdf_offset(int n) { return n == 4 ? 0 : (n == 6 ? 3 : (n == 7 ? 8 : 14)); }

// The table holds, per edge, the smallest fp32 value that is >= the fp64 edge: for an fp32 uniform u,
// u >= edge_fp64  <=>  u >= that fp32 value, so the exact decision needs no fp64 at run time.
__device__ __forceinline__ void ucdf_fill(float* tab) {
    const int ns[4] = {4, 6, 7, 49};
    for (int j = 0; j < 4; ++j) {
        const int n = ns[j];
        const double p = 1.0 / double(n);          // softmax of n equal logits: exp(0) / n
        double acc = 0.0;
        for (int k = 0; k < n - 1; ++k) {
            acc += p;                               // sequential, like numpy.cumsum
            tab[ucdf_offset(n) + k] = __double2float_ru(acc);
        }
    }
}

// Uniform (fully masked, quirk Q1) component: exact count from the table around floor(u*n).
__device__ __forceinline__ int ucdf_pick(const float* tab, int n, float u) {
    const float* t = tab + ucdf_offset(n);
    const float ud = u;
    // floor(fl32(u*n)) is within one of the exact count (table entries are (k+1)/n up to 1e-16 and the
    // fp32 product is off by at most n*2^-24), so one look at the two neighbouring edges settles it
    const int a = min(max(int(u * float(n)), 0), n - 1);
    const int up = (a < n - 1 && ud >= t[min(a, n - 2)]) ? 1 : 0;
    const int down = (a > 0 && ud < t[max(a - 1, 0)]) ? 1 : 0;
    return a + up - down;
}

// Warp-cooperative exact recompute of one component (all 32 lanes call it with the same args).
// Single normalisation p_i = exp(z'_i - m) / sum: equal to the oracle's doubly normalised softmax
// up to ~1e-16 relative, and bit-identical in the structured cases that can put a CDF edge exactly
// on an fp32 uniform (equal logits / one valid entry / fully masked: exp(0) == 1, p == 1/k).
__device__ __noinline__ int warp_sample_fp64(const float* z, const uint8_t* mk, int len, float u) {
    const int lane = threadIdx.x & 31;
    const bool in0 = lane < len, in1 = lane + 32 < len;
    const bool v0 = in0 && mk[lane] != 0, v1 = in1 && mk[lane + 32] != 0;
    const double ninf = -__longlong_as_double(0x7ff0000000000000ll);
    const double z0 = in0 ? (v0 ? double(z[lane]) : -1e8) : ninf;
    const double z1 = in1 ? (v1 ? double(z[lane + 32]) : -1e8) : ninf;
    double m = fmax(z0, z1);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    const double e0 = in0 ? exp(z0 - m) : 0.0, e1 = in1 ? exp(z1 - m) : 0.0;
    const double s = warp_sum(e0 + e1);
    const double p0 = e0 / s, p1 = e1 / s;
    const double ud = double(u);
    double cdf = 0.0;
    int a = 0;
    for (int i = 0; i < len - 1; ++i) {
        const double pi = __shfl_sync(0xffffffffu, (i < 32) ? p0 : p1, i & 31);
        cdf += pi;
        a += (ud >= cdf) ? 1 : 0;
    }
    const unsigned vm0 = __ballot_sync(0xffffffffu, v0), vm1 = __ballot_sync(0xffffffffu, v1);
    const bool landed = (a < 32) ? ((vm0 >> a) & 1u) : ((vm1 >> (a - 32)) & 1u);
    if ((vm0 | vm1) && !landed) a = vm1 ? 63 - __clz(int(vm1)) : 31 - __clz(int(vm0));
    return a;
}

// ---- per-component math --------------------------------------------------------------------
// z' = mask ? z : -1e8 (model.py:43); m = max z'; e_i = exp(z'_i - m); S = sum e; W = sum e*(z'-m);
// lse = m + log S (torch logsumexp); logp = z'_a - lse; H = log S - W/S (== -sum p*l over valid
// entries, model.py:46-52).  All code below is branch-free straight-line code over register arrays
// so that the independent chains of the seven components interleave (ILP); log / divide use the
// SFU approximations (lg2.approx: abs err 2^-22; rcp.approx: 1 ulp), far inside the 1e-5 budget.

__device__ __forceinline__ float fast_log(float x) { return lg2_approx(x) * kLn2; }

struct SegScalars {
    float lse;   // logsumexp of the masked logits
    float ent;   // masked entropy
    bool dead;   // every entry masked (quirk Q1): log_prob 0, entropy 0, gradient 0, uniform sample
};

__device__ __forceinline__ SegScalars seg_finish(float m, float S, float W) {
    SegScalars r;
    const float logS = fast_log(S);
    r.lse = m + logS;
    r.dead = (m < -5e7f);
    // fully masked: lse = fl(-1e8 + ln n) = -1e8 in fp32, so lse - m == 0 and W == 0  ->  0 (Q1)
    const float h0 = r.dead ? (r.lse - m) : logS;
    r.ent = h0 - __fdividef(W, S);
    return r;
}

__device__ __forceinline__ float sample_guard(float S, float W, int len) {
    // Rigorous fp32 error bound of (u*S - c_k) against exact arithmetic on the same fp32 inputs,
    // times 1.25: ex2.approx 2^-22 rel. + exponent-argument rounding 3*2^-24*|d| per term, n adds
    // in any order n*2^-24*S, the same again for S inside t = fl(u*S), and 2^-24*S for the product:
    //   S*(2^-21 + (2n+1)*2^-24) + 6*2^-24*|W|.
    return 5.9604644775390625e-08f * (S * (12.f + 2.5f * float(len)) + 8.f * fabsf(W));
}

// ---------------------------------------------------------------------------------------------
// A role: one thread per cell, the six small components (logit offsets 0..28).
// ---------------------------------------------------------------------------------------------
template <int MODE>
__device__ __forceinline__ void a_role_cell(float* zc, const uint8_t* mc, int64_t* ac, const float* uni,
                                            float g_lp, float g_h, int* status, const float* ucdf, bool active,
                                            float& lp, float& en) {
    constexpr int OFF[6] = {0, 6, 10, 14, 18, 22};
    constexpr int LEN[6] = {6, 4, 4, 4, 4, 7};
    float z[29], e[29];
#pragma unroll
    for (int j = 0; j < 14; ++j) {
        const float2 v = *reinterpret_cast<const float2*>(zc + 2 * j);
        const uint32_t h = *reinterpret_cast<const uint16_t*>(mc + 2 * j);
        z[2 * j] = (h & 0xffu) ? v.x : kMaskFill;
        z[2 * j + 1] = (h > 0xffu) ? v.y : kMaskFill;
    }
    z[28] = mc[28] ? zc[28] : kMaskFill;

    float m[6], S[6], W[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        m[k] = z[OFF[k]];
#pragma unroll
        for (int i = 1; i < LEN[k]; ++i) m[k] = fmaxf(m[k], z[OFF[k] + i]);
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        S[k] = 0.f;
        W[k] = 0.f;
#pragma unroll
        for (int i = 0; i < LEN[k]; ++i) {
            const float d = z[OFF[k] + i] - m[k];
            const float ei = ex2_approx(d * kLog2e);
            e[OFF[k] + i] = ei;
            S[k] += ei;
            W[k] = fmaf(ei, d, W[k]);
        }
    }
    SegScalars sc[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) sc[k] = seg_finish(m[k], S[k], W[k]);

    int a[6];
    if (MODE == kSample) {
        unsigned near_bits = 0;
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            const float u = uni[k];
            const float t = u * S[k];
            const float guard = sample_guard(S[k], W[k], LEN[k]);
            float c = 0.f;
            int cnt = 0;
            bool near = false;
#pragma unroll
            for (int i = 0; i < LEN[k] - 1; ++i) {
                c += e[OFF[k] + i];
                cnt += (t >= c) ? 1 : 0;
                near = near || (fabsf(t - c) <= guard);
            }
            // uniform over all n entries when fully masked: exact table decision (branch-free select)
            const int cnt_dead = ucdf_pick(ucdf, LEN[k], u);
            cnt = sc[k].dead ? cnt_dead : cnt;
            near = near && !sc[k].dead;
            a[k] = cnt;
            near_bits |= (near && active) ? (1u << k) : 0u;
        }
        // rare: decide exactly as the fp64 definition does, the whole warp on one component at a time
        unsigned flagged = __ballot_sync(0xffffffffu, near_bits != 0);
        while (flagged) {
            const int src = __ffs(flagged) - 1;
            flagged &= flagged - 1;
            const unsigned nb = __shfl_sync(0xffffffffu, near_bits, src);
            const int delta = (src - int(threadIdx.x & 31));          // cells are consecutive in the tile
#pragma unroll
            for (int k = 0; k < 6; ++k) {
                if (nb & (1u << k)) {
                    const int r = warp_sample_fp64(zc + delta * kLogitsPerCell + OFF[k],
                                                   mc + delta * kLogitsPerCell + OFF[k], LEN[k],
                                                   uni[delta * kCompsPerCell + k]);
                    if (delta == 0) a[k] = r;
                }
            }
        }
        if (active) {
#pragma unroll
            for (int k = 0; k < 6; ++k) ac[k] = a[k];
        }
    } else {
        bool bad = false;
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            const long long v = ac[k];
            bad = bad || (v < 0) || (v >= LEN[k]);
            a[k] = int(v < 0 ? 0 : (v >= LEN[k] ? LEN[k] - 1 : v));
        }
        if (bad && active && status) atomicOr(status, kStatusBadAction);
    }
    bool taken_valid[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        // z'_a re-read from shared memory (dynamic index); sentinel when the taken entry is masked (Q2)
        taken_valid[k] = mc[OFF[k] + a[k]] != 0;
        const float za = taken_valid[k] ? zc[OFF[k] + a[k]] : kMaskFill;
        lp += active ? (za - sc[k].lse) : 0.f;
        en += active ? sc[k].ent : 0.f;
    }
    if (MODE == kEvalBwd && active) {
        // d/dz_i [g_lp*logp + g_h*H] = p_i*(-g_lp - g_h*(l_i + H)) + g_lp*[i==a]; masked entries have
        // p_i == 0; a fully masked component gets exactly 0 (verified vs autograd)
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            const float invS = sc[k].dead ? 0.f : __fdividef(1.f, S[k]);
            const float c0 = -g_lp - g_h * sc[k].ent;
#pragma unroll
            for (int i = 0; i < LEN[k]; ++i) {
                const float l = z[OFF[k] + i] - sc[k].lse;
                z[OFF[k] + i] = (e[OFF[k] + i] * invS) * fmaf(-g_h, l, c0);
            }
        }
        // z[] now holds dlogits; write it over the logits in the stage, then add the one-hot term
#pragma unroll
        for (int j = 0; j < 14; ++j) *reinterpret_cast<float2*>(zc + 2 * j) = make_float2(z[2 * j], z[2 * j + 1]);
        zc[28] = z[28];
#pragma unroll
        for (int k = 0; k < 6; ++k)
            if (taken_valid[k]) zc[OFF[k] + a[k]] += g_lp;
    }
}

// ---------------------------------------------------------------------------------------------
// B role: a lane pair per cell, the 49-way component (logit offsets 29..77).
// even lane: element 29 + pairs 30..53 (25 entries); odd lane: pairs 54..77 (24 entries).
// The 24 paired entries are summed as four independent blocks of six (ILP); the sampling CDF is
// base_of_block + running sum inside the block.
// ---------------------------------------------------------------------------------------------
template <int MODE>
__device__ __forceinline__ void b_role_cell(bool active, int half, float* zc, const uint8_t* mc, int64_t* ac,
                                            const float* uni, float g_lp, float g_h, int* status,
                                            const float* ucdf, float& lp, float& en) {
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
    float mb[4];
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        mb[b] = z[6 * b];
#pragma unroll
        for (int i = 1; i < 6; ++i) mb[b] = fmaxf(mb[b], z[6 * b + i]);
    }
    float m = fmaxf(fmaxf(mb[0], mb[1]), fmaxf(mb[2], mb[3]));
    if (has_x) m = fmaxf(m, zx);
    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));

    float Sb[4], Wb[4];
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        Sb[b] = 0.f;
        Wb[b] = 0.f;
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            const float d = z[6 * b + i] - m;
            const float ei = ex2_approx(d * kLog2e);
            e[6 * b + i] = ei;
            Sb[b] += ei;
            Wb[b] = fmaf(ei, d, Wb[b]);
        }
    }
    float Wx = 0.f;
    if (has_x) {
        const float d = zx - m;
        ex = ex2_approx(d * kLog2e);
        Wx = ex * d;
    }
    // bases of the four blocks in this lane's part of the CDF (the odd lane's start is filled below)
    const float own1 = ex + Sb[0], own2 = own1 + Sb[1], own3 = own2 + Sb[2];
    const float S_own = own3 + Sb[3];
    const float W_own = ((Wx + Wb[0]) + Wb[1]) + (Wb[2] + Wb[3]);
    const float S_other = __shfl_xor_sync(0xffffffffu, S_own, 1);
    const float W_other = __shfl_xor_sync(0xffffffffu, W_own, 1);
    const float S = has_x ? (S_own + S_other) : (S_other + S_own);     // same operand order on both lanes
    const float W = has_x ? (W_own + W_other) : (W_other + W_own);
    const SegScalars sc = seg_finish(m, S, W);

    int a = 0;
    if (MODE == kSample) {
        const float u = uni[6];
        const float t = u * S;
        const float guard = sample_guard(S, W, LEN);
        const float start = has_x ? 0.f : S_other;     // even lane's 25 entries come first
        // Two-level count.  The running CDF is monotone, so every edge of a block that ends at or
        // below t is counted and none of a block that starts above t; only the block that straddles
        // t is scanned entry by entry (its six entries are picked out of the register array with
        // selects), and an edge can be within `guard` of t only if it is in that block or is one of
        // its two boundaries.
        const float base[5] = {start + ex, start + own1, start + own2, start + own3, start + S_own};
        const int bsel = ((t >= base[1]) ? 1 : 0) + ((t >= base[2]) ? 1 : 0) + ((t >= base[3]) ? 1 : 0);
        float c = (bsel & 2) ? ((bsel & 1) ? base[3] : base[2]) : ((bsel & 1) ? base[1] : base[0]);
        int cnt = 6 * bsel + ((has_x && t >= ex) ? 1 : 0);
        bool near = fabsf(t - c) <= guard;
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            const float lo = (bsel & 1) ? e[6 + i] : e[i];
            const float hi = (bsel & 1) ? e[18 + i] : e[12 + i];
            c += (bsel & 2) ? hi : lo;
            // entry 48 (k = n-1: the odd lane's last one) is not an edge of the count
            const bool edge = has_x || i < 5 || bsel < 3;
            cnt += (edge && t >= c) ? 1 : 0;
            near = near || (edge && fabsf(t - c) <= guard);
        }
        cnt += __shfl_xor_sync(0xffffffffu, cnt, 1);
        near = (__shfl_xor_sync(0xffffffffu, int(near), 1) != 0) || near;
        a = cnt;
        // uniform over all 49 entries when fully masked: exact table decision (branch-free select)
        const int a_dead = ucdf_pick(ucdf, LEN, u);
        a = sc.dead ? a_dead : a;
        near = near && !sc.dead;
        unsigned flagged = __ballot_sync(0xffffffffu, near && has_x && active);
        while (flagged) {
            const int src = __ffs(flagged) - 1;
            flagged &= flagged - 1;
            const int delta = (src >> 1) - int((threadIdx.x & 31) >> 1);
            const int r = warp_sample_fp64(zc + delta * kLogitsPerCell + OFF, mc + delta * kLogitsPerCell + OFF,
                                           LEN, uni[delta * kCompsPerCell + 6]);
            if (delta == 0) a = r;
        }
        if (has_x && active) ac[6] = a;
    } else if (active) {
        const long long v = ac[6];
        const bool bad = (v < 0) || (v >= LEN);
        a = int(v < 0 ? 0 : (v >= LEN ? LEN - 1 : v));
        if (bad && has_x && status) atomicOr(status, kStatusBadAction);
    }
    const bool taken_valid = active && (mc[OFF + a] != 0);
    if (active && has_x) {
        const float za = taken_valid ? zc[OFF + a] : kMaskFill;
        lp += za - sc.lse;
        en += sc.ent;
    }
    if (MODE == kEvalBwd) {
        __syncwarp();       // the even lane has read z'_a (possibly in the odd lane's range) above
        if (active) {
            const float invS = sc.dead ? 0.f : __fdividef(1.f, S);
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
    float* ucdf = reinterpret_cast<float*>(done + NSTAGE);        // sample mode only

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < NSTAGE; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&done[s], Cfg::kConsWarps);
        }
        fence_mbar_init();
        if (MODE == kSample) ucdf_fill(ucdf);
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
                if ((j % tpf) == tpf - 1 && !p.per_cell) {
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
            }
        };

        if (MODE != kEvalBwd) {
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
            const bool active = cell_in_tile < nc;
            const long long cell_global = tile_frame(j) * p.cells + tile_cell0(j) + cell_in_tile;
            float g_lp = 0.f, g_h = 0.f;
            if (MODE == kEvalBwd) {
                // upstream gradients: one pair per frame, or (cell-list mode) one pair per cell
                const long long gi = p.per_cell ? (active ? cell_global : 0) : tile_frame(j);
                g_lp = __ldg(p.g_logp + gi);
                g_h = __ldg(p.g_ent + gi);
            }
            mbar_wait(&full[s], uint32_t((j / NSTAGE) & 1), p.status);

            float* zc = reinterpret_cast<float*>(st) + cell_in_tile * kLogitsPerCell;
            const uint8_t* mc = st + Cfg::kOffMask + cell_in_tile * kLogitsPerCell;
            // eval: the taken actions staged in shared memory; sample: the global output row of this cell
            int64_t* ac = (MODE == kSample)
                              ? p.action_out + (tile_frame(j) * p.cells + tile_cell0(j) + cell_in_tile) * kCompsPerCell
                              : reinterpret_cast<int64_t*>(st + Cfg::kOffAct) + cell_in_tile * kCompsPerCell;
            const float* uc = reinterpret_cast<const float*>(st + Cfg::kOffUni) + cell_in_tile * kCompsPerCell;
            float lp = 0.f, en = 0.f;
            // lanes past the end of a ragged tile run on stale stage contents and discard the result
            if (role == 0) {
                a_role_cell<MODE>(zc, mc, ac, uc, g_lp, g_h, p.status, ucdf, active, lp, en);
            } else {
                b_role_cell<MODE>(active, half, zc, mc, ac, uc, g_lp, g_h, p.status, ucdf, lp, en);
            }
            if (MODE == kEvalFwd && p.per_cell) {
                // cell-list mode: no per-frame sum; each cell's two partial results go out as they are
                const long long total = p.n_frames * p.cells;
                if (active && (role == 0 || half == 0)) {
                    float* o = p.cell_out + (role == 0 ? 0 : 2) * total + cell_global;
                    o[0] = lp;
                    o[total] = en;
                }
            } else if (MODE != kEvalBwd) {
                lp = warp_sum(lp);
                en = warp_sum(en);
                if (lane == 0) {
                    part[(s * Cfg::kConsWarps + cw) * 2] = lp;
                    part[(s * Cfg::kConsWarps + cw) * 2 + 1] = en;
                }
            }
            if (MODE == kEvalBwd) fence_proxy_async_smem();   // in-place dlogits -> async-proxy bulk store
            __syncwarp();
            if (lane == 0) mbar_arrive(&done[s]);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Host launchers
// ---------------------------------------------------------------------------------------------
static std::atomic<int> g_head_ctas_per_sm{0};      // 0 = as many as fit (occupancy query)
static std::atomic<int> g_head_grid_override{0};
static std::atomic<int> g_head_variant{-1};         // tile / pipeline shape, see dispatch_head; -1 = per-mode default
void head_set_tuning(int ctas_per_sm, int grid_override, int variant) {
    g_head_ctas_per_sm = ctas_per_sm;
    g_head_grid_override = grid_override;
    g_head_variant = variant;
}

template <int TC, int MODE, int NSTAGE>
static int launch_head(HeadParams p, cudaStream_t stream) {
    using Cfg = HeadCfg<TC, MODE>;
    constexpr size_t smem = head_smem_bytes<TC, MODE, NSTAGE>();
    static int occ[64] = {};
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    if (dev >= 64) dev = 63;
    if (occ[dev] == 0) {
        MB_CUDA(cudaFuncSetAttribute(head_kernel<TC, MODE, NSTAGE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     int(smem)));
        int n = 0;
        MB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, head_kernel<TC, MODE, NSTAGE>, Cfg::kThreads,
                                                              smem));
        occ[dev] = n > 0 ? n : 1;
    }
    p.tiles_per_frame = (p.cells + TC - 1) / TC;
    const int want_per_sm = g_head_ctas_per_sm.load(), want_grid = g_head_grid_override.load();
    const int per_sm = (want_per_sm > 0 && want_per_sm < occ[dev]) ? want_per_sm : occ[dev];
    long long grid = (long long)sm_count(dev) * per_sm;
    if (want_grid > 0) grid = want_grid;
    if (grid > p.n_frames) grid = p.n_frames;
    head_kernel<TC, MODE, NSTAGE><<<dim3((unsigned)grid), dim3(Cfg::kThreads), smem, stream>>>(p);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}

// Tile / pipeline shapes selectable with mb_set_tuning("head_variant", v) for benchmarking.
template <int MODE>
static int dispatch_head(const HeadParams& p, cudaStream_t stream) {
    int v = g_head_variant;
    if (v < 0) v = (MODE == kEvalFwd) ? 1 : (MODE == kEvalBwd ? 0 : 7);   // measured best on B200 (profiles/)
    switch (v) {
        case 1: return launch_head<64, MODE, 2>(p, stream);
        case 2: return launch_head<64, MODE, 4>(p, stream);
        case 3: return launch_head<32, MODE, 3>(p, stream);
        case 4: return launch_head<32, MODE, 5>(p, stream);
        case 5: return launch_head<128, MODE, 3>(p, stream);
        case 6: return launch_head<128, MODE, 2>(p, stream);
        case 7: return launch_head<32, MODE, 2>(p, stream);
        default: return launch_head<64, MODE, 3>(p, stream);
    }
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
    p.status = status; p.n_frames = n_frames; p.cells = cells; 
    return dispatch_head<kEvalFwd>(p, stream);
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
    return dispatch_head<kEvalBwd>(p, stream);
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
    return dispatch_head<kSample>(p, stream);
}

// Cell-list mode: the inputs are an arbitrary list of cells ([n_cells, 78] logits / mask, [n_cells, 7]
// actions) instead of whole frames -- what a sparse actor head produces when it computes logits only
// for cells that have a valid action.  Results are per cell; the caller scatters them to frames.
int head_eval_cells_fwd(const float* logits, const uint8_t* mask, const int64_t* action, float* cell_out4,
                        long long n_cells, int* status, cudaStream_t stream) {
    if (n_cells < 0 || n_cells % 64 != 0) {
        set_error("head_eval_cells_fwd: n_cells=%lld must be a non-negative multiple of 64 (pad with masked cells)",
                  n_cells);
        return MB_ERR_INVALID;
    }
    if (n_cells == 0) return MB_OK;
    if (!logits || !mask || !action || !cell_out4) {
        set_error("head_eval_cells_fwd: null pointer");
        return MB_ERR_INVALID;
    }
    HeadParams p{};
    p.logits = logits; p.mask = mask; p.action_in = action; p.cell_out = cell_out4; p.per_cell = 1;
    p.status = status; p.n_frames = n_cells / 64; p.cells = 64;
    return dispatch_head<kEvalFwd>(p, stream);
}

int head_eval_cells_bwd(const float* logits, const uint8_t* mask, const int64_t* action, const float* g_logp_cell,
                        const float* g_ent_cell, float* dlogits, long long n_cells, int* status,
                        cudaStream_t stream) {
    if (n_cells < 0 || n_cells % 64 != 0) {
        set_error("head_eval_cells_bwd: n_cells=%lld must be a non-negative multiple of 64 (pad with masked cells)",
                  n_cells);
        return MB_ERR_INVALID;
    }
    if (n_cells == 0) return MB_OK;
    if (!logits || !mask || !action || !g_logp_cell || !g_ent_cell || !dlogits) {
        set_error("head_eval_cells_bwd: null pointer");
        return MB_ERR_INVALID;
    }
    HeadParams p{};
    p.logits = logits; p.mask = mask; p.action_in = action; p.g_logp = g_logp_cell; p.g_ent = g_ent_cell;
    p.dlogits = dlogits; p.per_cell = 1; p.status = status; p.n_frames = n_cells / 64; p.cells = 64;
    return dispatch_head<kEvalBwd>(p, stream);
}

}  // namespace mb
