// V-trace targets / advantages fused with the pg / baseline / entropy loss reduction and their
// gradients, for sm_100a.
//
// Replaces the inline block of reference libs/utils.py:277-329 (PPO_learn): importance weights
// rho = c = min(1, exp(lp_new - lp_old)), the reverse-time Python loop of utils.py:302-304, vs,
// pg_advantages, and the three .mean() reductions + total (utils.py:320-329), plus the autograd
// of those means (d/dlp_new = adv/n, d/dV = -(vs - V)/n, d/dentropy = -coef/n).
//
// Layout (reference D2): every input is [R, T] with TIME contiguous.  The recurrence is serial in
// time and independent per row, so one thread owns one row.  A thread walking its row backwards
// in time straight from global memory would touch 32 different 128-B lines per warp request;
// instead a CTA (compute warp + TMA warp) stages [32 rows x 32|16 steps] tiles of the inputs through
// TMA (cp.async.bulk.tensor.2d, SASS UTMALDG; 128-B / 64-B swizzle so that per-thread float4 reads
// are bank conflict free) into a multi-stage shared-memory ring, walks the time tiles from the end,
// overwrites the tiles in place with vs / adv / d_logp / d_values and writes them back with TMA
// stores (UTMASTG).  Loss sums are carried per thread (fp32 inside 32 steps, fp64 across),
// reduced by a fixed shuffle tree, written per CTA to a workspace and folded by the last CTA to
// arrive (ticket counter) in a fixed order: deterministic, one launch, no float atomics.
// Three kernels share the arithmetic (vt_step): the TMA kernel, a generic thread-per-row kernel
// for layouts TMA cannot describe, and a time-major kernel for [T, R] arrays.
//
// Roofline: HBM-bound, 29 B per [r,t] element forward (+8 B with the two gradient outputs).
// Shapes below ~1e7 elements are launch/latency bound and L2 resident.
#include <atomic>

#include "common.cuh"

#include <cuda.h>
#include <math.h>
#include <string.h>

#include <mutex>
#include <type_traits>

namespace mb {

struct VtParams {
    const float* lp_old;
    const float* lp_new;
    const float* values;
    const float* rewards;
    const float* entropy;
    const uint8_t* done;
    const float* bootstrap;
    float* vs;
    float* adv;
    float* dlp;      // may be null
    float* dv;       // may be null
    float* loss4;    // [pg, value, entropy, total]
    double* partials;        // [grid][3]
    unsigned int* ticket;    // zero on entry, reset to zero by the last CTA
    long long rows;
    int T;
    long long in_stride;     // elements between consecutive rows of the inputs
    long long out_stride;    // elements between consecutive rows of the outputs
    float gamma, rho_bar, c_bar, pg_rho_bar, ent_coef, inv_n;
    int debug;               // diagnostics (mb_set_tuning "vtrace_debug"): 1 = skip the arithmetic, 2 = no entropy, 4 = no stores
};

struct VtMaps {
    CUtensorMap lp_old, lp_new, values, rewards, done, vs, adv, dlp, dv;
};

struct VtCarry {
    float acc;       // vs_{t+1} - V_{t+1}   (utils.py:300: starts at 0)
    float v_next;    // V_{t+1}              (utils.py:296: bootstrap at the end)
    float vs_next;   // vs_{t+1}             (utils.py:313-314)
};

struct VtStepOut {
    float vs, adv;
};

// One time step of utils.py:277-318 for one row.  Every product and sum is rounded separately
// (__fmul_rn / __fadd_rn: no FMA contraction), in the reference's left-to-right order, so the
// only arithmetic difference to the CPU torch path is exp (ex2.approx of the scaled log-ratio:
// 2 ulp).  Branch-free: `ok` (step inside [0,T) and row inside [0,R)) only gates the carry, so the
// four steps of a chunk form one basic block and their independent parts overlap.
__device__ __forceinline__ float vt_ratio_exp(float lp_old, float lp_new) {
    return ex2_approx(__fmul_rn(__fsub_rn(lp_new, lp_old), kLog2e));           // utils.py:285,290
}

__device__ __forceinline__ VtStepOut vt_step(const VtParams& p, float ratio_exp, float v, float rew, bool done,
                                             bool ok, VtCarry& c) {
    const float disc = done ? 0.0f : p.gamma;                                   // utils.py:277
    const float rho = fminf(ratio_exp, p.rho_bar);                              // utils.py:293
    const float cs = fminf(ratio_exp, p.c_bar);                                 // utils.py:294
    const float delta =
        __fmul_rn(rho, __fsub_rn(__fadd_rn(rew, __fmul_rn(disc, c.v_next)), v));        // utils.py:298
    const float acc = __fadd_rn(delta, __fmul_rn(__fmul_rn(disc, cs), c.acc));  // utils.py:303
    VtStepOut o;
    o.vs = __fadd_rn(acc, v);                                                   // utils.py:310
    const float pg_rho = fminf(rho, p.pg_rho_bar);                              // utils.py:316
    o.adv = __fmul_rn(pg_rho, __fsub_rn(__fadd_rn(rew, __fmul_rn(disc, c.vs_next)), v));  // utils.py:318
    c.acc = ok ? acc : c.acc;
    c.v_next = ok ? v : c.v_next;
    c.vs_next = ok ? o.vs : c.vs_next;
    return o;
}

// Folds the per-CTA partial sums; executed by one warp of the last CTA to arrive.
__device__ __forceinline__ void vt_finalize(const VtParams& p, int lane) {
    double s[3] = {0.0, 0.0, 0.0};
    for (unsigned int b = lane; b < gridDim.x; b += 32) {
#pragma unroll
        for (int i = 0; i < 3; ++i) s[i] += __ldcg(p.partials + size_t(b) * 3 + i);
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) s[i] = warp_sum(s[i]);
    if (lane == 0) {
        const double n_inv = double(p.inv_n);
        const float pg = float(s[0] * n_inv);                    // utils.py:320
        const float vl = float(0.5 * s[1] * n_inv);              // utils.py:323
        const float en = float(s[2] * n_inv);                    // utils.py:326
        p.loss4[0] = pg;
        p.loss4[1] = vl;
        p.loss4[2] = en;
        p.loss4[3] = __fsub_rn(__fadd_rn(pg, vl), __fmul_rn(p.ent_coef, en));   // utils.py:329
        *p.ticket = 0u;
    }
}

// ---------------------------------------------------------------------------------------------
// TMA kernel: one warp per CTA, 32 rows, time tiles of TS (32 or 16) steps walked from the end.
// The recurrence makes every thread a long dependent chain, so throughput comes from many resident
// CTAs per SM: a stage is 17 B x 32 rows x TS steps (17 KB / 8.5 KB), NSTAGE of them per CTA.
// Entropy only feeds a mean, so it is not staged: the warp streams its rows' entropy with plain
// coalesced float4 loads issued at the top of every iteration and summed at its end.
// ---------------------------------------------------------------------------------------------
constexpr int kVtRows = 32;

template <int TS, int ROWS = kVtRows>
struct VtTile {
    static_assert(TS == 32 || TS == 16, "tile steps");
    static_assert(ROWS == 32 || ROWS == 16, "rows per group");
    static constexpr int kF32 = ROWS * TS * 4;             // 32 rows: 4096 / 2048 B; rows of 128 / 64 B, swizzled
    static constexpr int kU8 = ROWS * TS;                  // 32 rows: 1024 / 512 B; rows of 32 / 16 B, no swizzle
    static constexpr int kStage = 4 * kF32 + 1024;         // lp_old, lp_new, values, rewards, done (padded)
    static constexpr int kChunks = TS / 4;                 // float4 chunks per row
    static constexpr int kRowPitch = TS * 4;
    // byte offset of float4 chunk j of row `lane` under SWIZZLE_128B (TS=32) / SWIZZLE_64B (TS=16)
    __device__ static __forceinline__ uint32_t off(int lane, int j) {
        return TS == 32 ? uint32_t(lane) * 128u + (uint32_t(j ^ (lane & 7)) << 4)
                        : uint32_t(lane) * 64u + (uint32_t(j ^ ((lane >> 1) & 3)) << 4);
    }
};

// DONE_TMA = false: the rows' byte pitch is not a multiple of 16 (e.g. T = 20), which a tensor map
// cannot describe for the u8 `done` array; each thread then reads its row's done bytes directly.
// ROWS = 16: half the lanes of the compute warp idle, but a launch with few rows splits into twice as many groups
// (16 384 rows are 512 groups of 32: 3.46 per SM, so a quarter of the SMs carry 4 while the rest carry 3).
template <int NSTAGE, int TS, bool DONE_TMA = true, int ROWS = kVtRows>
__global__ void __launch_bounds__(64) vtrace_tma_kernel(const __grid_constant__ VtMaps maps, const VtParams p) {
    using Tile = VtTile<TS, ROWS>;
    constexpr int kStageBytes = Tile::kStage;              // multiple of 1024
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full[NSTAGE];     // TMA loads of a stage have landed
    __shared__ __align__(8) uint64_t done[NSTAGE];     // the compute warp has overwritten the stage with outputs
    __shared__ double ent_sum;
    // swizzled tiles need 1024-B alignment
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);

    // warp 0 computes (one thread per row); warp 1 is the TMA warp: its lane 0 issues every load and
    // store, so descriptor issue, store drain and refill never sit on the compute warp's critical path
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const bool want_grads = (p.dlp != nullptr);
    // persistent: this CTA owns the 32-row groups blockIdx.x, blockIdx.x + gridDim.x, ... and runs
    // their time tiles as ONE stream through the ring, so the pipeline never drains between groups
    const int n_groups = int((p.rows + ROWS - 1) / ROWS);
    const int my_groups = (n_groups - int(blockIdx.x) + int(gridDim.x) - 1) / int(gridDim.x);

    if (threadIdx.x == 0) {
        for (int s = 0; s < NSTAGE; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&done[s], 1);
        }
        fence_mbar_init();
        tma_prefetch_desc(&maps.lp_old);
        tma_prefetch_desc(&maps.lp_new);
        tma_prefetch_desc(&maps.values);
        tma_prefetch_desc(&maps.rewards);
        if (DONE_TMA) tma_prefetch_desc(&maps.done);
    }
    __syncthreads();

    const int nt = (p.T + TS - 1) / TS;

    const long long total_q = (long long)my_groups * nt;
    // lane 0 of the TMA warp: loads the tile with consumption index x = (group x / nt, tile x % nt counted from
    // the END of the rows) into stage x % NSTAGE
    auto issue_load = [&](long long x) {
        const int s = int(x % NSTAGE);
        const int xg = int(x / nt), xk = int(x % nt);
        const int t0 = (nt - 1 - xk) * TS;
        const int row0 = (int(blockIdx.x) + xg * int(gridDim.x)) * ROWS;
        uint8_t* st = smem + size_t(s) * kStageBytes;
        mbar_arrive_expect_tx(&full[s], uint32_t(4 * Tile::kF32 + (DONE_TMA ? Tile::kU8 : 0)));
        tma_load_2d(st + 0 * Tile::kF32, &maps.lp_old, t0, row0, &full[s]);
        tma_load_2d(st + 1 * Tile::kF32, &maps.lp_new, t0, row0, &full[s]);
        tma_load_2d(st + 2 * Tile::kF32, &maps.values, t0, row0, &full[s]);
        tma_load_2d(st + 3 * Tile::kF32, &maps.rewards, t0, row0, &full[s]);
        if (DONE_TMA) tma_load_2d(st + 4 * Tile::kF32, &maps.done, t0, row0, &full[s]);
    };
    // Tiles are CONSUMED from the end of a row towards its start, but they are REQUESTED in blocks of kBlk
    // consecutive tiles (256 B of a row), lowest address first: with the tensor maps' 256-B L2 promotion a miss on
    // the lower half of a 256-B block brings in the upper half as well, which only an ascending request order can
    // use (measured, scripts/micro/tma_multi.cu: 5.96 TB/s descending, 6.68 TB/s in ascending pairs, 6.72 TB/s
    // ascending throughout).  Block b of a group holds its tiles [b*kBlk, b*kBlk + kBlk) in consumption order.
    constexpr int kBlk = (256 / (TS * 4)) < (NSTAGE / 2) ? (256 / (TS * 4)) : (NSTAGE / 2 > 0 ? NSTAGE / 2 : 1);
    long long lq = 0;                     // first consumption index that has not been requested yet
    auto block_len = [&](long long q0) { return int(min((long long)kBlk, (long long)(nt - int(q0 % nt)))); };
    auto issue_block = [&]() {            // requests tiles [lq, lq + n) in DEscending consumption = AScending address order
        const int n = block_len(lq);
        for (int i = n - 1; i >= 0; --i) issue_load(lq + i);
        lq += n;
    };

    if (warp == 1) {
        // ===================== TMA warp =====================
        // lane 0 issues every tensor copy; ALL lanes stream the group's entropy, which only feeds a mean: the group's
        // [rows_here x T] block is a list of float4 slots (nq per row), lane l owns slots l, l+32, ... and fetches the
        // tile's share (kEntSlots per lane) at the top of an iteration; the values are summed at its end, after the
        // wait for the compute warp and the TMA work, so their latency is off every critical path (on the compute
        // warp they stalled the recurrence once per tile: 15 % of its stall samples in profiles/r01_ncu_stalls.md)
        constexpr int kEntSlots = ROWS * TS / 128;
        const int nq = (p.T + 3) >> 2;
        const int tail = p.T - 4 * (nq - 1);               // valid floats in a row's last slot (1..4)
        int ent_r = 0, ent_c = 0, e_row0 = 0, e_rows = 0;
        float s_en = 0.0f;
        double s_en_d = 0.0;
        if (lane == 0)
            while (lq < total_q && lq + block_len(lq) <= NSTAGE) issue_block();
        int sg = 0, sk = 0;                                // store cursor: tile sk of group sg
        float4 eprev[kEntSlots];                           // the previous iteration's loads: summed one iteration late
#pragma unroll
        for (int i = 0; i < kEntSlots; ++i) eprev[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        for (long long q = 0; q < total_q; ++q) {
            const int s = int(q % NSTAGE);
            if (sk == 0) {
                e_row0 = (int(blockIdx.x) + sg * int(gridDim.x)) * ROWS;
                e_rows = int(min((long long)ROWS, p.rows - e_row0));
                ent_r = 0;
                ent_c = lane;
                while (ent_c >= nq) { ent_c -= nq; ++ent_r; }
                s_en_d += double(s_en);
                s_en = 0.0f;
            }
            float4 ebuf[kEntSlots];
#pragma unroll
            for (int i = 0; i < kEntSlots; ++i) {
                ebuf[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (ent_r < e_rows && !(p.debug & 2)) {
                    const float* src = p.entropy + (long long)(e_row0 + ent_r) * p.in_stride + 4 * ent_c;
                    float4 v;
                    if (ent_c == nq - 1 && tail < 4) {     // never read past T in a row's last slot
                        v = make_float4(__ldg(src), tail > 1 ? __ldg(src + 1) : 0.f, tail > 2 ? __ldg(src + 2) : 0.f, 0.f);
                    } else {
                        // volatile asm: stays ahead of the mbarrier wait below
                        asm volatile("ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%4];"
                                     : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                                     : "l"(src));
                    }
                    ebuf[i] = v;
                }
                ent_c += 32;
                while (ent_c >= nq) { ent_c -= nq; ++ent_r; }
            }
            if (lane == 0) {
                mbar_wait(&done[s], uint32_t((q / NSTAGE) & 1), nullptr);
                uint8_t* st = smem + size_t(s) * kStageBytes;
                const int t0 = (nt - 1 - sk) * TS;
                // in place: vs over rewards, adv over lp_old, d_logp over lp_new, d_values over values
                if (!(p.debug & 4)) tma_store_2d(&maps.vs, t0, e_row0, st + 3 * Tile::kF32);
                if (!(p.debug & 4)) tma_store_2d(&maps.adv, t0, e_row0, st + 0 * Tile::kF32);
                if (want_grads && !(p.debug & 4)) {
                    tma_store_2d(&maps.dlp, t0, e_row0, st + 1 * Tile::kF32);
                    tma_store_2d(&maps.dv, t0, e_row0, st + 2 * Tile::kF32);
                }
                bulk_commit();
                // a block may be requested once the stages it lands in are free: its last tile x = lq + n - 1 reuses
                // the stage of tile x - NSTAGE, whose store must have finished reading shared memory (<= q - 1)
                if (lq < total_q && lq + block_len(lq) - 1 - NSTAGE <= q - 1) {
                    bulk_wait_read<1>();
                    issue_block();
                }
            }
            __syncwarp();
#pragma unroll
            for (int i = 0; i < kEntSlots; ++i) {
                s_en += (eprev[i].x + eprev[i].y) + (eprev[i].z + eprev[i].w);
                eprev[i] = ebuf[i];
            }
            if (++sk == nt) { sk = 0; ++sg; }
        }
#pragma unroll
        for (int i = 0; i < kEntSlots; ++i) s_en += (eprev[i].x + eprev[i].y) + (eprev[i].z + eprev[i].w);
        if (lane == 0) bulk_wait_all<0>();     // output tiles are in global memory before the CTA retires
        const double tot = warp_sum(s_en_d + double(s_en));
        if (lane == 0) ent_sum = tot;
        __syncthreads();
        return;
    }
    // ===================== compute warp =====================

    VtCarry carry;
    carry.acc = 0.0f;
    carry.v_next = 0.0f;
    carry.vs_next = 0.0f;
    double s_pg = 0.0, s_v = 0.0;
    float t_pg = 0.0f, t_v = 0.0f;
    // bootstrap of the NEXT group is fetched one group ahead so its latency is never exposed
    float boot_next = 0.0f;
    {
        const long long r = (long long)blockIdx.x * ROWS + lane;
        if (lane < ROWS && r < p.rows) boot_next = __ldg(p.bootstrap + r);
    }
    int row0 = 0, rows_here = 0;
    bool row_ok = false;
    const int lrow = lane & (ROWS - 1);                    // idle lanes (ROWS = 16) read a live row and never write

    int k = 0, g = 0;                                      // consumer-side cursor: tile k of group g
    for (long long q = 0; q < total_q; ++q) {
        const int s = int(q % NSTAGE);
        if (k == 0) {                                      // a new group of 32 rows starts
            row0 = (int(blockIdx.x) + g * int(gridDim.x)) * ROWS;
            rows_here = int(min((long long)ROWS, p.rows - row0));
            row_ok = lane < rows_here;
            carry.acc = 0.0f;                              // utils.py:300
            carry.v_next = boot_next;                      // utils.py:296: V_T = bootstrap
            carry.vs_next = boot_next;                     // utils.py:313-314
            const long long rn = (long long)(row0 + int(gridDim.x) * ROWS) + lane;
            boot_next = (g + 1 < my_groups && lane < ROWS && rn < p.rows) ? __ldg(p.bootstrap + rn) : 0.0f;
        }
        mbar_wait(&full[s], uint32_t((q / NSTAGE) & 1), nullptr);

        uint8_t* st = smem + size_t(s) * kStageBytes;
        const int t0 = (nt - 1 - k) * TS;
        uint8_t* q_old = st + 0 * Tile::kF32;
        uint8_t* q_new = st + 1 * Tile::kF32;
        uint8_t* q_val = st + 2 * Tile::kF32;
        uint8_t* q_rew = st + 3 * Tile::kF32;
        const uint8_t* q_dn = st + 4 * Tile::kF32;
        // Only the first tile of a group (the END of its rows) can reach past T, and rows past R exist
        // only in the last group: every other tile runs the body without any range checks.
        auto tile_body = [&](auto guard_tag) {
            constexpr bool kGuard = decltype(guard_tag)::value;
            // software pipeline over the chunks of the tile: the four LDS.128 (and the done word) of
            // chunk j-1 are issued before chunk j is computed, so shared-memory latency and the next
            // chunk's exp() overlap the serial part of the recurrence
            auto load_dn = [&](int j) -> uint32_t {
                if (DONE_TMA) return *reinterpret_cast<const uint32_t*>(q_dn + lrow * TS + j * 4);
                uint32_t d = 0;
                const uint8_t* drow = p.done + (long long)(row0 + lane) * p.in_stride + t0 + 4 * j;
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    if (row_ok && t0 + 4 * j + i < p.T) d |= uint32_t(__ldg(drow + i) != 0) << (8 * i);
                return d;
            };
            float4 n_old, n_new, n_val, n_rew;
            uint32_t n_dn;
            {
                const uint32_t off0 = Tile::off(lrow, Tile::kChunks - 1);
                n_old = *reinterpret_cast<const float4*>(q_old + off0);
                n_new = *reinterpret_cast<const float4*>(q_new + off0);
                n_val = *reinterpret_cast<const float4*>(q_val + off0);
                n_rew = *reinterpret_cast<const float4*>(q_rew + off0);
                n_dn = load_dn(Tile::kChunks - 1);
            }
#pragma unroll
            for (int j = Tile::kChunks - 1; j >= 0; --j) {
                const uint32_t off = Tile::off(lrow, j);
                const float4 a_old = n_old, a_new = n_new, a_val = n_val, a_rew = n_rew;
                const uint32_t dn = n_dn;
                if (j > 0) {
                    const uint32_t offn = Tile::off(lrow, j - 1);
                    n_old = *reinterpret_cast<const float4*>(q_old + offn);
                    n_new = *reinterpret_cast<const float4*>(q_new + offn);
                    n_val = *reinterpret_cast<const float4*>(q_val + offn);
                    n_rew = *reinterpret_cast<const float4*>(q_rew + offn);
                    n_dn = load_dn(j - 1);
                }
                const float lo[4] = {a_old.x, a_old.y, a_old.z, a_old.w};
                const float ln[4] = {a_new.x, a_new.y, a_new.z, a_new.w};
                const float vv[4] = {a_val.x, a_val.y, a_val.z, a_val.w};
                const float rw[4] = {a_rew.x, a_rew.y, a_rew.z, a_rew.w};
                float o_vs[4], o_adv[4], o_dlp[4], o_dv[4], re[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) re[i] = vt_ratio_exp(lo[i], ln[i]);
#pragma unroll
                for (int i = 3; i >= 0; --i) {
                    const bool ok = !kGuard || (row_ok && (t0 + 4 * j + i < p.T));
                    const bool dflag = ((dn >> (8 * i)) & 0xffu) != 0;
                    const VtStepOut o = vt_step(p, re[i], vv[i], rw[i], dflag, ok, carry);
                    const float dvs = __fsub_rn(o.vs, vv[i]);
                    o_vs[i] = o.vs;             // out-of-range rows / steps are clipped by the TMA store
                    o_adv[i] = o.adv;
                    o_dlp[i] = o.adv * p.inv_n;
                    o_dv[i] = -dvs * p.inv_n;
                    t_pg = __fadd_rn(t_pg, ok ? __fmul_rn(ln[i], o.adv) : 0.0f);
                    t_v = __fadd_rn(t_v, ok ? __fmul_rn(dvs, dvs) : 0.0f);
                }
                // in place: vs over rewards, adv over lp_old, d_logp over lp_new, d_values over values
                if (ROWS == 32 || lane < ROWS) {
                    *reinterpret_cast<float4*>(q_rew + off) = make_float4(o_vs[0], o_vs[1], o_vs[2], o_vs[3]);
                    *reinterpret_cast<float4*>(q_old + off) = make_float4(o_adv[0], o_adv[1], o_adv[2], o_adv[3]);
                }
                if (want_grads && (ROWS == 32 || lane < ROWS)) {
                    *reinterpret_cast<float4*>(q_new + off) = make_float4(o_dlp[0], o_dlp[1], o_dlp[2], o_dlp[3]);
                    *reinterpret_cast<float4*>(q_val + off) = make_float4(o_dv[0], o_dv[1], o_dv[2], o_dv[3]);
                }
            }
        };
        if (p.debug & 1) {
        } else if ((k == 0) || rows_here < 32)
            tile_body(std::true_type{});
        else
            tile_body(std::false_type{});
        // the pg / value sums are folded into fp64 once per 32 steps (same grouping as the generic kernel)
        if (TS == 32 || ((nt - 1 - k) & 1) == 0) {
            s_pg += double(t_pg);
            s_v += double(t_v);
            t_pg = 0.0f;
            t_v = 0.0f;
        }
        fence_proxy_async_smem();       // generic-proxy writes -> visible to the TMA store
        __syncwarp();
        if (lane == 0) mbar_arrive(&done[s]);
        if (++k == nt) { k = 0; ++g; }
    }

    s_pg = warp_sum(s_pg);
    s_v = warp_sum(s_v);
    __syncthreads();                    // the TMA warp's entropy sum is in shared memory
    unsigned int tk = 0;
    if (lane == 0) {
        p.partials[size_t(blockIdx.x) * 3 + 0] = s_pg;
        p.partials[size_t(blockIdx.x) * 3 + 1] = s_v;
        p.partials[size_t(blockIdx.x) * 3 + 2] = ent_sum;
        __threadfence();
        tk = atomicAdd(p.ticket, 1u);
    }
    tk = __shfl_sync(0xffffffffu, tk, 0);
    if (tk == gridDim.x - 1) {
        __threadfence();
        vt_finalize(p, lane);
    }
}

// ---------------------------------------------------------------------------------------------
// Generic kernel: any stride / alignment (e.g. T = 20, sliced [R, T+1] views).  One thread per
// row reading global memory directly; correct everywhere, fast nowhere.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) vtrace_generic_kernel(const VtParams p) {
    __shared__ double red[4][3];
    __shared__ unsigned int tk_s;
    const long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double s_pg = 0.0, s_v = 0.0, s_en = 0.0;
    if (row < p.rows) {
        const long long ib = row * p.in_stride, ob = row * p.out_stride;
        VtCarry carry;
        carry.acc = 0.0f;
        carry.v_next = p.bootstrap[row];
        carry.vs_next = carry.v_next;
        float t_pg = 0.f, t_v = 0.f, t_en = 0.f;
        for (int t = p.T - 1; t >= 0; --t) {
            const float v = p.values[ib + t], ln = p.lp_new[ib + t];
            const VtStepOut o = vt_step(p, vt_ratio_exp(p.lp_old[ib + t], ln), v, p.rewards[ib + t],
                                        p.done[ib + t] != 0, true, carry);
            p.vs[ob + t] = o.vs;
            p.adv[ob + t] = o.adv;
            const float dvs = __fsub_rn(o.vs, v);
            if (p.dlp) p.dlp[ob + t] = o.adv * p.inv_n;
            if (p.dv) p.dv[ob + t] = -dvs * p.inv_n;
            t_pg = __fadd_rn(t_pg, __fmul_rn(ln, o.adv));
            t_v = __fadd_rn(t_v, __fmul_rn(dvs, dvs));
            t_en = __fadd_rn(t_en, p.entropy[ib + t]);
            if ((t & 31) == 0) {
                s_pg += double(t_pg); s_v += double(t_v); s_en += double(t_en);
                t_pg = t_v = t_en = 0.f;
            }
        }
    }
    s_pg = warp_sum(s_pg);
    s_v = warp_sum(s_v);
    s_en = warp_sum(s_en);
    if (lane == 0) {
        red[warp][0] = s_pg; red[warp][1] = s_v; red[warp][2] = s_en;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 0; i < 3; ++i)
            p.partials[size_t(blockIdx.x) * 3 + i] = red[0][i] + red[1][i] + red[2][i] + red[3][i];
        __threadfence();
        tk_s = atomicAdd(p.ticket, 1u);
    }
    __syncthreads();
    if (tk_s == gridDim.x - 1 && warp == 0) {
        __threadfence();
        vt_finalize(p, lane);
    }
}

// ---------------------------------------------------------------------------------------------
// Time-major kernel: arrays laid out [T, R] (rows contiguous), which is the order the rollout
// buffers are written in (libs/utils.py:35-45: [T+1, n_envs, ...]) and the order get_batch
// flattens them (utils.py:197-205).  With one thread per row a warp's 32 loads of one time step
// are then a single 128-B line: perfectly coalesced, no shared-memory staging needed.  Each thread
// keeps the next block of kTmU time steps in flight in registers while it computes the current one.
// Same arithmetic (vt_step) and the same 32-step fp64 folding as the other two kernels.
// ---------------------------------------------------------------------------------------------
constexpr int kTmU = 8;

struct TmBlock {
    float lo[kTmU], ln[kTmU], v[kTmU], rw[kTmU], en[kTmU];
    uint32_t dn[kTmU];
};

__device__ __forceinline__ void tm_load(const VtParams& p, long long row, int t_hi, bool row_ok, TmBlock& b) {
#pragma unroll
    for (int i = 0; i < kTmU; ++i) {
        const int t = t_hi - 1 - i;
        const bool ok = row_ok && t >= 0;
        const long long idx = ok ? (long long)t * p.in_stride + row : 0;
        b.lo[i] = ok ? __ldg(p.lp_old + idx) : 0.f;
        b.ln[i] = ok ? __ldg(p.lp_new + idx) : 0.f;
        b.v[i] = ok ? __ldg(p.values + idx) : 0.f;
        b.rw[i] = ok ? __ldg(p.rewards + idx) : 0.f;
        b.en[i] = ok ? __ldg(p.entropy + idx) : 0.f;
        b.dn[i] = ok ? uint32_t(__ldg(p.done + idx)) : 0u;
    }
}

__global__ void __launch_bounds__(128) vtrace_tmajor_kernel(const VtParams p) {
    __shared__ double red[4][3];
    __shared__ unsigned int tk_s;
    const long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool row_ok = row < p.rows;
    double s_pg = 0.0, s_v = 0.0, s_en = 0.0;
    VtCarry carry;
    carry.acc = 0.0f;
    carry.v_next = row_ok ? __ldg(p.bootstrap + row) : 0.f;
    carry.vs_next = carry.v_next;
    float t_pg = 0.f, t_v = 0.f, t_en = 0.f;
    TmBlock cur, nxt;
    tm_load(p, row, p.T, row_ok, cur);
    for (int t_hi = p.T; t_hi > 0; t_hi -= kTmU) {
        if (t_hi - kTmU > 0) tm_load(p, row, t_hi - kTmU, row_ok, nxt);
#pragma unroll
        for (int i = 0; i < kTmU; ++i) {
            const int t = t_hi - 1 - i;
            const bool ok = row_ok && t >= 0;
            const VtStepOut o = vt_step(p, vt_ratio_exp(cur.lo[i], cur.ln[i]), cur.v[i], cur.rw[i], cur.dn[i] != 0, ok,
                                        carry);
            const float dvs = __fsub_rn(o.vs, cur.v[i]);
            if (ok) {
                const long long od = (long long)t * p.out_stride + row;
                p.vs[od] = o.vs;
                p.adv[od] = o.adv;
                if (p.dlp) {
                    p.dlp[od] = o.adv * p.inv_n;
                    p.dv[od] = -dvs * p.inv_n;
                }
            }
            t_pg = __fadd_rn(t_pg, ok ? __fmul_rn(cur.ln[i], o.adv) : 0.f);
            t_v = __fadd_rn(t_v, ok ? __fmul_rn(dvs, dvs) : 0.f);
            t_en = __fadd_rn(t_en, ok ? cur.en[i] : 0.f);
            if (ok && (t & 31) == 0) {
                s_pg += double(t_pg); s_v += double(t_v); s_en += double(t_en);
                t_pg = t_v = t_en = 0.f;
            }
        }
        cur = nxt;
    }
    s_pg = warp_sum(s_pg);
    s_v = warp_sum(s_v);
    s_en = warp_sum(s_en);
    if (lane == 0) {
        red[warp][0] = s_pg; red[warp][1] = s_v; red[warp][2] = s_en;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 0; i < 3; ++i)
            p.partials[size_t(blockIdx.x) * 3 + i] = red[0][i] + red[1][i] + red[2][i] + red[3][i];
        __threadfence();
        tk_s = atomicAdd(p.ticket, 1u);
    }
    __syncthreads();
    if (tk_s == gridDim.x - 1 && warp == 0) {
        __threadfence();
        vt_finalize(p, lane);
    }
}

// ---------------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(ptr);
    }
    return fn;
}

static std::atomic<int> g_vt_l2_promo{256};      // bytes: 0 (none), 64, 128, 256

static bool encode_map_2d(EncodeTiledFn enc, CUtensorMap* m, const void* base, bool is_u8, long long rows, int T,
                          long long stride_elems, int tile_steps, int box_rows) {
    const cuuint64_t gdim[2] = {cuuint64_t(T), cuuint64_t(rows)};
    const cuuint64_t gstride[1] = {cuuint64_t(stride_elems) * (is_u8 ? 1u : 4u)};
    const cuuint32_t box[2] = {cuuint32_t(tile_steps), cuuint32_t(box_rows)};
    const cuuint32_t estr[2] = {1, 1};
    // 256-B L2 promotion: a row's next (earlier) time tile sits in the other half of the same
    // 256-B block, so one DRAM page activation serves two consecutive tiles of the walk
    const CUtensorMapL2promotion promo = g_vt_l2_promo >= 256  ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B
                                         : g_vt_l2_promo >= 128 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B
                                         : g_vt_l2_promo >= 64  ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B
                                                                : CU_TENSOR_MAP_L2_PROMOTION_NONE;
    CUresult r = enc(m, is_u8 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                     const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     is_u8 ? CU_TENSOR_MAP_SWIZZLE_NONE
                           : (tile_steps == 32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B),
                     promo,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}

// Encoding a tensor map costs a driver call; the learner calls with the same buffers every step,
// so descriptors are kept in a small direct-mapped cache keyed by (pointer, shape, stride).
struct MapKey {
    const void* base;
    long long rows, stride;
    int T, is_u8, promo, ts, box_rows;
    bool operator==(const MapKey& o) const {
        return base == o.base && rows == o.rows && stride == o.stride && T == o.T && is_u8 == o.is_u8 &&
               promo == o.promo && ts == o.ts && box_rows == o.box_rows;
    }
};
struct MapSlot {
    MapKey key;
    CUtensorMap map;
    bool valid;
};
static MapSlot g_map_cache[256];
static std::mutex g_map_mutex;

static bool make_map_2d(EncodeTiledFn enc, CUtensorMap* m, const void* base, bool is_u8, long long rows, int T,
                        long long stride_elems, int tile_steps, int box_rows) {
    const MapKey key{base, rows, stride_elems, T, is_u8 ? 1 : 0, g_vt_l2_promo, tile_steps, box_rows};
    size_t h = reinterpret_cast<uintptr_t>(base) >> 8;
    h ^= size_t(rows) * 0x9E3779B97F4A7C15ull ^ size_t(T) * 0xC2B2AE3D27D4EB4Full ^ size_t(stride_elems) << 17 ^
         size_t(tile_steps) << 3 ^ size_t(box_rows) << 9;
    MapSlot& slot = g_map_cache[(h ^ (h >> 29)) & 255];
    std::lock_guard<std::mutex> lock(g_map_mutex);
    if (slot.valid && slot.key == key) {
        *m = slot.map;
        return true;
    }
    if (!encode_map_2d(enc, m, base, is_u8, rows, T, stride_elems, tile_steps, box_rows)) return false;
    slot.key = key;
    slot.map = *m;
    slot.valid = true;
    return true;
}

static std::atomic<int> g_vt_debug{0};
void vtrace_set_debug(int v) { g_vt_debug = v; }
static std::atomic<int> g_vt_force_generic{0};
static std::atomic<int> g_vt_variant{-1};        // -1 = by grid size; see the switch in vtrace_loss_launch
void vtrace_set_tuning(int force_generic, int variant, int l2_promo) {
    g_vt_force_generic = force_generic;
    g_vt_variant = variant;
    g_vt_l2_promo = l2_promo;
}

size_t vtrace_workspace_bytes(long long rows) {
    if (rows < 0) rows = 0;
    const size_t ctas = size_t((rows + kVtRows - 1) / kVtRows);     // the generic kernel uses fewer
    return 256 + (ctas + 1) * 3 * sizeof(double);
}

static inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <int NSTAGE, int TS, bool DONE_TMA = true, int ROWS = kVtRows>
static int launch_vt_tma(const VtMaps& maps, const VtParams& p, cudaStream_t stream) {
    constexpr size_t smem = size_t(NSTAGE) * VtTile<TS, ROWS>::kStage + 1024;
    static bool attr_set[64] = {};
    static int per_sm[64] = {};
    int dev = 0;
    MB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) {
        set_error("vtrace_loss: device index %d not supported (0..63)", dev);
        return MB_ERR_UNSUPPORTED;
    }
    if (!attr_set[dev]) {
        MB_CUDA(cudaFuncSetAttribute(vtrace_tma_kernel<NSTAGE, TS, DONE_TMA, ROWS>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        int n = 0;
        MB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, vtrace_tma_kernel<NSTAGE, TS, DONE_TMA, ROWS>, 64, smem));
        per_sm[dev] = n > 0 ? n : 1;
        attr_set[dev] = true;
    }
    const long long groups = (p.rows + ROWS - 1) / ROWS;
    const long long resident = (long long)sm_count(dev) * per_sm[dev];
    // one partial-sum slot per CTA: the workspace holds ceil(rows / 32) + 1 of them
    const long long slots = (p.rows + kVtRows - 1) / kVtRows + 1;
    long long grid = groups < resident ? groups : resident;
    if (grid > slots) grid = slots;
    vtrace_tma_kernel<NSTAGE, TS, DONE_TMA, ROWS><<<unsigned(grid), 64, smem, stream>>>(maps, p);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}

// Default pipeline shape by problem size (measured on B200, profiles/r02_sweep_vtrace_variants.jsonl).
static int vt_default_variant(long long rows, int T) {
    if (T <= 128) return 4;                                  // short rows: 4 stages x 16 steps, 6 CTAs per SM
    // long rows: 6 stages x 32 steps (2 CTAs per SM) when there are several waves of groups or less than one;
    // in between (about one wave) 3 x 32 keeps 4 CTAs per SM and every group resident at once
    return (rows >= 32768 || rows <= 8192) ? 8 : 1;
}

// pipeline shapes selectable with mb_set_tuning("vtrace_variant", v): stages, steps per tile, rows per group
struct VtShape { int stages, ts, rows; };
static const VtShape kVtShapes[16] = {{2, 32, 32}, {3, 32, 32}, {4, 32, 32}, {3, 16, 32}, {4, 16, 32}, {5, 16, 32},
                                      {6, 16, 32}, {8, 16, 32}, {6, 32, 32}, {12, 16, 32}, {8, 32, 32}, {4, 32, 16},
                                      {6, 32, 16}, {8, 32, 16}, {8, 16, 16}, {6, 16, 16}};

// `scratch` holds the ticket word at offset 0 (must be zero on entry; the kernel leaves it zero)
// and the per-CTA partial sums from offset 256.
// `clear_header`: zero the header on `stream` once the arguments have passed validation (callers whose scratch
// is not known to start at zero).
int vtrace_loss_launch(VtParams p, void* scratch, size_t scratch_bytes, cudaStream_t stream, bool clear_header = false) {
    if (p.rows < 0 || p.T <= 0) {
        set_error("vtrace_loss: rows=%lld T=%d out of range", p.rows, p.T);
        return MB_ERR_INVALID;
    }
    if (!p.loss4) {
        set_error("vtrace_loss: null loss4_out pointer");
        return MB_ERR_INVALID;
    }
    if (p.rows == 0) {
        MB_CUDA(cudaMemsetAsync(p.loss4, 0, 4 * sizeof(float), stream));
        return MB_OK;
    }
    if (!p.lp_old || !p.lp_new || !p.values || !p.rewards || !p.entropy || !p.done || !p.bootstrap || !p.vs ||
        !p.adv) {
        set_error("vtrace_loss: null input/output pointer");
        return MB_ERR_INVALID;
    }
    if ((p.dlp == nullptr) != (p.dv == nullptr)) {
        set_error("vtrace_loss: dlp_out and dV_out must both be given or both be null");
        return MB_ERR_INVALID;
    }
    if (p.in_stride < p.T || p.out_stride < p.T) {
        set_error("vtrace_loss: row stride smaller than T");
        return MB_ERR_INVALID;
    }
    if (!scratch || scratch_bytes < vtrace_workspace_bytes(p.rows) || (reinterpret_cast<uintptr_t>(scratch) & 7)) {
        set_error("vtrace_loss: workspace missing, misaligned or smaller than mb_vtrace_workspace_bytes(rows)=%zu",
                  vtrace_workspace_bytes(p.rows));
        return MB_ERR_INVALID;
    }
    if (clear_header) MB_CUDA(cudaMemsetAsync(scratch, 0, 256, stream));
    p.ticket = reinterpret_cast<unsigned int*>(scratch);
    p.partials = reinterpret_cast<double*>(reinterpret_cast<uint8_t*>(scratch) + 256);
    if (p.inv_n <= 0.0f) p.inv_n = float(1.0 / (double(p.rows) * double(p.T)));
    p.debug = g_vt_debug;

    EncodeTiledFn enc = get_encode_fn();
    bool tma_ok = enc && !g_vt_force_generic && (p.in_stride % 4 == 0) && (p.out_stride % 4 == 0) &&
                  al16(p.lp_old) && al16(p.lp_new) && al16(p.values) && al16(p.rewards) && al16(p.entropy) &&
                  al16(p.vs) && al16(p.adv) && (!p.dlp || (al16(p.dlp) && al16(p.dv)));
    const bool done_tma = (p.in_stride % 16 == 0) && al16(p.done);
    if (tma_ok) {
        int v = g_vt_variant;
        if (v < 0 || v > 15) v = vt_default_variant(p.rows, p.T);
        if (!done_tma) v = (kVtShapes[v].ts == 16) ? 4 : 2;      // the direct-`done` kernel exists in two shapes
        const VtShape sh = kVtShapes[v];
        const int ts = sh.ts, br = sh.rows;
        VtMaps maps;
        memset(&maps, 0, sizeof(maps));
        bool ok = make_map_2d(enc, &maps.lp_old, p.lp_old, false, p.rows, p.T, p.in_stride, ts, br) &&
                  make_map_2d(enc, &maps.lp_new, p.lp_new, false, p.rows, p.T, p.in_stride, ts, br) &&
                  make_map_2d(enc, &maps.values, p.values, false, p.rows, p.T, p.in_stride, ts, br) &&
                  make_map_2d(enc, &maps.rewards, p.rewards, false, p.rows, p.T, p.in_stride, ts, br) &&
                  (!done_tma || make_map_2d(enc, &maps.done, p.done, true, p.rows, p.T, p.in_stride, ts, br)) &&
                  make_map_2d(enc, &maps.vs, p.vs, false, p.rows, p.T, p.out_stride, ts, br) &&
                  make_map_2d(enc, &maps.adv, p.adv, false, p.rows, p.T, p.out_stride, ts, br);
        if (ok && p.dlp)
            ok = make_map_2d(enc, &maps.dlp, p.dlp, false, p.rows, p.T, p.out_stride, ts, br) &&
                 make_map_2d(enc, &maps.dv, p.dv, false, p.rows, p.T, p.out_stride, ts, br);
        if (ok && !done_tma) {
            if (ts == 16) return launch_vt_tma<4, 16, false>(maps, p, stream);
            return launch_vt_tma<4, 32, false>(maps, p, stream);
        }
        if (ok) {
            switch (v) {
                case 0: return launch_vt_tma<2, 32>(maps, p, stream);
                case 1: return launch_vt_tma<3, 32>(maps, p, stream);
                case 2: return launch_vt_tma<4, 32>(maps, p, stream);
                case 3: return launch_vt_tma<3, 16>(maps, p, stream);
                case 4: return launch_vt_tma<4, 16>(maps, p, stream);
                case 5: return launch_vt_tma<5, 16>(maps, p, stream);
                case 6: return launch_vt_tma<6, 16>(maps, p, stream);
                case 7: return launch_vt_tma<8, 16>(maps, p, stream);
                case 8: return launch_vt_tma<6, 32>(maps, p, stream);
                case 9: return launch_vt_tma<12, 16>(maps, p, stream);
                case 10: return launch_vt_tma<8, 32>(maps, p, stream);
                case 11: return launch_vt_tma<4, 32, true, 16>(maps, p, stream);
                case 12: return launch_vt_tma<6, 32, true, 16>(maps, p, stream);
                case 13: return launch_vt_tma<8, 32, true, 16>(maps, p, stream);
                case 14: return launch_vt_tma<8, 16, true, 16>(maps, p, stream);
                default: return launch_vt_tma<6, 16, true, 16>(maps, p, stream);
            }
        }
        // tensor-map encoding refused the shape: fall through to the generic kernel
    }
    const unsigned grid = unsigned((p.rows + 127) / 128);
    vtrace_generic_kernel<<<grid, 128, 0, stream>>>(p);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}

}  // namespace mb

#include "../../include/mbeast.h"

extern "C" int mb_vtrace_loss(const float* lp_old, const float* lp_new, const float* values, const float* bootstrap,
                              const float* rewards, const uint8_t* done, const float* entropy, int64_t rows,
                              int32_t T, int64_t in_row_stride, int64_t out_row_stride, float gamma, float rho_bar,
                              float c_bar, float pg_rho_bar, float ent_coef, float inv_n, float* vs_out,
                              float* adv_out, float* loss4_out, float* dlp_out, float* dV_out, void* workspace,
                              size_t workspace_bytes, void* stream) {
    mb::VtParams p;
    memset(&p, 0, sizeof(p));
    p.lp_old = lp_old; p.lp_new = lp_new; p.values = values; p.rewards = rewards; p.entropy = entropy;
    p.done = done; p.bootstrap = bootstrap; p.vs = vs_out; p.adv = adv_out; p.dlp = dlp_out; p.dv = dV_out;
    p.loss4 = loss4_out; p.rows = rows; p.T = T; p.in_stride = in_row_stride; p.out_stride = out_row_stride;
    p.gamma = gamma; p.rho_bar = rho_bar; p.c_bar = c_bar; p.pg_rho_bar = pg_rho_bar; p.ent_coef = ent_coef;
    p.inv_n = inv_n;
    return mb::vtrace_loss_launch(p, workspace, workspace_bytes, static_cast<cudaStream_t>(stream));
}

// ---------------------------------------------------------------------------------------------
// mb_vtrace_loss_static: the same computation for callers that cannot hand in a workspace.
//
// mb_vtrace_loss needs a few KB of scratch (a ticket word and three partial sums per CTA) and
// leaves it to the caller to allocate them.  Some users of this library are not in a position to
// allocate device memory where they call it: a plain C / ctypes host that links no allocator, a
// step loop that must not run into cudaMalloc's device-wide synchronisation, or a thread that runs
// while another part of the process has a stream under capture (the runtime refuses cudaMalloc
// then).  For them the scratch lives in the library image itself: one fixed array per GPU that has
// loaded the library, held until the process exits.
//
// The array's name is only a host-side handle; the kernel needs the device address behind it, which
// cudaGetSymbolAddress returns for the device that is current at the time.  The first call on each
// device does that look-up (and nothing else extra), so make the first call on a device outside any
// stream capture.  No call allocates or frees device memory.
//
// All calls on one GPU share the array and it keeps whatever the previous call left there.  The only
// part that has to start at zero, the ticket word, is cleared on the caller's stream before every
// launch.  Two calls that may run on the GPU at the same time (different streams, no dependency
// between them) would corrupt each other's sums: serialise them or use mb_vtrace_loss with separate
// workspaces.
// ---------------------------------------------------------------------------------------------
namespace mb {

constexpr long long kVtStaticMaxRows = 65536;            // 2048 CTAs x 24 B of partial sums + header
constexpr size_t kVtStaticScratchBytes = 64 * 1024;

static __device__ __align__(256) unsigned char g_vt_static_scratch[kVtStaticScratchBytes];

static std::atomic<int> g_vt_static_poison{0};    // test hook, see mb_set_tuning("vtrace_static_poison")

static int vt_static_scratch(void** out, cudaStream_t stream) {
    static std::mutex mu;
    static void* ptrs[64] = {};
    int dev = -1;
    MB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) {
        set_error("mb_vtrace_loss_static: device index %d not supported (0..63)", dev);
        return MB_ERR_UNSUPPORTED;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        // The look-up may load the module lazily, which is not allowed while a stream is being captured:
        // refuse with a clear message instead of letting the runtime invalidate the caller's capture.
        cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
        MB_CUDA(cudaStreamIsCapturing(stream, &cap));
        if (cap != cudaStreamCaptureStatusNone) {
            set_error("mb_vtrace_loss_static: the first call on a device looks up the scratch array's address and "
                      "must be made outside stream capture (device %d has not been set up yet)", dev);
            return MB_ERR_INVALID;
        }
        void* p = nullptr;
        MB_CUDA(cudaGetSymbolAddress(&p, g_vt_static_scratch));      // device address on the CURRENT device
        ptrs[dev] = p;
    }
    *out = ptrs[dev];
    return MB_OK;
}

void vtrace_static_poison(int on) { g_vt_static_poison = on; }

}  // namespace mb

extern "C" int mb_vtrace_loss_static(const float* lp_old, const float* lp_new, const float* values,
                                     const float* bootstrap, const float* rewards, const uint8_t* done,
                                     const float* entropy, int64_t rows, int32_t T, int64_t in_row_stride,
                                     int64_t out_row_stride, float gamma, float rho_bar, float c_bar,
                                     float pg_rho_bar, float ent_coef, float inv_n, float* vs_out, float* adv_out,
                                     float* loss4_out, float* dlp_out, float* dV_out, void* stream) {
    if (rows > mb::kVtStaticMaxRows) {
        mb::set_error("mb_vtrace_loss_static: rows=%lld exceeds the built-in scratch (at most %lld rows per call); "
                      "use mb_vtrace_loss with a caller-owned workspace",
                      (long long)rows, mb::kVtStaticMaxRows);
        return MB_ERR_INVALID;
    }
    static_assert(mb::kVtStaticScratchBytes >= 256 + (mb::kVtStaticMaxRows / mb::kVtRows + 1) * 3 * sizeof(double),
                  "static scratch too small for the advertised row limit");
    void* scratch = nullptr;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (int rc = mb::vt_static_scratch(&scratch, st)) return rc;
    if (mb::g_vt_static_poison) {       // test hook: leave the array as dirty as a previous user could have
        mb::g_vt_static_poison = 0;
        MB_CUDA(cudaMemsetAsync(scratch, 0xA5, mb::kVtStaticScratchBytes, st));
    }
    mb::VtParams p;
    memset(&p, 0, sizeof(p));
    p.lp_old = lp_old; p.lp_new = lp_new; p.values = values; p.rewards = rewards; p.entropy = entropy;
    p.done = done; p.bootstrap = bootstrap; p.vs = vs_out; p.adv = adv_out; p.dlp = dlp_out; p.dv = dV_out;
    p.loss4 = loss4_out; p.rows = rows; p.T = T; p.in_stride = in_row_stride; p.out_stride = out_row_stride;
    p.gamma = gamma; p.rho_bar = rho_bar; p.c_bar = c_bar; p.pg_rho_bar = pg_rho_bar; p.ent_coef = ent_coef;
    p.inv_n = inv_n;
    // the header (ticket word and its padding) must start at zero: cleared on `stream`, after the arguments
    // have been validated, right before the launch
    return mb::vtrace_loss_launch(p, scratch, mb::kVtStaticScratchBytes, st, /*clear_header=*/true);
}


// ---------------------------------------------------------------------------------------------
// mb_vtrace_loss_tmajor: the same computation on TIME-MAJOR arrays [T, rows].
// ---------------------------------------------------------------------------------------------
extern "C" int mb_vtrace_loss_tmajor(const float* lp_old, const float* lp_new, const float* values,
                                     const float* bootstrap, const float* rewards, const uint8_t* done,
                                     const float* entropy, int64_t rows, int32_t T, int64_t in_t_stride,
                                     int64_t out_t_stride, float gamma, float rho_bar, float c_bar,
                                     float pg_rho_bar, float ent_coef, float inv_n, float* vs_out, float* adv_out,
                                     float* loss4_out, float* dlp_out, float* dV_out, void* workspace,
                                     size_t workspace_bytes, void* stream) {
    using namespace mb;
    if (rows < 0 || T <= 0) {
        set_error("vtrace_loss_tmajor: rows=%lld T=%d out of range", (long long)rows, T);
        return MB_ERR_INVALID;
    }
    if (!loss4_out) {
        set_error("vtrace_loss_tmajor: null loss4_out pointer");
        return MB_ERR_INVALID;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (rows == 0) {
        MB_CUDA(cudaMemsetAsync(loss4_out, 0, 4 * sizeof(float), st));
        return MB_OK;
    }
    if (!lp_old || !lp_new || !values || !rewards || !entropy || !done || !bootstrap || !vs_out || !adv_out) {
        set_error("vtrace_loss_tmajor: null input/output pointer");
        return MB_ERR_INVALID;
    }
    if ((dlp_out == nullptr) != (dV_out == nullptr)) {
        set_error("vtrace_loss_tmajor: dlp_out and dV_out must both be given or both be null");
        return MB_ERR_INVALID;
    }
    if (in_t_stride < rows || out_t_stride < rows) {
        set_error("vtrace_loss_tmajor: time stride smaller than rows");
        return MB_ERR_INVALID;
    }
    if (!workspace || workspace_bytes < vtrace_workspace_bytes(rows) || (reinterpret_cast<uintptr_t>(workspace) & 7)) {
        set_error("vtrace_loss_tmajor: workspace missing, misaligned or smaller than mb_vtrace_workspace_bytes(rows)=%zu",
                  vtrace_workspace_bytes(rows));
        return MB_ERR_INVALID;
    }
    VtParams p;
    memset(&p, 0, sizeof(p));
    p.lp_old = lp_old; p.lp_new = lp_new; p.values = values; p.rewards = rewards; p.entropy = entropy;
    p.done = done; p.bootstrap = bootstrap; p.vs = vs_out; p.adv = adv_out; p.dlp = dlp_out; p.dv = dV_out;
    p.loss4 = loss4_out; p.rows = rows; p.T = T; p.in_stride = in_t_stride; p.out_stride = out_t_stride;
    p.gamma = gamma; p.rho_bar = rho_bar; p.c_bar = c_bar; p.pg_rho_bar = pg_rho_bar; p.ent_coef = ent_coef;
    p.inv_n = inv_n > 0.0f ? inv_n : float(1.0 / (double(rows) * double(T)));
    p.ticket = reinterpret_cast<unsigned int*>(workspace);
    p.partials = reinterpret_cast<double*>(reinterpret_cast<uint8_t*>(workspace) + 256);
    const unsigned grid = unsigned((rows + 127) / 128);
    vtrace_tmajor_kernel<<<grid, 128, 0, st>>>(p);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}
