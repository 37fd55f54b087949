// "Exact torch-RNG" sampler mode of the policy head, for sm_100a.
//
// Reference: model.py:177 -- Categorical.sample() -> torch.multinomial(probs, 1, True), which for one draw is
// argmax_i(p_i / q_i) with q_i ~ Exp(1) drawn per CATEGORY from torch's generator.  The inverse-CDF
// sampler of csrc/head.cu takes one uniform per component and is defined by the oracle; this mode takes the
// exponentials themselves, [n_frames, 78*cells] like the logits, so that a caller who draws them with torch's
// generator (torch.empty_like(logits).exponential_()) samples the way the reference does.
//
// One thread per (frame, cell): 78 logits, 78 mask bytes, 78 exponentials in, 7 actions out.  Acting mode runs on
// a handful of frames per call (n_envs = 6 in microbeast.py:114), so this is a latency kernel, not a bandwidth one;
// log-prob / entropy of the sampled actions come from mb_head_eval_fwd.
#include "common.cuh"

#include "../../include/mbeast.h"

namespace mb {
namespace {

template <int OFF, int LEN>
__device__ __forceinline__ int seg_sample_exp(const float* __restrict__ z, const uint8_t* __restrict__ m,
                                              const float* __restrict__ q) {
    float zz[LEN];
    float mx = -3.4e38f;
#pragma unroll
    for (int i = 0; i < LEN; ++i) {
        zz[i] = m[OFF + i] ? z[OFF + i] : kMaskFill;                   // model.py:43
        mx = fmaxf(mx, zz[i]);
    }
    // p_i = e_i / S; the common factor 1/S does not change the argmax, but it is applied so that the compared
    // values are the ones torch compares (probs / q)
    float e[LEN], S = 0.f;
#pragma unroll
    for (int i = 0; i < LEN; ++i) {
        e[i] = __expf(zz[i] - mx);
        S += e[i];
    }
    const float inv = 1.0f / S;
    int best = 0;
    float best_v = -1.0f;
#pragma unroll
    for (int i = 0; i < LEN; ++i) {
        const float v = (e[i] * inv) / q[OFF + i];
        if (v > best_v) {                                               // first maximum wins, like argmax
            best_v = v;
            best = i;
        }
    }
    return best;
}

__global__ void __launch_bounds__(128) head_sample_exp_kernel(const float* __restrict__ logits, const uint8_t* __restrict__ mask,
                                                              const float* __restrict__ expo, int64_t* __restrict__ action,
                                                              long long n_cells) {
    const long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (c >= n_cells) return;
    const float* z = logits + c * kLogitsPerCell;
    const uint8_t* m = mask + c * kLogitsPerCell;
    const float* q = expo + c * kLogitsPerCell;
    int64_t* a = action + c * kCompsPerCell;
    a[0] = seg_sample_exp<0, 6>(z, m, q);
    a[1] = seg_sample_exp<6, 4>(z, m, q);
    a[2] = seg_sample_exp<10, 4>(z, m, q);
    a[3] = seg_sample_exp<14, 4>(z, m, q);
    a[4] = seg_sample_exp<18, 4>(z, m, q);
    a[5] = seg_sample_exp<22, 7>(z, m, q);
    a[6] = seg_sample_exp<29, 49>(z, m, q);
}

}  // namespace
}  // namespace mb

extern "C" int mb_head_sample_exp(const float* logits, const uint8_t* mask, const float* exponentials, int64_t* action_out,
                                  int64_t n_frames, int32_t cells, void* stream) {
    if (n_frames < 0 || cells <= 0) {
        mb::set_error("mb_head_sample_exp: n_frames=%lld cells=%d out of range", (long long)n_frames, cells);
        return MB_ERR_INVALID;
    }
    if (n_frames == 0) return MB_OK;
    if (!logits || !mask || !exponentials || !action_out) {
        mb::set_error("mb_head_sample_exp: null pointer");
        return MB_ERR_INVALID;
    }
    const long long n_cells = (long long)n_frames * cells;
    mb::head_sample_exp_kernel<<<unsigned((n_cells + 127) / 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(
        logits, mask, exponentials, action_out, n_cells);
    MB_CUDA(cudaGetLastError());
    return MB_OK;
}
