// extern "C" surface of libmbeast_sm100.so (declared in include/mbeast.h) and host plumbing.
#include "common.cuh"

#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include "../../include/mbeast.h"

namespace mb {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
    set_error("CUDA error %d (%s) in %s", int(e), cudaGetErrorString(e), what);
    return MB_ERR_CUDA;
}

int sm_count(int device) {
    static int cached[64] = {};
    if (device >= 0 && device < 64 && cached[device] > 0) return cached[device];
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || n <= 0) n = 148;
    if (device >= 0 && device < 64) cached[device] = n;
    return n;
}

// implemented in head.cu / vtrace.cu
int head_eval_fwd(const float*, const uint8_t*, const int64_t*, float*, float*, long long, int, int*, cudaStream_t);
int head_eval_bwd(const float*, const uint8_t*, const int64_t*, const float*, const float*, float*, long long, int,
                  int*, cudaStream_t);
int head_sample(const float*, const uint8_t*, const float*, int64_t*, float*, float*, long long, int, int*,
                cudaStream_t);
int head_eval_cells_fwd(const float*, const uint8_t*, const int64_t*, float*, long long, int*, cudaStream_t);
int head_eval_cells_bwd(const float*, const uint8_t*, const int64_t*, const float*, const float*, float*, long long, int*,
                        cudaStream_t);
void head_set_tuning(int ctas_per_sm, int grid_override, int variant);
void vtrace_set_tuning(int force_generic, int variant, int l2_promo);
void gemm_set_tuning(int group_m, int stream_store);
void conv_set_tuning(int tensor_cores);
void conv_tc_set_debug(int bits);
void head_list_set_debug(int bits);
void vtrace_static_poison(int on);
void vtrace_set_debug(int v);
size_t vtrace_workspace_bytes(long long rows);

struct VtParams;
}  // namespace mb

// VtParams is defined in vtrace.cu; the C entry point lives there too (mb_vtrace_loss) so the
// struct stays private to that translation unit.

extern "C" {

int mb_version(void) { return 100; }

const char* mb_last_error(void) { return mb::g_err; }

int mb_head_eval_fwd(const float* logits, const uint8_t* mask, const int64_t* action, float* logp_out,
                     float* ent_out, int64_t n_frames, int32_t cells, int32_t* status, void* stream) {
    return mb::head_eval_fwd(logits, mask, action, logp_out, ent_out, n_frames, cells, status,
                             static_cast<cudaStream_t>(stream));
}

int mb_head_eval_bwd(const float* logits, const uint8_t* mask, const int64_t* action, const float* g_logp,
                     const float* g_ent, float* dlogits_out, int64_t n_frames, int32_t cells, int32_t* status,
                     void* stream) {
    return mb::head_eval_bwd(logits, mask, action, g_logp, g_ent, dlogits_out, n_frames, cells, status,
                             static_cast<cudaStream_t>(stream));
}

int mb_head_sample(const float* logits, const uint8_t* mask, const float* uniforms, int64_t* action_out,
                   float* logp_out, float* ent_out, int64_t n_frames, int32_t cells, int32_t* status,
                   void* stream) {
    return mb::head_sample(logits, mask, uniforms, action_out, logp_out, ent_out, n_frames, cells, status,
                           static_cast<cudaStream_t>(stream));
}

int mb_head_eval_cells_fwd(const float* logits, const uint8_t* mask, const int64_t* action, float* cell_out4,
                           int64_t n_cells, int32_t* status, void* stream) {
    return mb::head_eval_cells_fwd(logits, mask, action, cell_out4, n_cells, status, static_cast<cudaStream_t>(stream));
}

int mb_head_eval_cells_bwd(const float* logits, const uint8_t* mask, const int64_t* action, const float* g_logp_cell,
                           const float* g_ent_cell, float* dlogits_out, int64_t n_cells, int32_t* status,
                           void* stream) {
    return mb::head_eval_cells_bwd(logits, mask, action, g_logp_cell, g_ent_cell, dlogits_out, n_cells, status,
                                   static_cast<cudaStream_t>(stream));
}

size_t mb_vtrace_workspace_bytes(int64_t rows) { return mb::vtrace_workspace_bytes(rows); }

int mb_set_tuning(const char* key, int32_t value) {
    static int head_ctas = 0, head_grid = 0, head_variant = -1, vt_generic = 0, vt_variant = -1, vt_promo = 256;
    static int gemm_group_m = 0, gemm_stream_store = 0;
    if (!key) {
        mb::set_error("mb_set_tuning: null key");
        return MB_ERR_INVALID;
    }
    if (!strcmp(key, "head_ctas_per_sm")) head_ctas = value;
    else if (!strcmp(key, "head_grid")) head_grid = value;
    else if (!strcmp(key, "head_variant")) head_variant = value;
    else if (!strcmp(key, "vtrace_force_generic")) vt_generic = value;
    else if (!strcmp(key, "vtrace_variant")) vt_variant = value;
    else if (!strcmp(key, "vtrace_l2_promo")) vt_promo = value;
    else if (!strcmp(key, "vtrace_static_poison")) mb::vtrace_static_poison(value);
    else if (!strcmp(key, "vtrace_debug")) mb::vtrace_set_debug(value);
    else if (!strcmp(key, "gemm_group_m")) gemm_group_m = value;
    else if (!strcmp(key, "gemm_stream_store")) gemm_stream_store = value;
    else if (!strcmp(key, "conv_tc")) mb::conv_set_tuning(value);
    else if (!strcmp(key, "conv_tc_debug")) mb::conv_tc_set_debug(value);
    else if (!strcmp(key, "head_list_debug")) mb::head_list_set_debug(value);
    else {
        mb::set_error("mb_set_tuning: unknown key '%s'", key);
        return MB_ERR_INVALID;
    }
    mb::head_set_tuning(head_ctas, head_grid, head_variant);
    mb::vtrace_set_tuning(vt_generic, vt_variant, vt_promo);
    mb::gemm_set_tuning(gemm_group_m, gemm_stream_store);
    return MB_OK;
}

}  // extern "C"
