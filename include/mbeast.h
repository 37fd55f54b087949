/* microbeast_b200 -- C ABI of the B200 (sm_100a) learner hot path.
 *
 * The reference (Neos-codes/microbeast) has no FFI: its seam for this path is the Python call
 * surface of model.py / libs/utils.py.  These entry points are what a binding for that seam
 * calls; each comment names the reference code the entry point replaces.  Plain pointers and
 * sizes only (no torch types).  See INTEGRATION.md for the ctypes binding a maintainer adds.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless stated; tensors are contiguous row-major;
 *  - the library never allocates or frees device memory; callers own every buffer;
 *  - every call is an asynchronous launch on `stream` (a cudaStream_t passed as void*; NULL =
 *    the legacy default stream); no call synchronises or reads device results on the host;
 *  - return value 0 = ok; otherwise an MB_ERR_* code and mb_last_error() holds a message
 *    (thread-local string, valid until the next failing call on the same thread);
 *  - `status` (optional, may be NULL) is a device int32 the kernels OR bits into:
 *      MB_STATUS_BAD_ACTION  an action index outside [0, n_k) was seen (clamped; torch's
 *                            Categorical(validate_args=True) raises ValueError there)
 *      MB_STATUS_TIMEOUT     an internal pipeline wait gave up (kernel bug guard)
 *    callers zero it once and check it whenever they synchronise anyway.
 *
 * Geometry: a map of `cells` = h*w cells; 78 logits / mask bytes and 7 components per cell,
 * split 6/4/4/4/4/7/49 (reference microbeast.py:145, libs/utils.py:40-44).
 */
#ifndef MBEAST_H_
#define MBEAST_H_

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define MB_API __attribute__((visibility("default")))
#else
#define MB_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define MB_OK 0
#define MB_ERR_INVALID 1
#define MB_ERR_CUDA 2
#define MB_ERR_UNSUPPORTED 3

#define MB_STATUS_BAD_ACTION 1
#define MB_STATUS_TIMEOUT 2

/* Library ABI version (major*100 + minor). */
MB_API int mb_version(void);

/* Message of the last failing call on this thread ("" if none). */
MB_API const char* mb_last_error(void);

/* Learner branch of Agent.get_action after the actor Linear: reference model.py:180-214
 * (torch.split -> 7*cells CategoricalMasked objects, model.py:33-52 -> log_prob / entropy ->
 * sums over the 7*cells components of a frame).
 *   logits [n_frames, 78*cells] f32, mask [n_frames, 78*cells] u8 (non-zero = valid),
 *   action [n_frames, 7*cells] i64  ->  logp_out [n_frames] f32, ent_out [n_frames] f32. */
MB_API int mb_head_eval_fwd(const float* logits, const uint8_t* mask, const int64_t* action, float* logp_out,
                     float* ent_out, int64_t n_frames, int32_t cells, int32_t* status, void* stream);

/* Autograd of the above (reference: total_loss.backward(), libs/utils.py:334, through the
 * ~10*7*cells graph nodes built by model.py:190-196).  g_logp / g_ent [n_frames] are the
 * upstream gradients of logp_out / ent_out; dlogits_out [n_frames, 78*cells] f32 is written in
 * full (masked entries and fully masked components get exactly 0). */
MB_API int mb_head_eval_bwd(const float* logits, const uint8_t* mask, const int64_t* action, const float* g_logp,
                     const float* g_ent, float* dlogits_out, int64_t n_frames, int32_t cells, int32_t* status,
                     void* stream);

/* Actor branch of Agent.get_action: reference model.py:171-177,194-210 (Categorical.sample ->
 * torch.multinomial on the global RNG).  Sampling here is inverse-CDF from caller-supplied
 * uniforms [n_frames, 7*cells] f32 in [0,1) as defined in oracle/head.py:sample_invcdf
 * (a = #{k < n-1 : u >= cdf_k}, fp64-exact decision).  action_out [n_frames, 7*cells] i64,
 * logp_out [n_frames] f32 (log-prob of the sampled action), ent_out [n_frames] f32 or NULL. */
MB_API int mb_head_sample(const float* logits, const uint8_t* mask, const float* uniforms, int64_t* action_out,
                   float* logp_out, float* ent_out, int64_t n_frames, int32_t cells, int32_t* status,
                   void* stream);

/* The reference's own sampling rule with caller-supplied randomness: reference model.py:177, Categorical.sample() ->
 * torch.multinomial(probs, 1, True) = argmax_i(probs_i / q_i), q_i ~ Exp(1) per category.  exponentials
 * [n_frames, 78*cells] f32 > 0 (torch.empty_like(logits).exponential_() draws them the way torch.multinomial does);
 * action_out [n_frames, 7*cells] i64.  A fully masked component has probs = 1/n (quirk Q1) and so samples uniformly;
 * masked entries of a partly valid component have probability 0 and are never chosen.  Log-prob and entropy of the
 * result: mb_head_eval_fwd. */
MB_API int mb_head_sample_exp(const float* logits, const uint8_t* mask, const float* exponentials, int64_t* action_out,
                              int64_t n_frames, int32_t cells, void* stream);

/* Cell-list forms of the two eval entry points, for a sparse actor head that computes logits only for
 * cells with at least one valid action (a fully masked cell contributes exactly 0 to log-prob, entropy
 * and gradients whatever its logits are: reference model.py:43 + fp32 absorption, quirk Q1).
 *   logits / mask [n_cells, 78], action [n_cells, 7]; n_cells a multiple of 64 (pad with masked cells);
 *   cell_out4 [4, n_cells] f32 = log-prob of the six small components, their entropy, log-prob of the
 *   49-way component, its entropy (the caller adds the pairs and scatters them to frames);
 *   g_logp_cell / g_ent_cell [n_cells] upstream gradients per cell; dlogits_out [n_cells, 78]. */
MB_API int mb_head_eval_cells_fwd(const float* logits, const uint8_t* mask, const int64_t* action,
                                  float* cell_out4, int64_t n_cells, int32_t* status, void* stream);
MB_API int mb_head_eval_cells_bwd(const float* logits, const uint8_t* mask, const int64_t* action,
                                  const float* g_logp_cell, const float* g_ent_cell, float* dlogits_out,
                                  int64_t n_cells, int32_t* status, void* stream);

/* Bytes of scratch mb_vtrace_loss needs for `rows` rows (ticket word + per-CTA partial sums). */
MB_API size_t mb_vtrace_workspace_bytes(int64_t rows);

/* V-trace + losses: the inline block of PPO_learn, reference libs/utils.py:277-329, and the
 * gradients autograd derives from utils.py:320-329.
 *   inputs  [rows, T], time contiguous, `in_row_stride` elements between rows:
 *           lp_old (behaviour log-prob), lp_new (learner log-prob), values, rewards, entropy f32,
 *           done u8/bool;  bootstrap [rows] f32 (= V at slot T, utils.py:258)
 *   outputs [rows, T], `out_row_stride` elements between rows: vs_out, adv_out f32 and, when both
 *           are non-NULL, dlp_out = adv * inv_n and dV_out = -(vs - V) * inv_n
 *           (d total / d entropy is the constant -ent_coef * inv_n);
 *           loss4_out [4] f32 = {pg, value, entropy, total} with means taken as sum * inv_n.
 *   gamma, rho_bar, c_bar, pg_rho_bar, ent_coef: the reference hard-codes 0.99, 1, 1, 1, 0.01.
 *   inv_n: 1 / (global rows * T) for data-parallel shards; <= 0 means 1 / (rows * T).
 *   workspace: >= mb_vtrace_workspace_bytes(rows) bytes, 8-byte aligned; its first 4 bytes must
 *           be zero before the FIRST call (the kernel restores the zero itself), and two calls
 *           that may overlap on the GPU must not share a workspace.
 * Uses TMA tiles when the f32 pointers are 16-byte aligned and both row strides are multiples of 4
 * (`done` rides along as a TMA tile when in_row_stride % 16 == 0, else it is read directly); any
 * other layout takes a slower generic kernel with equal results. */
MB_API int mb_vtrace_loss(const float* lp_old, const float* lp_new, const float* values, const float* bootstrap,
                   const float* rewards, const uint8_t* done, const float* entropy, int64_t rows, int32_t T,
                   int64_t in_row_stride, int64_t out_row_stride, float gamma, float rho_bar, float c_bar,
                   float pg_rho_bar, float ent_coef, float inv_n, float* vs_out, float* adv_out, float* loss4_out,
                   float* dlp_out, float* dV_out, void* workspace, size_t workspace_bytes, void* stream);

/* mb_vtrace_loss on TIME-MAJOR arrays: every [rows, T] tensor of mb_vtrace_loss is given as [T, rows]
 * (rows contiguous, `in_t_stride` / `out_t_stride` elements between consecutive time steps).  This is
 * the order the reference's rollout buffers are written in and get_batch flattens them in
 * (libs/utils.py:35-45,197-205: [T+1, n_envs, ...]); the reference reshapes that to [R, T+1]
 * (utils.py:245,253), which this entry point makes unnecessary when a row is meant to be one
 * trajectory.  Same arithmetic, arguments, outputs (time-major) and workspace rules otherwise. */
MB_API int mb_vtrace_loss_tmajor(const float* lp_old, const float* lp_new, const float* values,
                                 const float* bootstrap, const float* rewards, const uint8_t* done,
                                 const float* entropy, int64_t rows, int32_t T, int64_t in_t_stride,
                                 int64_t out_t_stride, float gamma, float rho_bar, float c_bar,
                                 float pg_rho_bar, float ent_coef, float inv_n, float* vs_out, float* adv_out,
                                 float* loss4_out, float* dlp_out, float* dV_out, void* workspace,
                                 size_t workspace_bytes, void* stream);

/* mb_vtrace_loss without the workspace argument.  It has no counterpart in the reference: it exists
 * for users of this library who cannot allocate device memory where they call it (a C / ctypes host
 * with no allocator of its own, a step loop that must stay clear of cudaMalloc's device-wide
 * synchronisation, a thread running while another stream of the process is under capture, when the
 * runtime refuses cudaMalloc).  The scratch is a fixed 64 KiB array inside the library image, one
 * per GPU that loads the library, kept until the process exits.
 *   - rows <= 65536 per call; more is rejected with MB_ERR_INVALID (use mb_vtrace_loss then);
 *   - the first call on each device looks the array's device address up (cudaGetSymbolAddress, with
 *     that device current); make that first call outside any stream capture -- a first call whose
 *     `stream` is being captured is refused with MB_ERR_INVALID and leaves the capture intact;
 *   - device indices 0..63 only (MB_ERR_UNSUPPORTED beyond);
 *   - the call never allocates or frees device memory; it clears the part of the scratch that must
 *     start at zero on `stream` before the launch;
 *   - every call on a GPU shares the one array: calls that may overlap on the GPU race with each
 *     other -- keep them on one stream or order them with events.
 * All other arguments, outputs and return codes are those of mb_vtrace_loss. */
MB_API int mb_vtrace_loss_static(const float* lp_old, const float* lp_new, const float* values,
                                 const float* bootstrap, const float* rewards, const uint8_t* done,
                                 const float* entropy, int64_t rows, int32_t T, int64_t in_row_stride,
                                 int64_t out_row_stride, float gamma, float rho_bar, float c_bar,
                                 float pg_rho_bar, float ent_coef, float inv_n, float* vs_out, float* adv_out,
                                 float* loss4_out, float* dlp_out, float* dV_out, void* stream);

/* Wire-format expansion of the rollout store (no reference counterpart; it restores the dtypes of
 * reference libs/utils.py:43-44 on the device after a compact host->device copy).
 *   mask_bits [n_frames, ceil(78*cells/8)] u8, bit i of byte j = mask entry 8j+i of the frame
 *        ->  mask_out [n_frames, 78*cells] u8 (0/1);
 *   action_u8 [n_frames, 7*cells] u8  ->  action_out [n_frames, 7*cells] i64.
 * Either pair may be NULL/NULL to skip it. */
MB_API int mb_unpack_rollout(const uint8_t* mask_bits, const uint8_t* action_u8, uint8_t* mask_out,
                             int64_t* action_out, int64_t n_frames, int32_t cells, void* stream);

/* Observation planes on the wire (no reference counterpart; replaces the float cast of reference
 * env_packer.py:8-10 and the permute of model.py:156-160 after a 27x smaller copy than float32):
 *   obs_bits [n_frames, cells] u32, bit c = feature plane c of the cell (planes are 0/1)
 *        ->  obs_out [n_frames, planes, cells] f32, i.e. NCHW with cells = h*w. */
MB_API int mb_unpack_obs(const uint32_t* obs_bits, float* obs_out, int64_t n_frames, int32_t cells,
                         int32_t planes, void* stream);

/* active[c] = 1 if cell c of mask [n_cells,78] has any valid action, else 0 (the cell list of the
 * sparse actor head; the reference has no counterpart -- it evaluates every cell,
 * model.py:98-104, and fully masked cells contribute exactly 0, quirk Q1). */
MB_API int mb_active_cells(const uint8_t* mask, uint8_t* active, int64_t n_cells, void* stream);

/* Weight gradient of a 3x3 / stride 1 / pad 1 convolution in fp32 (the encoder's convolutions,
 * reference model.py:56-109, differentiated by loss.backward() at libs/utils.py:334):
 *   dW[co,ci,ky,kx] = sum_{n,y,x} dy[n,co,y,x] * x[n,ci,y+ky-1,x+kx-1],  x [n,cin,h,w], dy [n,cout,h,w],
 * both NCHW contiguous, 16-byte aligned.  The kernel writes one partial result per CTA:
 * partials [n_blocks][cout*cin*9 + cout] -- dW, then the bias gradient sum_{n,y,x} dy[n,co,y,x]; the caller
 * adds the slots up (fixed order => deterministic).
 * mb_conv3x3_wgrad_blocks returns n_blocks for a shape on the current device, 0 when the shape is not
 * covered (the caller keeps its own path: square maps 2..16, (cin,cout) in {(27,16),(16,16),(16,32),
 * (32,32)} as instantiated in csrc/convgrad.cu), -1 on error. */
MB_API int mb_conv3x3_wgrad_blocks(int64_t n_frames, int32_t cin, int32_t cout, int32_t h, int32_t w);
MB_API int mb_conv3x3_wgrad(const float* x, const float* dy, float* partials, int32_t n_blocks, int64_t n_frames,
                            int32_t cin, int32_t cout, int32_t h, int32_t w, void* stream);

/* Forward-type 3x3 / stride 1 / pad 1 convolution in fp32 (reference model.py:61-62,88: nn.Conv2d(.., 3, padding=1)):
 *   out[n,co,y,x] = bias[co] + sum_{ci,ky,kx} wk[ci][co][ky*3+kx] * in[n,ci,y+ky-1,x+kx-1]
 * in [n,cin,h,w], out [n,cout,h,w] NCHW contiguous, 16-byte aligned; bias [cout] or NULL.  The kernel weights
 * come in one of three layouts: wmode 0 = wk [cin][cout][9] as written above; 1 = a torch Conv2d weight
 * [cout][cin][3][3] (forward pass); 2 = the data gradient of a convolution whose torch weight is
 * [cin][cout][3][3] (here cin / cout are those of THIS operation: the gradient's channels in, the input's
 * channels out): the kernel flips the taps itself, pass no bias.  mb_conv3x3_fwd_supported returns 1 for the
 * shapes instantiated in csrc/convgrad.cu, else 0 (the caller keeps cuDNN). */
MB_API int mb_conv3x3_fwd_supported(int32_t cin, int32_t cout, int32_t h, int32_t w);
MB_API int mb_conv3x3_fwd(const float* in, const float* wk, const float* bias, float* out, int64_t n_frames,
                          int32_t cin, int32_t cout, int32_t h, int32_t w, int32_t wmode, void* stream);

/* The encoder's FIRST convolution (reference model.py:88,121: Conv2d(27, 16, 3, padding=1) on the 27 binary feature
 * planes, model.py:156-160) straight from bit-packed observations: obs_bits [n_frames, h*w] i32, bit c = plane c (the
 * rollout store's wire format, see mb_unpack_obs); weight [16, 27, 3, 3] (torch layout), bias [16] or NULL;
 * out [n_frames, 16, h, w] f32.  out = bias + the sum of the weight columns of the set bits under each tap (any bit
 * pattern; fp32 additions only).  planes must be 27, cout 16, h == w in {8, 16}; other shapes: MB_ERR_UNSUPPORTED. */
MB_API int mb_conv1_bits_fwd(const int32_t* obs_bits, const float* weight, const float* bias, float* out, int64_t n_frames,
                             int32_t planes, int32_t cout, int32_t h, int32_t w, void* stream);

/* Its weight and bias gradients (what autograd of model.py:88 gives for a 0/1 input), again from the bits:
 * dy [n_frames, 16, h, w] f32 (16-byte aligned); accumulate_out [16*27*9 + 16] f32, ZEROED by the caller (or holding a
 * running total): the kernel ADDS dW [16][27][3][3] and then db [16] into it with atomics.  Each (plane, frame) costs
 * work in proportion to its set cells -- or its unset ones when more than half are set -- so the expanded planes are
 * never formed in the backward pass either.  Same shape limits as mb_conv1_bits_fwd. */
MB_API int mb_conv1_bits_wgrad(const int32_t* obs_bits, const float* dy, float* accumulate_out, int64_t n_frames,
                               int32_t planes, int32_t cout, int32_t h, int32_t w, void* stream);

/* The same two kernels with the element-wise neighbours of a residual block (reference model.py:67-73:
 * relu -> conv0 -> relu -> conv1 -> + x, and its autograd) folded in, so that ReLU, the residual add and ReLU's
 * backward cost no pass over the activations of their own:
 *   relu_in != 0   the convolution reads max(in, 0) (forward: conv(relu(x)));
 *   mask_src       [like out] or NULL: out is zeroed where mask_src <= 0 (backward: the data gradient through the
 *                  ReLU that preceded the convolution, mask_src = that ReLU's input);
 *   add_src        [like out] or NULL: added to out after the mask (forward: the skip connection; backward: the
 *                  gradient that bypasses the block);
 *   relu_x != 0    mb_conv3x3_wgrad_relu differentiates conv(relu(x)) with respect to the weights: x is rectified
 *                  as it is read;
 *   accumulate_out [cout*cin*9 + cout] or NULL: every CTA adds its share to this (caller-zeroed) result with atomics
 *                  instead of writing a slot of `partials` (which may then be NULL): no summation pass afterwards, but
 *                  the order of the fp32 additions varies from run to run. */
MB_API int mb_conv3x3_fwd_fused(const float* in, const float* wk, const float* bias, float* out, int64_t n_frames,
                                int32_t cin, int32_t cout, int32_t h, int32_t w, int32_t wmode, int32_t relu_in,
                                const float* mask_src, const float* add_src, void* stream);
MB_API int mb_conv3x3_wgrad_relu(const float* x, const float* dy, float* partials, int32_t n_blocks, int64_t n_frames,
                                 int32_t cin, int32_t cout, int32_t h, int32_t w, int32_t relu_x, float* accumulate_out,
                                 void* stream);

/* Actor Linear forward (reference model.py:136,167: logits = actor(hidden)) on the tcgen05 tensor cores at
 * fp32 accuracy: c[m,n] = a[m,k] * b[n,k]^T + bias[n] with both operands given as a TF32 "hi" part and a TF32
 * "lo" remainder (mb_split_tf32: x = hi + lo up to 2^-22 |x|) and three TF32 products accumulated in fp32:
 * a_hi*b_hi + a_lo*b_hi + a_hi*b_lo.  a_* [m,k], b_* [n,k], c [m,n] row-major fp32, 16-byte aligned;
 * k % 32 == 0, n % 4 == 0; bias [n] or NULL; status = optional device word (pipeline-timeout bit). */
MB_API int mb_split_tf32(const float* x, float* hi, float* lo, int64_t n, void* stream);
MB_API int mb_linear_tf32x3(const float* a_hi, const float* a_lo, const float* b_hi, const float* b_lo,
                            const float* bias, float* c, int64_t m, int64_t n, int32_t k, int32_t* status,
                            void* stream);

/* Actor Linear + policy head in one tensor-core kernel (reference model.py:167 followed by model.py:168-214; the
 * learner, libs/utils.py:237-329, reads only the per-frame log-prob / entropy sums, so the [n_frames, 78*cells] logits
 * need not exist in memory).  Three steps:
 *   mb_pack_cells         mask [n_frames, 78*cells] u8 + action [n_frames, 7*cells] i64 -> cell_meta [cells][n_frames]
 *                         records of 16 B (78 mask bits, seven 6-bit action indices); out-of-range actions are clamped
 *                         and flagged in `status` (torch validate_args raises there, model.py:38).  alive_bits (optional)
 *                         [cells][ceil(n_frames/32)] u32: bit i of word w = frame 32w+i has a valid action in the cell
 *                         (what mb_cells_bwd builds its lists from).
 *   mb_pack_actor_weight  weight [78*cells, k], bias [78*cells] -> TF32 hi / lo copies w_hi, w_lo [80*cells, k] with
 *                         every cell's 78 rows padded to 80, bias_p [80*cells].  Once per weight update.
 *   mb_actor_head_fwd     a_hi, a_lo [n_frames, k] = the TF32 hi / lo parts of `hidden` (mb_split_tf32); three TF32
 *                         products accumulated in fp32 on tcgen05 (as mb_linear_tf32x3), the epilogue adds the bias and
 *                         evaluates the head of each (frame, cell) from tensor memory:
 *                         cell_out [cells][n_frames][2] f32 = (log-prob, entropy) of the cell's seven components;
 *                         logp_out / ent_out [n_frames] (both or neither) = their sums in cell order;
 *                         logits_out [n_frames, 78*cells] or NULL = policy_logits (model.py:208) on request; with
 *                         logits_alive_only != 0 only cells that have a valid action are written (the rest of the
 *                         buffer is left untouched) -- all a backward pass needs, at a fraction of the traffic.
 * k % 32 == 0; every pointer 16-byte aligned. */
MB_API int mb_pack_cells(const uint8_t* mask, const int64_t* action, void* cell_meta, void* alive_bits, int64_t n_frames,
                         int32_t cells, int32_t* status, void* stream);
MB_API int mb_pack_actor_weight(const float* weight, const float* bias, float* w_hi, float* w_lo, float* bias_p,
                                int32_t cells, int32_t k, void* stream);
MB_API int mb_actor_head_fwd(const float* a_hi, const float* a_lo, const float* w_hi, const float* w_lo,
                             const float* bias_p, const void* cell_meta, float* cell_out, float* logits_out,
                             int32_t logits_alive_only, float* logp_out, float* ent_out, int64_t n_frames,
                             int32_t cells, int32_t k, int32_t* status, void* stream);

/* The same forward over the (frame, cell) pairs that have a valid action only: log-prob and entropy of every other pair
 * are exactly 0 (quirk Q1) and the learner never reads policy_logits (libs/utils.py:237-329), so their 78 logits are
 * not formed at all -- 6 % of the work at microRTS densities.  alive_bits [cells][ceil(n_frames / 32)] as written by
 * mb_pack_cells; workspace of mb_actor_head_list_workspace_bytes(n_frames, cells) bytes (device memory, 16-byte
 * aligned: the cell-ordered lists are built in it on the device, no host synchronisation); cell_out is zeroed by the
 * call and filled for the live pairs; logits_out [n_frames, 78 * cells] or NULL receives the logits of the live cells
 * (what mb_cells_bwd needs), other cells are left untouched.  Logits, cell_out, logp_out, ent_out are bit-identical
 * to mb_actor_head_fwd's.  cells <= 1024, k % 32 == 0. */
MB_API size_t mb_actor_head_list_workspace_bytes(int64_t n_frames, int32_t cells);
MB_API int mb_actor_head_list_fwd(const float* a_hi, const float* a_lo, const float* w_hi, const float* w_lo,
                                  const float* bias_p, const void* cell_meta, const void* alive_bits, float* cell_out,
                                  float* logits_out, float* logp_out, float* ent_out, void* workspace, size_t workspace_bytes,
                                  int64_t n_frames, int32_t cells, int32_t k, int32_t* status, void* stream);

/* Backward of actor Linear + policy head over the cells that can have a gradient (reference: loss.backward(),
 * libs/utils.py:334, through model.py:167-214; a cell without a valid action has exactly zero gradient, quirk Q1).
 * Everything runs on the device, nothing synchronises with the host: per cell the frames with a valid action are
 * listed in ascending order, dlogits of those (frame, cell) pairs is formed from the pair's logits (`logits`
 * [n_frames, 78*cells]: only the cells with a valid action are read, which is what mb_actor_head_fwd writes with
 * logits_alive_only), its cell record and the frame's upstream gradients g_logp / g_ent [n_frames], and two fp32 FMA
 * products give
 *   d_hidden [n_frames, 256] = sum over the frame's alive cells of dlogits . W_cell      (zeroed, then vector atomics)
 *   d_weight [78*cells, 256] = sum over the cell's alive frames of dlogits^T hidden_row  (zeroed, then vector atomics)
 *   d_bias   [78*cells]      = sum of dlogits.
 * hidden [n_frames, 256], weight [78*cells, 256], cell_meta and alive_bits from mb_pack_cells.  k must be 256,
 * cells <= 1024.
 * workspace: >= mb_cells_bwd_workspace_bytes(n_frames, cells) bytes (the lists, sized for the worst case), 16-byte
 * aligned, contents irrelevant.  Atomic accumulation order varies from run to run (as torch's index_add_ does). */
MB_API size_t mb_cells_bwd_workspace_bytes(int64_t n_frames, int32_t cells);
MB_API int mb_cells_bwd(const float* hidden, const float* weight, const float* logits, const void* cell_meta,
                        const void* alive_bits, const float* g_logp, const float* g_ent, float* d_hidden, float* d_weight,
                        float* d_bias,
                        void* workspace, size_t workspace_bytes, int64_t n_frames, int32_t cells, int32_t k, void* stream);

/* Backward of max_pool2d(kernel 3, stride 2, padding 1) (reference model.py:96) from the argmax
 * indices of the forward: dy, idx [planes, (h+1)/2, (w+1)/2] (idx = flat position in the h*w input
 * plane, as torch.nn.functional.max_pool2d(..., return_indices=True) returns it), dx [planes, h, w]. */
MB_API int mb_maxpool3s2_bwd(const float* dy, const int64_t* idx, float* dx, int64_t planes, int32_t h, int32_t w,
                             void* stream);

/* max_pool2d(kernel 3, stride 2, padding 1) forward (reference model.py:96) with a one-byte argmax: y, code
 * [planes, (h+1)/2, (w+1)/2]; code = ky * 3 + kx of the winner inside its window, first maximum in row-major scan
 * order (ATen's tie rule; NaN wins like there).  mb_maxpool3s2_bwd_code is the matching backward:
 * dx [planes, h, w] = sum of dy over the windows whose winner the pixel is. */
MB_API int mb_maxpool3s2_fwd(const float* x, float* y, uint8_t* code, int64_t planes, int32_t h, int32_t w, void* stream);
MB_API int mb_maxpool3s2_bwd_code(const float* dy, const uint8_t* code, float* dx, int64_t planes, int32_t h, int32_t w,
                                  void* stream);

/* Tuning knobs for benchmarking (not part of the reference surface).  Keys:
 *   "head_ctas_per_sm" (0 = as many as fit), "head_grid" (0 = auto), "head_variant" (tile /
 *   pipeline shape: 0 = 64 cells x 3 stages, 1 = 64x2, 2 = 64x4, 3 = 32x3, 4 = 32x5, 5 = 128x3, 6 = 128x2, 7 = 32x2; -1 = per-kernel default),
 *   "vtrace_static_poison" (1: the next mb_vtrace_loss_static call first fills its built-in scratch with
 *   0xA5 on the caller's stream -- a test hook that stands in for whatever an earlier user left there),
 *   "vtrace_force_generic" (0/1), "vtrace_l2_promo" (0/64/128/256 bytes),
 *   "vtrace_variant" (stages x steps per tile: -1 = by grid size, 0 = 2x32, 1 = 3x32, 2 = 4x32,
 *   3 = 3x16, 4 = 4x16, 5 = 5x16, 6 = 6x16, 7 = 8x16, 8 = 6x32, 9 = 12x16, 10 = 8x32; with 16-row instead of
 *   32-row groups: 11 = 4x32, 12 = 6x32, 13 = 8x32, 14 = 8x16, 15 = 6x16).
 *   "gemm_group_m" (mb_linear_tf32x3 tile order: 0 = row-major, g > 0 = bands of g tile rows walked column by
 *   column), "gemm_stream_store" (0/1: store the output with the streaming cache hint).
 * Returns MB_ERR_INVALID for an unknown key. */
MB_API int mb_set_tuning(const char* key, int32_t value);

#ifdef __cplusplus
}
#endif
#endif /* MBEAST_H_ */
