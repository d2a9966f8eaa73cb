/*
 * ngp_b200.h — C ABI of the B200-native neural-field hot path (libngp_b200.so).
 *
 * This is the drop-in boundary: every entry point takes plain device pointers, sizes and scalars plus
 * the CUDA stream to launch on, allocates nothing, returns 0 on success and non-zero on failure
 * (message via ngp_last_error()).  It replaces the pybind11 `_backend.<fn>(at::Tensor...)` functions of
 * the reference's torch extensions; each declaration cites the reference interface it replaces
 * (paths relative to the reference root).  The reference launches on the legacy default stream without a
 * device guard; here the caller selects the device (cudaSetDevice / torch.cuda.device) and passes a stream.
 * THE CURRENT DEVICE IS PART OF THE CONTRACT: kernels launch on it, so every buffer must live on it.  Each entry
 * point verifies this for its first device pointer (cudaPointerGetAttributes) and fails with NGP_ERR_ARG instead of
 * launching on the wrong GPU; host pointers are rejected the same way.  ngp_set_device_check(0) removes the check.
 *
 * Conventions: all pointers are device pointers unless named *_host; float = fp32; `table_dtype`
 * 0 = fp32, 1 = fp16.  Row-major, contiguous.  B/N/M are element counts.
 */
#ifndef NGP_B200_H
#define NGP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void *ngp_stream_t; /* cudaStream_t */

#define NGP_OK 0
#define NGP_ERR_ARG 1
#define NGP_ERR_CUDA 2
#define NGP_ERR_UNSUPPORTED 3

#define NGP_DTYPE_F32 0
#define NGP_DTYPE_F16 1

/* thread-local last error string (never NULL) */
const char *ngp_last_error(void);
/* ABI version, bumped on signature changes */
int ngp_abi_version(void);
/* device-contract check described above: on != 0 enables (default; env NGP_B200_CHECK_DEVICE=0 disables at start),
 * on < 0 only queries.  Returns the previous setting. */
int ngp_set_device_check(int on);

/* ---- hash-grid encoder: gridencoder/src/gridencoder.h:12-13, bindings.cpp:5-8 ------------------------- */

/* Per-level scale exp2f(l*S)*H-1 evaluated ON DEVICE exactly as every kernel here evaluates it
 * (reference: gridencoder.cu:126).  scales[L] device buffer.  Debug/test aid for index parity. */
int ngp_grid_level_scales(float *scales, int L, float S, uint32_t H, ngp_stream_t stream);

/* grid_encode_forward (gridencoder.cu:423-446).  inputs [B,D] in [0,1]; embeddings [rows,C]; offsets [L+1] i32.
 * out_layout 0: outputs [L,B,C] (reference layout); 1: outputs [B,L*C] (what grid.py:52 permutes to).
 * outputs / dy_dx have the table dtype.  dy_dx [B,L,D,C] written iff calc_grad_inputs && dy_dx != NULL.
 * indices (optional, may be NULL): u32 [B,L,2^D] level-local row indices, for the bit-exact parity test. */
int ngp_grid_encode_forward(const float *inputs, const void *embeddings, int table_dtype, const int32_t *offsets,
                            void *outputs, int out_layout, int64_t B, int D, int C, int L, float S, uint32_t H,
                            int calc_grad_inputs, void *dy_dx, uint32_t gridtype, int align_corners,
                            uint32_t *indices, ngp_stream_t stream);

/* grid_encode_backward, input part (gridencoder.cu:317-343, 468-476): grad_inputs [B,D] fp32 (w.r.t. the [0,1]
 * inputs).  The reference contracts grad with a stored dy_dx; here the corners are re-gathered instead, so no
 * dy_dx buffer is needed (same arithmetic order).  grad_layout as out_layout above. */
int ngp_grid_encode_backward_input(const void *grad, int grad_layout, const float *inputs, const void *embeddings,
                                   int table_dtype, const int32_t *offsets, float *grad_inputs, int64_t B, int D,
                                   int C, int L, float S, uint32_t H, uint32_t gridtype, int align_corners,
                                   ngp_stream_t stream);

/* grid_encode_backward, table part (gridencoder.cu:227-314, 448-466): grad_embeddings [rows,C] fp32 is
 * OVERWRITTEN (no pre-zeroing needed) with a deterministic sorted segmented reduction (no float atomics).
 * total_rows = embeddings.shape[0] (= offsets[L]; a host scalar, so nothing is read back from the device).
 * workspace: ngp_grid_backward_table_workspace_bytes(B,D,L) bytes of device scratch. */
int64_t ngp_grid_backward_table_workspace_bytes(int64_t B, int D, int L);
int ngp_grid_encode_backward_table(const void *grad, int grad_layout, int grad_dtype, const float *inputs,
                                   const int32_t *offsets, int64_t total_rows, float *grad_embeddings,
                                   int64_t B, int D, int C, int L, float S, uint32_t H, uint32_t gridtype,
                                   int align_corners, void *workspace, int64_t workspace_bytes, ngp_stream_t stream);

/* The same gradient by the reference's own method (float atomics in arrival order, gridencoder.cu:227-314): NOT
 * bit-reproducible, no workspace, skips (point, level) pairs whose upstream gradient is exactly zero.  Opt-in
 * (gridencoder.table_grad.DETERMINISTIC = False); grad_embeddings is overwritten. */
int ngp_grid_encode_backward_table_atomic(const void *grad, int grad_layout, int grad_dtype, const float *inputs,
                                          const int32_t *offsets, int64_t total_rows, float *grad_embeddings,
                                          int64_t B, int D, int C, int L, float S, uint32_t H, uint32_t gridtype,
                                          int align_corners, ngp_stream_t stream);

/* The same gradient (gridencoder.cu:227-314), bit-reproducible WITHOUT a sort: contributions are scaled by one power of two
 * (from max |grad| and the bound 2^D B on contributions per row, so no sum overflows), rounded to int64 and added with
 * 64-bit integer atomics — integer addition is associative, any arrival order gives the same bits — then converted back.
 * Resolution 2^-(61 - ceil(log2(2^D B))) of max |grad| per contribution.  An inf / NaN anywhere in grad makes the whole
 * result NaN (integers cannot carry it; it is what an AMP inf check needs to see).  workspace: 16-byte aligned,
 * ngp_grid_backward_table_fixed_workspace_bytes(total_rows, C) = 256 + 8 total_rows C bytes; grad_embeddings is overwritten. */
int64_t ngp_grid_backward_table_fixed_workspace_bytes(int64_t total_rows, int C);
int ngp_grid_encode_backward_table_fixed(const void *grad, int grad_layout, int grad_dtype, const float *inputs,
                                         const int32_t *offsets, int64_t total_rows, float *grad_embeddings,
                                         int64_t B, int D, int C, int L, float S, uint32_t H, uint32_t gridtype,
                                         int align_corners, void *workspace, int64_t workspace_bytes, ngp_stream_t stream);

/* ---- fused field query: NeRFNetwork.density(), nerf/network.py:126-143 (+ grid.py:140-156, activation.py) --- */

/* sigma = exp(o[0]), geo = o[1:n_out], o = W1 relu(W0 enc((x+bound)/(2 bound))).  fp32 table + fp32 MLP,
 * D=3, C=2, L=16, hidden=64, n_out=16.  W0 [64,32], W1 [16,64] row-major (nn.Linear.weight).
 * geo may be NULL.  relu_mask (optional): u64 [B] saved for ngp_field_query_backward. */
int ngp_field_query_forward(const float *x, float bound, const float *embeddings, const int32_t *offsets, float S,
                            uint32_t H, uint32_t gridtype, int align_corners, const float *W0, const float *W1,
                            float *sigma, float *geo, uint64_t *relu_mask, int64_t B, ngp_stream_t stream);

/* dL/dx [B,3] from dL/dsigma [B] (+ optional dL/dgeo [B,15]); needs sigma and relu_mask from the forward. */
int ngp_field_query_backward(const float *x, float bound, const float *embeddings, const int32_t *offsets, float S,
                             uint32_t H, uint32_t gridtype, int align_corners, const float *W0, const float *W1,
                             const float *sigma, const uint64_t *relu_mask, const float *grad_sigma,
                             const float *grad_geo, float *grad_x, int64_t B, ngp_stream_t stream);

/* Density value AND gradient in one pass (csrc/fieldvg.cu): sigma [B], geo [B,15] (may be NULL) and
 * grad_x [B,3] = u_i * d sigma_i / d x_i with the per-point upstream factor u_i = a_i + b_i * sigma_i
 * (coef_a / coef_b [B] may be NULL -> the scalars a0 / b0).  a0 = 1, b0 = 0 gives the Jacobian rows; coef_b = 2 c_i the
 * planner's collision cost sum_i c_i sigma_i^2 (nav/quad_plot.py:224-250); b0 = 2/B the gradient of mean(sigma^2).
 * Equals ngp_field_query_forward followed by ngp_field_query_backward with grad_sigma = u, in one table pass. */
int ngp_field_query_value_grad(const float *x, float bound, const float *embeddings, const int32_t *offsets, float S,
                               uint32_t H, uint32_t gridtype, int align_corners, const float *W0, const float *W1,
                               const float *coef_a, const float *coef_b, float a0, float b0, float *sigma, float *geo,
                               float *grad_x, int64_t B, ngp_stream_t stream);

/* Density forward that also returns the Jacobian rows jac [B,3] = d sigma_i / d x_i (one table pass, csrc/dfield5.cu): what
 * autograd's NeRFNetwork.density(x) runs when x requires a gradient (the planner's collision cost, nav/quad_plot.py:224-250,
 * differentiates sigma w.r.t. the body-cloud points).  relu_mask (optional, u64 [B]) is what ngp_field_query_backward needs
 * if a gradient also arrives through geo_feat; geo may be NULL. */
int ngp_field_query_forward_jacobian(const float *x, float bound, const float *embeddings, const int32_t *offsets, float S,
                                     uint32_t H, uint32_t gridtype, int align_corners, const float *W0, const float *W1,
                                     float *sigma, float *geo, uint64_t *relu_mask, float *jac, int64_t B,
                                     ngp_stream_t stream);

/* grad_x [B,3] = grad_sigma_i * jac_i: the backward of sigma after ngp_field_query_forward_jacobian (no table pass).
 * grad_x may be jac itself (in place). */
int ngp_field_query_backward_jacobian(const float *jac, const float *grad_sigma, float *grad_x, int64_t B,
                                      ngp_stream_t stream);

/* NeRFNetwork.density of the FFMLP variant, nerf/network_ff.py:76-89: h = sigma_net(encoder(x, bound)) as ONE kernel
 * (csrc/ffmlp5.cu): the hash-grid features are rounded to fp16 (what ffmlp.py:20-24 does to the encoder output) and handed to
 * the FFMLP layers in shared memory; the [B, 32] feature tensor never exists.  fp32 table, D=3, C=2, L=16; weights = the
 * flat fp16 FFMLP vector [64 x 32][64 x 64]*(num_layers-1)[16 x 64]; outputs [B, 16] fp16 (column 0 = log sigma, 1..15 =
 * geo_feat).  Inference only (no saved activations): training goes through ngp_grid_encode_forward + ngp_ffmlp_forward. */
int ngp_field_query_ff_forward(const float *x, float bound, const float *embeddings, const int32_t *offsets, float S,
                               uint32_t H, uint32_t gridtype, int align_corners, const void *weights, int num_layers,
                               int activation, void *outputs, int64_t B, ngp_stream_t stream);

/* ---- fused per-sample field for the ray paths: NeRFNetwork.forward(x, d), nerf/network.py:95-124 ------------------- */

/* Implementation of the fused colour-field / render kernels: 1 = warp-specialised tcgen05 (TMEM accumulators and A
 * operands, 3xTF32; the default), 0 = the mma.sync 3xTF32 kernels kept as an independent fp32 cross-check.  Process-wide;
 * the initial value can also be set with the environment variable NGP_B200_FIELD_IMPL=mma|tc5.  Returns the previous
 * value; impl < 0 only queries.  (No reference counterpart: the reference has one implementation.) */
int ngp_set_field_impl(int impl);

/* sigma [M] = exp(o[0]), rgb [M,3] = sigmoid(colour_net([SH4(d) | o[1:16]])), o = sigma_net(encode(x)); fp32 table and
 * weights of the configuration nav uses: D=3, L=16, C=2, sigma net 32-64-16 (Ws0 [64,32], Ws1 [16,64]), colour net
 * 31-64-64-3 (Wc0 [64,31], Wc1 [64,64], Wc2 [3,64]), all nn.Linear.weight layouts.  masks (optional): u16 [M,3,4] ReLU
 * bits saved for ngp_field_color_backward. */
int ngp_field_color_forward(const float *xyzs, const float *dirs, float bound, const float *embeddings,
                            const int32_t *offsets, float S, uint32_t H, uint32_t gridtype, int align_corners,
                            const float *Ws0, const float *Ws1, const float *Wc0, const float *Wc1, const float *Wc2,
                            float *sigma, float *rgb, void *masks, int64_t M, ngp_stream_t stream);
/* dL/dxyzs [M,3] and dL/ddirs [M,3] (may be NULL) from dL/dsigma [M], dL/drgb [M,3] (frozen field). */
int ngp_field_color_backward(const float *xyzs, const float *dirs, float bound, const float *embeddings,
                             const int32_t *offsets, float S, uint32_t H, uint32_t gridtype, int align_corners,
                             const float *Ws0, const float *Ws1, const float *Wc0, const float *Wc1, const float *Wc2,
                             const float *sigma, const float *rgb, const void *masks, const float *grad_sigma,
                             const float *grad_rgb, float *grad_xyzs, float *grad_dirs, int64_t M, ngp_stream_t stream);

/* ---- ray ops: raymarching/src/raymarching.h:7-18, bindings.cpp:5-19 -------------------------------------------- */

/* near_far_from_aabb (raymarching.cu:93-158).  rays_o/rays_d [N,3], aabb[6] -> nears/fars [N]; miss => FLT_MAX. */
int ngp_near_far_from_aabb(const float *rays_o, const float *rays_d, const float *aabb, int64_t N, float min_near,
                           float *nears, float *fars, ngp_stream_t stream);

/* morton3D / morton3D_invert (raymarching.cu:216-262).  coords [N,3] i32 <-> indices [N] i32. */
int ngp_morton3D(const int32_t *coords, int64_t N, int32_t *indices, ngp_stream_t stream);
int ngp_morton3D_invert(const int32_t *indices, int64_t N, int32_t *coords, ngp_stream_t stream);

/* packbits (raymarching.cu:269-302).  grid [C*H^3] f32 (16-byte aligned), N = C*H^3/8 -> bitfield u8 [N]. */
int ngp_packbits(const float *grid, int64_t N, float thresh, uint8_t *bitfield, ngp_stream_t stream);

/* march_rays_train (raymarching.cu:313-495).  Deterministic: rays[n] = (n, offset, count) with offsets the exclusive
 * scan of counts in ray order starting at counter[0]; counter[0] += total samples, counter[1] += N.  xyzs/dirs [M,3],
 * deltas [M,2] must be pre-zeroed by the caller (raymarching.py:209-211); a ray with offset+count >= M writes nothing.
 * workspace: ngp_march_rays_train_workspace_bytes(N) bytes of device scratch: per-ray counts, scan totals and first-sample
 * parameters (about 8 N bytes) and, for N > 8192 (one thread per ray), room for 16 runs of consecutive samples per ray
 * (132 N bytes more, of which only the runs that exist are touched): the count pass leaves each ray's runs there and the
 * write pass rebuilds the samples from them instead of marching the occupancy grid a second time. */
int64_t ngp_march_rays_train_workspace_bytes(int64_t N);
int ngp_march_rays_train(const float *rays_o, const float *rays_d, const uint8_t *grid, float bound, float dt_gamma,
                         uint32_t max_steps, int64_t N, uint32_t C, uint32_t H, int64_t M, const float *nears,
                         const float *fars, float *xyzs, float *dirs, float *deltas, int32_t *rays, int32_t *counter,
                         uint32_t perturb, void *workspace, int64_t workspace_bytes, ngp_stream_t stream);
/* Write-only second half of the two-call form: after ngp_march_rays_train(..., M = 0, xyzs = dirs = deltas = NULL, ...)
 * counted the samples (counter[0]; `rays` holds (ray, offset, count)), this writes them into buffers of exactly that size
 * without repeating the count pass.  `rays` / `workspace`: those of the count call, untouched in between. */
int ngp_march_rays_train_write(const float *rays_o, const float *rays_d, const uint8_t *grid, float bound, float dt_gamma,
                               uint32_t max_steps, int64_t N, uint32_t C, uint32_t H, int64_t M, const float *nears,
                               const float *fars, float *xyzs, float *dirs, float *deltas, const int32_t *rays,
                               uint32_t perturb, const void *workspace, ngp_stream_t stream);

/* EXTENSION (no reference counterpart; the reference's cuda_ray path is not differentiable w.r.t. the pose): gradient
 * of a loss w.r.t. rays_o / rays_d through the samples of ngp_march_rays_train, with the sample abscissae held constant
 * as the reference's torch sampler does (renderer.py:159-160).  grad_xyzs [M,3], grad_dirs [M,3]
 * (may be NULL) -> grad_rays_o, grad_rays_d [N,3] (overwritten).  nears / deltas / rays / perturb / max_steps as given
 * to / returned by ngp_march_rays_train. */
int ngp_march_rays_train_backward(const float *grad_xyzs, const float *grad_dirs, const float *rays_o,
                                  const float *rays_d, const float *nears, const float *deltas, const int32_t *rays,
                                  float bound, uint32_t max_steps, uint32_t perturb, int64_t N, int64_t M,
                                  float *grad_rays_o, float *grad_rays_d, ngp_stream_t stream);

/* composite_rays_train_forward / _backward (raymarching.cu:505-593, 606-699).  grad_sigmas [M], grad_rgbs [M,3] must be
 * pre-zeroed (raymarching.py:272-273). */
int ngp_composite_rays_train_forward(const float *sigmas, const float *rgbs, const float *deltas, const int32_t *rays,
                                     int64_t M, int64_t N, float *weights_sum, float *depth, float *image,
                                     ngp_stream_t stream);
int ngp_composite_rays_train_backward(const float *grad_weights_sum, const float *grad_image, const float *sigmas,
                                      const float *rgbs, const float *deltas, const int32_t *rays,
                                      const float *weights_sum, const float *image, int64_t M, int64_t N,
                                      float *grad_sigmas, float *grad_rgbs, ngp_stream_t stream);

/* march_rays / composite_rays (raymarching.cu:706-825, 828-922): inference stepping of the alive rays.
 * xyzs/dirs [n_alive*n_step,3], deltas [n_alive*n_step,2] pre-zeroed (delta == 0 marks an unused slot). */
int ngp_march_rays(int64_t n_alive, uint32_t n_step, const int32_t *rays_alive, const float *rays_t,
                   const float *rays_o, const float *rays_d, float bound, float dt_gamma, uint32_t max_steps, uint32_t C,
                   uint32_t H, const uint8_t *grid, const float *nears, const float *fars, float *xyzs, float *dirs,
                   float *deltas, uint32_t perturb, ngp_stream_t stream);
int ngp_composite_rays(int64_t n_alive, uint32_t n_step, int32_t *rays_alive, float *rays_t, const float *sigmas,
                       const float *rgbs, const float *deltas, float *weights_sum, float *depth, float *image,
                       ngp_stream_t stream);

/* ---- pose -> rays closure of the estimator: get_rays (nerf/utils.py:52-116) at the sampled pixels only, for a batch of
 * hypotheses (nav/estimator_helpers.py:311-314 builds all H*W rays of one pose and indexes ~1k of them) ------------------- */

/* poses [P,4,4] cam2world (row-major), pix i64 [R] flat pixel indices (j * W + i) shared by the hypotheses, pinhole
 * (fx, fy, cx, cy) -> rays_o, rays_d [P*R, 3] (hypothesis-major): dir = ((i + .5 - cx) / fx, (j + .5 - cy) / fy, 1) / |.|,
 * rays_d = R dir, rays_o = t.  Backward: grad_poses [P,4,4] (row 3 = 0) from grad_o / grad_d [P*R,3] (either may be NULL),
 * one fixed-order reduction over the R rays of each hypothesis. */
int ngp_pose_rays_forward(const float *poses, const int64_t *pix, int P, int R, int W, float fx, float fy, float cx, float cy,
                          float *rays_o, float *rays_d, ngp_stream_t stream);
int ngp_pose_rays_backward(const float *grad_o, const float *grad_d, const int64_t *pix, int P, int R, int W, float fx, float fy,
                           float cx, float cy, float *grad_poses, ngp_stream_t stream);

/* The estimator's 12-state (position 3, velocity 3, rotation vector 3, angular velocity 3) -> the cam2world pose its
 * measurement renders: nav/estimator_helpers.py:293-327 (measurement_fn) with nav/math_utils.py:159-174 (vec_to_rot_matrix,
 * axis = v / (1e-10 + |v|)) and :26-38 (Blender -> NGP axes): pose = [flip_yz (rot_x(pi/2) R) neg_yz | flip_yz p].
 * states [P,12] -> poses [P,4,4]; the backward is the analytic derivative (dL/dposes [P,4,4] -> dL/dstates [P,12], zero for
 * the velocities; zero for v = 0 as autograd gives).  cos_half_pi / sin_half_pi: cos / sin of the float32 pi / 2 the
 * reference evaluates.  One launch each instead of ~25 / ~40 eager torch ops. */
int ngp_state_to_pose_forward(const float *states, int P, float cos_half_pi, float sin_half_pi, float *poses, ngp_stream_t stream);
int ngp_state_to_pose_backward(const float *states, const float *grad_poses, int P, float cos_half_pi, float sin_half_pi,
                               float *grad_states, ngp_stream_t stream);

/* ---- fused inference renderer: NeRFRenderer.run_cuda eval branch, nerf/renderer.py:325-374 ----------------------- */

/* One persistent kernel for: near_far_from_aabb -> loop { march_rays -> NeRFNetwork.forward(x, d) -> composite_rays ->
 * compact alive rays } -> image + (1 - weights_sum) * bg, clamp(depth - near, 0) / (far - near).  Per-sample data never
 * reach HBM.  rays_o/rays_d [N,3]; aabb [6]; grid = density_bitfield u8 [C*Hgrid^3/8]; field arguments as in
 * ngp_field_color_forward; bg_color NULL (= 1), 3 floats (bg_stride 0) or [N,3] (bg_stride 3).  Outputs image [N,3],
 * depth [N], weights_sum [N] (may be NULL), counts i32 [N] (may be NULL): the number of samples composited for each ray —
 * the integer output of the path (= what march_rays / composite_rays process for the ray before it dies).
 * workspace: ngp_render_rays_infer_workspace_bytes() bytes, 8-byte aligned; after the call its second u64 holds the number
 * of samples evaluated.  dt_gamma/max_steps as in march_rays; a ray takes at most max_steps samples; no perturbation.
 * Implementation: the warp-specialised tcgen05 kernel (csrc/render5.cu) unless ngp_set_field_impl(0) selected mma.sync. */
int64_t ngp_render_rays_infer_workspace_bytes(void);
int ngp_render_rays_infer(const float *rays_o, const float *rays_d, int64_t N, const float *aabb, float min_near,
                          const uint8_t *grid, float bound, float dt_gamma, uint32_t max_steps, uint32_t C, uint32_t Hgrid,
                          const float *embeddings, const int32_t *offsets, float S, uint32_t H, uint32_t gridtype,
                          int align_corners, const float *Ws0, const float *Ws1, const float *Wc0, const float *Wc1,
                          const float *Wc2, float density_scale, const float *bg_color, int bg_stride, float *image,
                          float *depth, float *weights_sum, int32_t *counts, void *workspace, ngp_stream_t stream);

/* The same renderer with the occupancy marching taken OUT of the persistent kernel (csrc/render5.cu, tcgen05 implementation
 * only): marching is a chain of dependent steps that only thread-level parallelism hides, so it runs as a thread-per-ray
 * pre-pass at full occupancy and the persistent kernel only evaluates and composites.  Same samples bit for bit, same
 * outputs and per-ray counts as ngp_render_rays_infer; samples behind a ray's early termination are marched but never
 * evaluated.  Two calls because the sample lists are sized by their total:
 *   ngp_render_plan: plan_counts i32 [N] = samples on each ray (up to max_steps, before any termination), plan_ends i64 [N]
 *     = their inclusive prefix sums (plan_ends[N-1] = total); workspace = ngp_render_plan_workspace_bytes(N) bytes;
 *   ngp_render_rays_infer_planned: arguments of ngp_render_rays_infer + the plan + samples = scratch for 2 * total floats
 *     (8-byte aligned; the (t, dt) lists are written by the call itself) + the renderer's workspace. */
int64_t ngp_render_plan_workspace_bytes(int64_t N);
int ngp_render_plan(const float *rays_o, const float *rays_d, int64_t N, const float *aabb, float min_near,
                    const uint8_t *grid, float bound, float dt_gamma, uint32_t max_steps, uint32_t C, uint32_t Hgrid,
                    int32_t *plan_counts, int64_t *plan_ends, void *workspace, int64_t workspace_bytes, ngp_stream_t stream);
int ngp_render_rays_infer_planned(const float *rays_o, const float *rays_d, int64_t N, const float *aabb, float min_near,
                                  const uint8_t *grid, float bound, float dt_gamma, uint32_t max_steps, uint32_t C,
                                  uint32_t Hgrid, const float *embeddings, const int32_t *offsets, float S, uint32_t H,
                                  uint32_t gridtype, int align_corners, const float *Ws0, const float *Ws1,
                                  const float *Wc0, const float *Wc1, const float *Wc2, float density_scale,
                                  const float *bg_color, int bg_stride, float *image, float *depth, float *weights_sum,
                                  int32_t *counts, const int32_t *plan_counts, const int64_t *plan_ends, float *samples,
                                  int64_t samples_capacity, void *workspace, ngp_stream_t stream);

/* The pre-marched renderer as ONE call without a host read: the pre-pass marches every ray once and records its RUNS of
 * consecutive samples, (first t, number of samples), at most seg_cap per ray — inside a run the marcher's recurrence is
 * t += dt(t), which the persistent kernel's feeder warps replay, so a run is all they need (8 bytes per run instead of per
 * sample, no count pass, no scan, no sizing by the host).  A ray with more than seg_cap runs (their number is known on
 * the device only) is rendered by the marcher-inside kernel over an overflow list in the same call.  Same samples, outputs
 * and per-ray counts as ngp_render_rays_infer.  workspace: ngp_render_rays_infer_segments_workspace_bytes(N, seg_cap)
 * = 512 + 8 N + 8 N seg_cap bytes, 8-byte aligned; after the call its second u64 holds the number of samples evaluated and
 * u64[26] the number of overflow rays.  seg_cap in 1..1024 (32 covers the occupancy grids of trained scenes). */
int64_t ngp_render_rays_infer_segments_workspace_bytes(int64_t N, int seg_cap);
int ngp_render_rays_infer_segments(const float *rays_o, const float *rays_d, int64_t N, const float *aabb, float min_near,
                                   const uint8_t *grid, float bound, float dt_gamma, uint32_t max_steps, uint32_t C,
                                   uint32_t Hgrid, const float *embeddings, const int32_t *offsets, float S, uint32_t H,
                                   uint32_t gridtype, int align_corners, const float *Ws0, const float *Ws1,
                                   const float *Wc0, const float *Wc1, const float *Wc2, float density_scale,
                                   const float *bg_color, int bg_stride, float *image, float *depth, float *weights_sum,
                                   int32_t *counts, int seg_cap, void *workspace, int64_t workspace_bytes,
                                   ngp_stream_t stream);

/* ---- training-step tail: torch.optim.Adam(betas=(0.9, 0.99), eps=1e-15) (main_nerf.py:126) + EMA (nerf/utils.py:814) - */

/* One fused pass over a parameter tensor: g *= grad_scale (all-reduce mean), Adam moments and update with torch's
 * single-tensor arithmetic, optional EMA shadow update (ema may be NULL; ema_decay is the decay of THIS update).  step
 * counts from 1.  torch.amp.GradScaler (nerf/utils.py:784-791): amp_scale (device float, may be NULL) is the loss scale the
 * gradients are divided by; found_inf (device float, may be NULL) non-zero skips the step entirely.  All array pointers
 * 16-byte aligned, n fp32 elements. */
int ngp_adam_step(float *param, const float *grad, float *exp_avg, float *exp_avg_sq, float *ema, int64_t n, float lr,
                  float beta1, float beta2, float eps, int64_t step, float grad_scale, float ema_decay,
                  const float *amp_scale, const float *found_inf, ngp_stream_t stream);

/* Weight gradient of a bias-free Linear of the field MLPs (nn.Linear backward inside Trainer.train_step,
 * nerf/network.py:33-68): dW [n_out, n_in] = dY^T X with dY [M, n_out], X [M, n_in] row-major fp32, n_out, n_in <= 64.
 * fp32 accuracy (3xTF32 split products), deterministic (per-CTA partials summed in CTA order).  dW is overwritten. */
int64_t ngp_linear_wgrad_workspace_bytes(int64_t M);
int ngp_linear_wgrad(const float *dY, const float *X, float *dW, int64_t M, int n_out, int n_in, void *workspace,
                     int64_t workspace_bytes, ngp_stream_t stream);

/* Allocation-free variant of ngp_linear_wgrad.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same
 * inputs, outputs and arithmetic; the per-CTA partials live in a file-scope static __device__ array of the library
 * (4 MiB, held by every GPU that loads the library until the process exits) instead of a caller workspace.
 * Limit: M <= 65536 samples (NGP_LINEAR_WGRAD_STATIC_MAX_M); a larger M returns NGP_ERR_UNSUPPORTED with a message.
 * Calls no cudaMalloc / cudaFree / allocating routine.  All calls on a device share the array (it keeps the last call's
 * contents; nothing needs clearing, every partial read is written by the same call): calls that OVERLAP on the GPU
 * (different streams, concurrent graph replays) race — serialise them.  The first call on a device looks the array's
 * device address up (cudaGetSymbolAddress); make one call before starting a capture. */
#define NGP_LINEAR_WGRAD_STATIC_MAX_M 65536
int ngp_linear_wgrad_static(const float *dY, const float *X, float *dW, int64_t M, int n_out, int n_in, ngp_stream_t stream);

/* Forward / input-gradient product of the same layers (csrc/linear.cu), W = nn.Linear.weight [n_out, n_in] row-major:
 * transpose_w = 0: Y [M, n_out] = X [M, n_in] W^T;  transpose_w = 1: Y [M, n_in] = X [M, n_out] W.  n_out, n_in <= 64,
 * 3xTF32 tensor-core products with fp32 accumulation (relative error ~1e-6 against an fp32 GEMM). */
int ngp_linear_small(const float *X, const float *W, float *Y, int64_t M, int n_out, int n_in, int transpose_w,
                     ngp_stream_t stream);

/* ---- occupancy-grid maintenance: NeRFRenderer.update_extra_state full sweep, nerf/renderer.py:457-486, 519-530 ---- */

/* xyz [G^3, 3]: jittered query position of every cell of `cascade`, in Morton order (so the densities come back in grid
 * order); jitter scales the uniform [-1, 1) half-cell noise (0 = cell centres), seed selects the noise. */
int ngp_occ_grid_positions(int cascade, uint32_t G, float bound, float jitter, uint32_t seed, float *xyz,
                           ngp_stream_t stream);
/* grid[i] = max(grid[i] * decay, sigma[i] * density_scale) where grid[i] >= 0; block_sums [ngp_occ_grid_update_blocks(n)]
 * receives partial sums of max(grid, 0) (their total / n is the mean density of renderer.py:521). */
int64_t ngp_occ_grid_update_blocks(int64_t n);
int ngp_occ_grid_update(float *grid, const float *sigma, int64_t n, float decay, float density_scale, float *block_sums,
                        ngp_stream_t stream);

/* Partial refresh of update_extra_state (renderer.py:488-517), one cascade: N uniformly random cells + N cells drawn with
 * replacement from the occupied ones (grid_cas [G^3] > 0; ordered, reproducible compaction) -> cell_idx i32 [2N] (Morton
 * index; -1 where nothing is occupied) and their jittered query positions xyz [2N,3].  workspace:
 * ngp_occ_grid_partial_workspace_bytes(G) bytes.  Then, with sigma [2N] = density(xyz):
 * ngp_occ_grid_scatter_update: tmp [n] (scratch) = -1, tmp[cell] = max of the queries that hit it (x density_scale),
 * grid = max(grid * decay, tmp) where both >= 0 (renderer.py:519-520), block_sums as in ngp_occ_grid_update. */
int64_t ngp_occ_grid_partial_workspace_bytes(uint32_t G);
int ngp_occ_grid_partial_select(const float *grid_cas, int cascade, uint32_t G, float bound, float jitter, uint32_t seed,
                                uint32_t N, void *workspace, int32_t *cell_idx, float *xyz, ngp_stream_t stream);
int ngp_occ_grid_scatter_update(float *grid_cas, const int32_t *cell_idx, const float *sigma, uint32_t nq, int64_t n,
                                float decay, float density_scale, float *tmp, float *block_sums, ngp_stream_t stream);
/* NeRFRenderer.mark_untrained_grid (renderer.py:381-444): grid [C, G^3] cells that none of the B cam2world poses [B,4,4]
 * (intrinsics fx, fy, cx, cy) sees are set to -1. */
int ngp_occ_grid_mark_untrained(const float *poses, int B, float fx, float fy, float cx, float cy, uint32_t C, uint32_t G,
                                float bound, float *grid, ngp_stream_t stream);
/* The planner's A* occupancy (nav/quad_plot.py:64-83): xyz [side^3, 3] = meshgrid('ij') of torch.linspace(lo, hi, side)
 * per axis (lo_host / hi_host: 3 HOST floats each); out u8 [(side/k)^3] = MaxPool3d(k)(sigma [side^3]) > thresh. */
int ngp_lattice_positions(const float *lo_host, const float *hi_host, uint32_t side, float *xyz, ngp_stream_t stream);
int ngp_maxpool_threshold(const float *sigma, uint32_t side, uint32_t k, float thresh, uint8_t *out, ngp_stream_t stream);

/* ---- spherical harmonics: shencoder/src/shencoder.h:10,13 (degree <= 4) ----------------------------------------- */

/* sh_encode_forward (shencoder.cu:386-412).  inputs [B,3] -> outputs [B,degree^2]; dy_dx [B,3,degree^2] optional. */
int ngp_sh_encode_forward(const float *inputs, float *outputs, int64_t B, int D, int degree, int calc_grad_inputs,
                          float *dy_dx, ngp_stream_t stream);
/* sh_encode_backward (shencoder.cu:414-440): grad_inputs [B,3] overwritten; Jacobian recomputed from inputs. */
int ngp_sh_encode_backward(const float *grad, const float *inputs, int64_t B, int D, int degree, float *grad_inputs,
                           ngp_stream_t stream);

/* ---- fully fused MLP: ffmlp/src/ffmlp.h:8-14, bindings.cpp:5-11 ---------------------------------------------------- */

/* ffmlp_forward / ffmlp_inference (ffmlp.cu:629-700).  inputs [B,input_dim] f16; weights flat f16
 * [64 x input_dim][64 x 64]*(num_layers-1)[16 x 64], each block row-major [out,in]; outputs [B,16] f16.
 * forward_buffer [num_layers,B,64] f16 receives the post-activation hidden layers (training); NULL = inference
 * (the reference's inference_buffer scratch is not needed).  hidden_dim must be 64, output_dim 16 (the padded width),
 * input_dim a multiple of 16 (<= 256); activation 0 = ReLU | 6 = None, output_activation 6 = None (ffmlp.py:107-108).
 * B needs no padding to a multiple of 128. */
int ngp_ffmlp_forward(const void *inputs, const void *weights, int64_t B, int input_dim, int output_dim,
                      int hidden_dim, int num_layers, int activation, int output_activation, void *forward_buffer,
                      void *outputs, ngp_stream_t stream);

/* ffmlp_backward (ffmlp.cu:754-895).  grad [B,16] f16 (zero in the padded columns); forward_buffer from
 * ngp_ffmlp_forward.  Writes grad_weights (flat f16, same layout as weights; may be NULL), grad_inputs [B,input_dim] f16
 * iff calc_grad_inputs (any input_dim; the reference needs CUTLASS unless input_dim == 64), and, when not NULL,
 * backward_buffer [num_layers,B,64] f16 (entry k = gradient at hidden layer num_layers-1-k, as the reference).
 * Nothing needs pre-zeroing.  Weight gradients are summed per CTA in TMEM (fp32) and reduced in CTA order
 * (bit-reproducible).  workspace: ngp_ffmlp_backward_workspace_bytes(B, input_dim, num_layers) bytes. */
int64_t ngp_ffmlp_backward_workspace_bytes(int64_t B, int input_dim, int num_layers);
int ngp_ffmlp_backward(const void *grad, const void *inputs, const void *weights, const void *forward_buffer,
                       int64_t B, int input_dim, int output_dim, int hidden_dim, int num_layers, int activation,
                       int output_activation, int calc_grad_inputs, void *backward_buffer, void *grad_inputs,
                       void *grad_weights, void *workspace, int64_t workspace_bytes, ngp_stream_t stream);

/* ---- measurement aid (not on the product path) ----------------------------------------------------------------- */

/* L2-gather roofline probe: blocks x 256 threads each issue iters x 8 independent pseudo-random
 * loads of elem_bytes (4, 8 or 16) from buf[0:buf_bytes].  pattern 0 = independent rows, 1 = the 8 hashed corners of a
 * random cell.  Useful bytes = blocks*256*iters*8*elem_bytes; the caller times it with CUDA events. */
int ngp_probe_gather(const void *buf, int64_t buf_bytes, int elem_bytes, int iters, int pattern, int blocks,
                     uint32_t *sink, ngp_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* NGP_B200_H */
