/*
 * apb200.h -- C-ABI of libapb200.so: hand-written sm_100a CUDA kernels for the PointNeXt
 * set-abstraction hot path of AdaptPoint/OpenPoints.
 *
 * Drop-in boundary: each entry point replaces one function of the reference's pybind module
 * `pointnet2_batch_cuda` (openpoints/cpp/pointnet2_batch/src/pointnet2_api.cpp:10-24); the
 * reference wrapper it stands in for is cited per function.  Differences from the reference ABI,
 * all deliberate:
 *   - plain device pointers + int sizes instead of at::Tensor;
 *   - an explicit cudaStream_t (passed as void*) instead of the legacy default stream;
 *   - an int status is returned (0 = ok, otherwise a cudaError_t or APB_E_*); nothing calls
 *     exit() (the reference launchers fprintf + exit(-1), e.g. sampling_gpu.cu:46-50);
 *   - the current CUDA device of the calling thread is used (caller sets it, as torch does).
 * Ownership is unchanged: the caller allocates every buffer, pre-zeroes gradient accumulators and
 * the ball-query index buffer exactly as the reference Python does, the callee keeps no state.
 * Layouts are the reference's: coordinates (B,N,3) point-major f32, features (B,C,N)
 * channel-major f32, indices int32, all dense/contiguous.
 */
#ifndef APB200_H_
#define APB200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void *apb_stream_t; /* a cudaStream_t */

#define APB_OK 0
#define APB_E_INVALID 10001     /* bad argument (null pointer, negative size, unsupported shape) */
#define APB_E_UNSUPPORTED 10002 /* shape outside what the kernel family covers */
#define APB_E_SCRATCH_LIMIT 10003 /* problem needs more than the library's built-in scratch array (no-allocation entry points) */

/* Library / build identification: "apb200 <version> sm_100a". */
const char *apb_version(void);
/* Human-readable text for a status returned by any apb_* call. */
const char *apb_error_string(int status);

/* The reference's FPS thread-block size rule, cuda_utils.h:10-14 (opt_n_threads). It fixes the
 * tie-break order of apb_fps_f32. */
int apb_fps_block_size(int n);

/* furthest_point_sampling_wrapper(b,n,m,points,temp,idx)   -- sampling.cpp:39-48,
 * kernel sampling_gpu.cu:100-216.
 *   xyz  (B,N,3) f32;  idx (B,m) int32 out;
 *   temp (B,N) f32 or NULL: if given it is the initial running min-distance (the reference Python
 *        fills 1e10, subsample.py:94) and receives the final one, as in the reference; if NULL
 *        1e10 is assumed and nothing is written back.
 * Indices are bit-identical to the reference kernel, including its tie rule (smallest
 * bit-reversed (k mod block_size), then smallest k). */
int apb_fps_f32(const float *xyz, float *temp, int32_t *idx, int B, int N, int m, apb_stream_t stream);

/* Benchmarks only: override the forward grouping/gather work split (channels staged per CTA in
 * {0,1,2,4}, threads per CTA in {0,256,512,1024}; 0 = automatic; channels -1 = rows longer than shared
 * memory use the direct L2 gather instead of the cluster kernel).  No reference counterpart. */
int apb_tune_wgrad(int force_mt, int min_chunks);
int apb_tune_group(int channels_per_cta, int threads);
/* Benchmarks only: backward path of grouping / three_interpolate: 0 = automatic, 1 = scattered L2
 * reductions (the reference's scheme), 2 = accumulator rows in shared memory.  No reference counterpart. */
int apb_tune_group_grad(int mode);
/* Benchmarks only: slice size in KB (16..200, 0 = automatic) of the sliced grouping kernels used when a row is
 * longer than one SM's shared memory.  No reference counterpart. */
int apb_tune_group_seg(int kb);

/* gather_points_wrapper(b,c,n,npoints,points,idx,out)      -- sampling.cpp:11-22 */
int apb_gather_f32(const float *points, const int32_t *idx, float *out, int B, int C, int N, int m,
                   apb_stream_t stream);
/* gather_points_grad_wrapper(b,c,n,npoints,grad_out,idx,grad_points) -- sampling.cpp:25-36;
 * grad_points (B,C,N) must be pre-zeroed (accumulated into). */
int apb_gather_grad_f32(const float *grad_out, const int32_t *idx, float *grad_points, int B, int C,
                        int N, int m, apb_stream_t stream);

/* ball_query_wrapper(b,n,m,radius,nsample,new_xyz,xyz,idx) -- ball_query.cpp:29-39,
 * kernel ball_query_gpu.cu:15-51.  idx (B,M,K) is fully written (queries without a hit get
 * zeros, which is what the reference's pre-zeroed buffer holds). */
int apb_ball_query_f32(const float *new_xyz, const float *xyz, int32_t *idx, int B, int N, int M,
                       float radius, int nsample, apb_stream_t stream);
/* The same query through a uniform cell grid built per call (csrc/query_grid.cu): O(M * neighbourhood) instead of
 * O(M * N) distance evaluations, bit-identical rows (the nsample smallest in-radius indices == the first nsample in
 * index order).  workspace: apb_ball_query_grid_ws_bytes(B, N) bytes of device memory, 256-byte aligned, contents
 * irrelevant; nsample <= 96. */
long long apb_ball_query_grid_ws_bytes(int B, int N);
int apb_ball_query_grid_f32(const float *new_xyz, const float *xyz, int32_t *idx, void *workspace, long long ws_bytes,
                            int B, int N, int M, float radius, int nsample, apb_stream_t stream);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  apb_ball_query_grid_f32 with
 * the workspace in a 16 MiB array built into the library (file-scope __device__ array in csrc/query_grid.cu, held by every
 * GPU that loads the library until exit; apb_ball_query_grid_static_ws_bytes() returns its size): allocates nothing, takes
 * only inputs and outputs, same results bit for bit.  Returns APB_E_SCRATCH_LIMIT when apb_ball_query_grid_ws_bytes(B, N)
 * exceeds the array (about B * (262 KB + 20 N bytes): e.g. B = 8, N = 24000 or B = 32, N = 8192 fit).  The device current
 * at the call must be the one that owns the buffers and the stream.  All calls on a device share the array: calls that
 * overlap on the GPU (streams with no ordering between them) race -- serialise them. */
long long apb_ball_query_grid_static_ws_bytes(void);
int apb_ball_query_grid_static_f32(const float *new_xyz, const float *xyz, int32_t *idx, int B, int N, int M,
                                   float radius, int nsample, apb_stream_t stream);

/* kNN over squared distances, order (distance, index) ascending.  Replaces KNN.forward =
 * torch.cdist + topk in group.py:12-28 (no native function in the reference).
 * idx (B,M,K) int32; dist2 (B,M,K) f32 or NULL. K <= 128. */
int apb_knn_f32(const float *new_xyz, const float *xyz, int32_t *idx, float *dist2, int B, int N, int M,
                int K, apb_stream_t stream);

/* group_points_wrapper(b,c,n,npoints,nsample,points,idx,out) -- group_points.cpp:25-35 */
int apb_group_f32(const float *points, const int32_t *idx, float *out, int B, int C, int N, int M, int K,
                  apb_stream_t stream);
/* Grouping fused with the max and min over the K neighbours (no reference kernel; replaces
 * grouping_operation -> normalise -> torch.max of PointsetGrouper / Producefactor, generator_component4_15.py:368-431,
 * 560-563):  vmax/vmin (B,C,M) f32, amax/amin (B,C,M) int32 = the point index in [0,N) that attains them (first
 * neighbour slot on ties).  APB_E_UNSUPPORTED when one feature row (4N bytes) exceeds shared memory. */
int apb_group_minmax_f32(const float *points, const int32_t *idx, float *vmax, float *vmin, int32_t *amax,
                         int32_t *amin, int B, int C, int N, int M, int K, apb_stream_t stream);
/* group_points_grad_wrapper(b,c,n,npoints,nsample,grad_out,idx,grad_points) -- group_points.cpp:11-22;
 * grad_points (B,C,N) pre-zeroed. */
int apb_group_grad_f32(const float *grad_out, const int32_t *idx, float *grad_points, int B, int C, int N,
                       int M, int K, apb_stream_t stream);

/* three_nn_wrapper(b,n,m,unknown,known,dist2,idx)          -- interpolate.cpp:17-27 */
int apb_three_nn_f32(const float *unknown, const float *known, float *dist2, int32_t *idx, int B, int n,
                     int m, apb_stream_t stream);
/* three_interpolate_wrapper(b,c,m,n,points,idx,weight,out) -- interpolate.cpp:30-42 */
int apb_three_interp_f32(const float *points, const int32_t *idx, const float *weight, float *out, int B,
                         int C, int m, int n, apb_stream_t stream);
/* three_interpolate_grad_wrapper(b,c,n,m,grad_out,idx,weight,grad_points) -- interpolate.cpp:45-57;
 * grad_points (B,C,m) pre-zeroed. */
int apb_three_interp_grad_f32(const float *grad_out, const int32_t *idx, const float *weight,
                              float *grad_points, int B, int C, int n, int m, apb_stream_t stream);

/* Fused AdaptPoint / PointWOLF deformation (no native function in the reference: replaces the torch op chain of
 * AdaptPoint_Augmentor.forward, generator_component4_15.py:148-173 + 204-234 + 294 + 313-327, and
 * PointWOLF_classversion.__call__, online_aug/pointwolf.py:27-178).
 *   xyz (B,N,3); anchor (B,M,3) = xyz[fps idx]; affine (B,M,12) = row-major A_m = R_m S_m (9) then trl_m (3);
 *   proj_axis (B,3) in {0,1}; mask (B,N) or NULL; sigma > 0; M <= 8.
 *   out (B,N,3) = mask * s * (q - mean q),  q = kernel regression of the M local transforms (see deform.cu);
 *   stats (B,8) or NULL: mean xyz, s, max norm r*, argmax n* (int bits) -- consumed by the backward. */
int apb_deform_mask_fwd_f32(const float *xyz, const float *anchor, const float *affine, const float *proj_axis,
                            const float *mask, float sigma, float *out, float *stats, int B, int N, int M,
                            apb_stream_t stream);
/* Backward of the above w.r.t. affine (grad_affine (B,M,12), fully written) and mask (grad_mask (B,N) or NULL).
 * xyz / anchor are data in the reference's loop (train_autoaug.py:136-148) and get no gradient. */
int apb_deform_mask_bwd_f32(const float *xyz, const float *anchor, const float *affine, const float *proj_axis,
                            const float *mask, float sigma, const float *stats, const float *grad_out,
                            float *grad_affine, float *grad_mask, int B, int N, int M, apb_stream_t stream);

/* 1x1 convolution (the "shared MLP" layer: Conv1d/Conv2d kernel 1, openpoints/models/layers/conv.py:8-21, :24-103)
 * with folded BatchNorm / bias and optional ReLU, as a bf16 tcgen05 GEMM with fp32 accumulation in TMEM:
 *   y[b,:,p] = act(scale * (W . x[b,:,p]) + shift)
 *   x (B,Cin,P) f32 channel-major;  w_bf16 (Cout,Kp) bf16 bits, K-major, columns >= Cin zero, Kp % 64 == 0;
 *   scale/shift (Cout) f32 or both NULL;  y (B,Cout,P) f32.  Cout <= 512.  No native counterpart in the
 * reference (it calls cuDNN through nn.Conv2d); tolerance vs fp32: 2e-2 relative (bf16 inputs). */
int apb_pwconv_fwd_bf16(const float *x, const uint16_t *w_bf16, const float *scale, const float *shift, int relu,
                        float *y, int B, int Cin, int Cout, int P, int Kp, apb_stream_t stream);

/* Same layer with TF32 operands (fp32 inputs read as tf32 by the TMA unit, fp32 accumulate):
 *   w (Cout,Kp) f32 K-major, columns >= Cin zero, Kp % 32 == 0; any Cout.  This is the training-path precision: it
 * equals the reference's default cuDNN behaviour on Ampere+ (torch.backends.cudnn.allow_tf32 = True).
 * Both operands go global -> shared memory by TMA (csrc/tc_cm_gemm.cu: the channel-major activation is the MN-major
 * operand of tcgen05.mma, no thread touches it); needs P % 4 == 0 and 16-byte aligned x, w -- other shapes take the
 * register-staged kernel apb_pwconv_fwd_tf32_rows (Cout <= 512). */
int apb_pwconv_fwd_tf32(const float *x, const float *w, const float *scale, const float *shift, int relu, float *y,
                        int B, int Cin, int Cout, int P, int Kp, apb_stream_t stream);
int apb_pwconv_fwd_tf32_rows(const float *x, const float *w, const float *scale, const float *shift, int relu, float *y,
                             int B, int Cin, int Cout, int P, int Kp, apb_stream_t stream);
/* The TMA-fed kernel with its epilogue variants (APB_E_UNSUPPORTED unless P % 4 == 0 and x, w 16-byte aligned):
 *   point_major != 0 : y is (B*P, Cout) row-major (the per-support-point operand U of the fused set abstraction);
 *   stats != NULL    : (Cout,2) f64, += sum and sum of squares over (b,p) of scale*(W.x)+shift BEFORE the ReLU -- the
 *                      batch statistics of the BatchNorm that follows the conv (openpoints/models/layers/norm.py:57-71),
 *                      so no separate statistics pass reads y again;
 *   pmax != NULL     : (B,Cout) u64 pre-zeroed; y is not written; pmax[b,c] = max over p of scale*(W.x)+shift packed
 *                      as (order-preserving value bits << 32 | 0xFFFFFFFF - first arg-max p): the conv -> ReLU -> max
 *                      over the points head of the discriminator (point_discriminator.py:166-191) without the
 *                      (B,Cout,P) activation.  apb_pmax_decode unpacks n entries into value / arg. */
int apb_pwconv_ex_tf32(const float *x, const float *w, const float *scale, const float *shift, int relu, float *y,
                       int point_major, double *stats, unsigned long long *pmax, int B, int Cin, int Cout, int P,
                       int Kp, apb_stream_t stream);
int apb_pmax_decode(const unsigned long long *packed, float *val, int32_t *arg, long long n, apb_stream_t stream);

/* Weight gradient of the 1x1 convolution above (TF32 tcgen05, both operands TMA-fed):
 *   grad_w[co,ci] += sum_{b,p} grad_y[b,co,p] * x[b,ci,p];  grad_w (Cout,Cin) f32 pre-zeroed by the caller.
 *   grad_y (B,Cout,P), x (B,Cin,P) f32 channel-major, P % 4 == 0 (else APB_E_UNSUPPORTED).
 * The data gradient is apb_pwconv_fwd_bf16 on grad_y with the transposed weights. */
int apb_pwconv_wgrad_tf32(const float *grad_y, const float *x, float *grad_w, int B, int Cin, int Cout, int P,
                          apb_stream_t stream);
/* Same with a row pitch ldw >= Cin for grad_w (Cout, ldw): with ldw % 4 == 0 the split-K partial sums are added with
 * 128-bit reductions (4 columns per instruction) instead of scalar atomics; columns Cin .. ldw-1 receive zeros. */
int apb_pwconv_wgrad_ld_tf32(const float *grad_y, const float *x, float *grad_w, int ldw, int B, int Cin, int Cout,
                             int P, apb_stream_t stream);

/* Fused set-abstraction aggregation for BatchNorm-folded blocks: group -> relative position (/radius) ->
 * shared MLP (1 or 2 layers) -> max over the K neighbours, without materialising any (B,C,M,K) tensor.
 * Replaces QueryAndGroup.forward + get_aggregation_feautres('dp_fj') + convs + pool of SetAbstraction.forward
 * (pointnext.py:164-166) and LocalAggregation.forward (pointnext.py:66-77).
 *   xyz (B,N,3), new_xyz (B,M,3), feat (B,Cin,N) f32, idx (B,M,K) int32 (from apb_ball_query_f32 / apb_knn_f32);
 *   first conv W1 (C1, 3+Cin) split as  w1p (3,C1) f32 = W1[:, :3]^T / radius  and  w1f_bf16 (C1,K1p) = W1[:, 3:];
 *   s1,t1 (C1) folded BN1 (+bias); relu1;
 *   second conv (NULL for one-layer blocks) w2_bf16 (C2,K2p), s2,t2 (C2), relu2;  K must be 32 when present;
 *   u_ws: caller-provided workspace of B*N*C1 floats;  out (B,Cout,M) f32 with Cout = C2 (or C1). */
int apb_sa_fused_fwd_bf16(const float *xyz, const float *new_xyz, const float *feat, const int32_t *idx,
                          const uint16_t *w1f_bf16, const float *w1p, const float *s1, const float *t1, int relu1,
                          const uint16_t *w2_bf16, const float *s2, const float *t2, int relu2, float *u_ws,
                          float *out, int B, int N, int M, int K, int Cin, int C1, int K1p, int C2, int K2p,
                          apb_stream_t stream);

/* ---- BatchNorm (batch statistics) (+ReLU) and BatchNorm + max over K, channel-major (B,C,P) f32 -------------------
 * No native counterpart in the reference: it runs nn.BatchNorm{1,2}d (openpoints/models/layers/norm.py:57-71) and
 * torch.max over the neighbour dim (pointnext.py:134) through torch/cuDNN.  stats / stats2 are (C,2) fp64
 * accumulators, zeroed by the caller (and all-reduced across ranks by the caller for SyncBatchNorm).
 *   apb_bn_stats_f32       stats[c]  += (sum x, sum x^2)
 *   apb_bn_apply_f32       y = act(gamma * (x - mean) * rstd + beta); writes save_mean/save_rstd (C); updates
 *                          running_mean/var (may be NULL) with `momentum` and the unbiased variance; count = elements
 *                          per channel over all ranks
 *   apb_bn_bwd_reduce_f32  stats2[c] += (sum dy', sum dy' * xhat), dy' = dy * [y_relu > 0] if y_relu given
 *   apb_bn_bwd_apply_f32   dx = gamma * rstd * (dy' - mean(dy') - xhat * mean(dy' xhat))
 *   apb_kmax_stats_f32     y (B,C,M,K): vmax/vmin (B,C,M), arg = argmax | argmin << 16 (first index on ties),
 *                          stats[c] += (sum, sum of squares) (stats may be NULL)
 *   apb_bnmax_bwd_f32      dy[b,c,m,k] = gsel[b,c,m] * [k == arg[b,c,m]] + a[c] + bcoef[c] * y[b,c,m,k]
 *                          (a / bcoef NULL = 0; y may be NULL when bcoef is) */
/* out = act(scale[c] * x + shift[c]) * [mask_relu > 0]: eval-mode BatchNorm (+ReLU) forward (mask NULL) and its
 * data gradient (x := dy, shift NULL, relu 0, mask := the forward output when ReLU was applied). */
int apb_affine_f32(const float *x, const float *mask_relu, const float *scale, const float *shift, int relu, float *out,
                   int B, int C, long long P, apb_stream_t stream);
int apb_bn_stats_f32(const float *x, double *stats, int B, int C, long long P, apb_stream_t stream);
int apb_bn_apply_f32(const float *x, const double *stats, const float *gamma, const float *beta, float eps,
                     double count, int relu, float *y, float *save_mean, float *save_rstd, float *running_mean,
                     float *running_var, float momentum, int B, int C, long long P, apb_stream_t stream);
int apb_bn_bwd_reduce_f32(const float *x, const float *dy, const float *y_relu, const float *save_mean,
                          const float *save_rstd, double *stats2, int B, int C, long long P, apb_stream_t stream);
int apb_bn_bwd_apply_f32(const float *x, const float *dy, const float *y_relu, const float *save_mean,
                         const float *save_rstd, const float *gamma, const double *stats2, double count, float *dx,
                         int B, int C, long long P, apb_stream_t stream);
int apb_kmax_stats_f32(const float *y, float *vmax, float *vmin, int32_t *arg, double *stats, int B, int C, int M,
                       int K, apb_stream_t stream);
/* vmax[r] = max_p y[r,p], arg[r] = first p attaining it, y (rows,P) row-major.  Replaces the ReLU + torch.max over
 * the points of the discriminator's group_all stage (point_discriminator.py:166-191): max(relu(y)) = relu(max(y)). */
int apb_rowmax_f32(const float *y, float *vmax, int32_t *arg, long long rows, int P, apb_stream_t stream);
int apb_bnmax_bwd_f32(const float *y, const float *gsel, const int32_t *arg, const float *a, const float *bcoef,
                      float *dy, int B, int C, int M, int K, apb_stream_t stream);

/* grouping_operation on bf16 features (bit-exact copy: out[b,c,j,s] = features[b,c,idx[b,j,s]]), 16-byte stores of
 * eight bf16 per channel; features (B,C,N), out (B,C,M,K) are bf16 bit patterns.  The bf16 counterpart of
 * group_points_wrapper (pointnet2_api.cpp:13) for the bf16 shared-MLP path. */
int apb_group_bf16(const uint16_t *features, const int32_t *idx, uint16_t *out, int B, int C, int N, int M, int K,
                   apb_stream_t stream);

/* Spectral normalisation with one power iteration (torch.nn.utils.parametrizations.spectral_norm in training mode, as
 * used by PointDiscriminator1, openpoints/models_adaptpoint/point_discriminator.py:23-34): w_mat (h,w) row-major,
 * u (h), v (w) updated in place; t_ws (h) workspace; scratch (4+w) f32 pre-zeroed; u_saved, v_saved, sigma (1), w_sn (h,w)
 * outputs: w_sn = w_mat / sigma with sigma = u^T w_mat v.  3 launches.  Backward (u, v constant, as in torch):
 * grad_w = (grad - <grad, w_sn> u v^T) / sigma, scratch (4) f32 pre-zeroed, 2 launches. */
int apb_spectral_norm_fwd_f32(const float *w_mat, float *u, float *v, float *t_ws, float *scratch, float *u_saved,
                              float *v_saved, float *sigma, float *w_sn, float eps, int h, int w, apb_stream_t stream);
int apb_spectral_norm_bwd_f32(const float *grad, const float *w_sn, const float *u_saved, const float *v_saved,
                              const float *sigma, float *scratch, float *grad_w, int h, int w, apb_stream_t stream);

/* Eval-mode BatchNorm fold (scale = gamma*rsqrt(var+eps), shift = beta - mean*scale, rstd) and the split of a (C,2)
 * f64 sum pair into two f32 vectors: one launch each instead of the 4-5 element-wise launches torch would issue per
 * BatchNorm layer (openpoints/models/layers/norm.py:57-71 in eval mode / its parameter gradients). */
int apb_bn_fold_f32(const float *gamma, const float *beta, const float *running_mean, const float *running_var, float eps,
                    float *scale, float *shift, float *rstd, int C, apb_stream_t stream);
int apb_sums_split_f32(const double *sums, float *first, float *second, int C, apb_stream_t stream);
/* Training-mode BatchNorm whose normalised tensor is never written: (C,2) f64 sums over `count` elements -> mean, rstd,
 * scale = gamma*rstd, shift = beta - mean*scale, and the running-statistics update (NULL, NULL: skip)
 * (torch.nn.BatchNorm2d forward in training mode, openpoints/models/layers/norm.py:57-71). */
int apb_bn_finalize_f32(const double *stats, double count, float eps, float momentum, const float *gamma, const float *beta,
                        float *running_mean, float *running_var, float *scale, float *shift, float *mean, float *rstd, int C,
                        apb_stream_t stream);

/* The per-channel arithmetic around BatchNorm -> max over K, in three kernels instead of ~40 element-wise launches:
 *   apb_bnmax_finalize_f32  : (stats (C,2) f64 over `count` elements | NULL = running statistics) -> mean, rstd,
 *                             scale = gamma*rstd, running-statistics update (training), then per (b,c,m):
 *                             sel = scale >= 0 ? vmax : vmin, ksel = its k (arg = kmax | kmin << 16),
 *                             out = act(scale*sel + beta - mean*scale).
 *   apb_bnmax_bwd_reduce_f32: gs = g * [out_relu > 0] * scale; sums (C,2) f64 += [sum g*mask, sum g*mask*(sel-mean)*rstd].
 *   apb_bnmax_bwd_coef_f32  : dbeta, dgamma from the LOCAL sums; a, bcoef (dense BatchNorm-backward terms
 *                             dy += a + bcoef*y) from the sums over all ranks (NULL, NULL in eval mode). */
int apb_bnmax_finalize_f32(const double *stats, double count, float eps, float momentum, const float *gamma,
                           const float *beta, float *running_mean, float *running_var, const float *vmax,
                           const float *vmin, const int32_t *arg, int relu, float *out, float *sel, int32_t *ksel,
                           float *scale, float *mean, float *rstd, int B, int C, int M, apb_stream_t stream);
int apb_bnmax_bwd_reduce_f32(const float *g, const float *out_relu, const float *sel, const float *scale,
                             const float *mean, const float *rstd, float *gs, double *sums, int B, int C, int M,
                             apb_stream_t stream);
int apb_bnmax_bwd_coef_f32(const double *tot, const double *local, const float *scale, const float *mean,
                           const float *rstd, double count, float *a, float *bcoef, float *dgamma, float *dbeta, int C,
                           apb_stream_t stream);

/* Training-mode set abstraction for one-conv blocks without the (B,C,M,K) tensors, forward and backward
 * (csrc/sa_train.cu; replaces the chain QueryAndGroup -> cat -> Conv2d -> BatchNorm2d -> ReLU -> max over K of
 * openpoints/models/backbone/pointnext.py:66-77, :164-166 for LocalAggregation and sa_layers = 1 SetAbstraction).
 * u_pm (B,N,C) = Wf . f per support point, point-major (apb_pwconv_ex_tf32 with point_major = 1); wp (C,3) = xyz part of
 * the conv weight (already divided by the radius when normalize_dp);  y[b,m,k,c] = u_pm[b,idx,c] + wp[c,:].(p[idx]-q[m]).
 *   apb_group_ystats_f32 : vmax/vmin (B,C,M) = max/min over k of y, arg (B,C,M) = kmax | kmin << 16, stats (C,2) f64 +=
 *                          sum y, sum y^2 (NULL: skip).  vmax = vmin = arg = NULL: statistics only.  C % 4 == 0.
 *   apb_sa1_bwd_count_f32: cnt (B,N) f32 += occurrences of n in idx[b]; sdp (B,N,3) += sum of (p[n]-q[m]) over them;
 *                          mom (12) f64 += [sum dp (3), sum dp dp^T (9)].  All pre-zeroed by the caller.
 *   apb_sa1_bwd_sparse_f32: for every (b,c,m): du_cm[b,c,idx[b,m,ksel]] += gs; dwp[c,:] += gs*dp (f64, NULL: skip);
 *                          dxyz[b,n,:] += gs*wp[c,:], dq[b,m,:] -= gs*wp[c,:] (NULL: no coordinate gradient).
 *   apb_sa1_bwd_dense_f32: du_cm[b,c,n] += cnt*a[c] + bc[c]*(cnt*u_pm[b,n,c] + wp[c,:].sdp[b,n]);
 *                          usdp (C,3) f64 += sum_n u_pm[b,n,c]*sdp[b,n,:]   (the dense BatchNorm-backward terms). */
int apb_group_ystats_f32(const float *u_pm, const float *xyz, const float *new_xyz, const int32_t *idx, const float *wp,
                         float *vmax, float *vmin, int32_t *arg, double *stats, int B, int N, int M, int K, int C,
                         apb_stream_t stream);
int apb_sa1_bwd_count_f32(const float *xyz, const float *new_xyz, const int32_t *idx, float *cnt, float *sdp, double *mom,
                          int B, int N, int M, int K, apb_stream_t stream);
int apb_sa1_bwd_sparse_f32(const float *gs, const int32_t *ksel, const int32_t *idx, const float *xyz,
                           const float *new_xyz, const float *wp, float *du_cm, double *dwp, float *dxyz, float *dq, int B,
                           int N, int M, int K, int C, apb_stream_t stream);
int apb_sa1_bwd_dense_f32(const float *u_pm, const float *cnt, const float *sdp, const float *a, const float *bc,
                          const float *wp, float *du_cm, double *usdp, int B, int N, int C, apb_stream_t stream);
/* Batch statistics of y (sum y, sum y^2 per channel, added to stats (C,2) f64) in closed form from cnt / sdp / mom of
 * apb_sa1_bwd_count_f32: one pass over the B*N*C values of u_pm instead of the B*M*K*C gathered ones (the statistics
 * of the first BatchNorm2d of a two-conv block, openpoints/models/layers/conv.py:24-60 in training mode). */
int apb_sa_ustats_f32(const float *u_pm, const float *cnt, const float *sdp, const float *wp, const double *mom,
                      double *stats, int B, int N, int C, apb_stream_t stream);

/* Training-mode set abstraction for TWO-conv blocks (PointNeXt-S SetAbstraction, sa_layers = 2) without the (B,C,M,K)
 * tensors, forward and backward (csrc/sa_train2.cu, tcgen05 / TMEM / TMA; replaces QueryAndGroup -> cat -> Conv2d ->
 * BatchNorm2d -> ReLU -> Conv2d -> BatchNorm2d [-> ReLU] -> max over K of pointnext.py:164-166, layers/conv.py:24-60).
 * K == 32, C1 % 4 == 0, C2 % 4 == 0, C1, C2 <= 512 (backward rows pass and wgrad: C1 <= 256).  u_pm, wp as above; (s1, t1) (C1) = scale / shift of
 * BatchNorm 1;  h1[r,:] = relu1(s1*(u_pm[idx_r,:] + wp.dp_r) + t1),  y2[r,:] = W2 . h1[r,:].
 *   apb_sa2_fwd_tf32     : w2 (C2, K2p) fp32, K2p = C1 rounded up to 32, zero padded.  vmax/vmin (B,C2,M) = max/min over k
 *                          of y2, arg = kmax | kmin << 16, stats (C2,2) f64 += sum y2, sum y2^2 (NULL: skip).
 *   apb_sa2_bwd_rows_tf32: gs_pm (B*M, C2) f32 = gradient reaching the selected y2 value, ksel_pm (B*M, C2) u8 = its k;
 *                          wb (C1, Kp) = [ W2^T | G ], W2^T zero-padded to kc_sparse*32 columns, G = W2^T diag(b2) W2 (C1,C1)
 *                          zero-padded to C1 rounded up to 32 (eval mode: Kp = kc_sparse*32, no G block, vvec = NULL);
 *                          vvec (C1) = W2^T a2.  dz1 = (wb . [onehot(gs) | h1] + vvec) * [z1 > 0];  du_pm (B,N,C1) pre-zeroed
 *                          += s1*dz1 at idx;  sums1 (C1,2) f64 += sum dz1, sum dz1*(y1-mean1)*rstd1;  sdzdp (C1,3) f64 +=
 *                          sum dz1*dp;  sh1 (C1) f64 += sum h1  (sums1 NULL: none of the three);  dxyz (B,N,3) += s1*dz1 . wp,
 *                          dq (B,M,3) -= the same (NULL: no coordinate gradient).
 *   apb_sa2_wgrad_tf32   : d_out (C2 [+ C1], C1 rounded up to 32) f32 pre-zeroed: rows [0,C2) += sum_r onehot(gs)[r]^T h1[r],
 *                          with_gram: rows [C2, C2+C1) += sum_r h1[r] h1[r]^T. */
int apb_tune_sa2(int flags); /* benchmarks only (results are then wrong): bits 1 / 2 / 4 / 8 / 16 / 32 switch off the dU REDs / epilogue U loads / channel sums / producer U loads / producer h1 arithmetic / epilogue row metadata of the backward rows kernel */
int apb_sa2_fwd_tf32(const float *u_pm, const float *xyz, const float *new_xyz, const int32_t *idx, const float *wp,
                     const float *s1, const float *t1, int relu1, const float *w2, float *vmax, float *vmin, int32_t *arg,
                     double *stats, int B, int N, int M, int K, int C1, int C2, int K2p, apb_stream_t stream);
int apb_sa2_bwd_rows_tf32(const float *u_pm, const float *xyz, const float *new_xyz, const int32_t *idx, const float *wp,
                          const float *s1, const float *t1, int relu1, const float *gs_pm, const uint8_t *ksel_pm,
                          const float *wb, int Kp, int kc_sparse, const float *vvec, const float *mean1, const float *rstd1,
                          float *du_pm, double *sums1, double *sdzdp, double *sh1, float *dxyz, float *dq, int B, int N,
                          int M, int K, int C1, int C2, apb_stream_t stream);
int apb_sa2_wgrad_tf32(const float *u_pm, const float *xyz, const float *new_xyz, const int32_t *idx, const float *wp,
                       const float *s1, const float *t1, int relu1, const float *gs_pm, const uint8_t *ksel_pm, float *d_out,
                       int with_gram, int B, int N, int M, int K, int C1, int C2, apb_stream_t stream);

/* Multi-head self-attention of the imitator (generator_component4_15.py:434-480, Anchor_selfattention.forward:
 * attn = softmax(q k^T / sqrt(d)); out = attn v), head_dim = 16 only (APB_E_UNSUPPORTED otherwise), fp32, scores never
 * materialised.  q,k,v,o,dout,dq,dk,dv: (B, M, H*16) token-major, 16-byte aligned; lse, delta: (B,H,M) f32 (lse =
 * log2-domain log-sum-exp written by the forward pass; delta = caller-allocated workspace of the backward pass). */
int apb_attn_fwd_f32(const float *q, const float *k, const float *v, float *o, float *lse, int B, int M, int H,
                     int head_dim, apb_stream_t stream);
int apb_attn_bwd_f32(const float *q, const float *k, const float *v, const float *o, const float *dout,
                     const float *lse, float *dq, float *dk, float *dv, float *delta, int B, int M, int H,
                     int head_dim, apb_stream_t stream);
/* Same attention for FEW tokens and any head width (the imitator's anchor head: 4 anchors, head_dim 256,
 * generator_component4_15.py:534-585): M <= 32 tokens, p = (B,H,M,M) soft-max probabilities kept for the backward. */
int apb_attn_small_fwd_f32(const float *q, const float *k, const float *v, float *o, float *p, int B, int M, int H,
                           int head_dim, apb_stream_t stream);
int apb_attn_small_bwd_f32(const float *q, const float *k, const float *v, const float *p, const float *dout,
                           float *dq, float *dk, float *dv, int B, int M, int H, int head_dim, apb_stream_t stream);

/* ---- small all-reduce over NVLink peer memory (SyncBN statistics) -------------------------------------------
 * Replaces the per-layer collective of torch SyncBatchNorm the reference switches on for world_size > 1
 * (examples/classification/main.py:27, openpoints/models/layers/norm.py:57-71).  Every rank creates one workspace,
 * the 64-byte CUDA IPC handles are exchanged by the host (torch.distributed), every rank opens its peers' handles,
 * and apb_peer_allreduce then sums buf[0..n) over the ranks in ONE kernel (P2P stores + flags; CUDA-graph
 * replayable; n <= apb_peer_max_elems()).  dtype: 0 = f32, 1 = f64.  ws[r] = rank r's workspace as mapped in this
 * process (ws[rank] = the pointer apb_peer_ws_create returned). */
int apb_peer_ws_bytes(void);
int apb_peer_max_elems(void);
int apb_peer_ws_create(void **ws_out, unsigned char *handle_out64);
int apb_peer_ws_open(const unsigned char *handle64, void **ws_out);
int apb_peer_ws_close(void *ws, int own);
int apb_peer_allreduce(void *buf, int n, int dtype, void *const *ws, int world, int rank, apb_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* APB200_H_ */
