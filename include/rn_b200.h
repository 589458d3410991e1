/*
 * rn_b200.h — C ABI of librn_b200.so: hand-written sm_100a CUDA kernels for the synchronous
 * data-parallel ResNet training step of michaelwfc/distributed-tensorflow-resnet.
 *
 * The reference has NO plugin / FFI interface for this path (it is pure TF-1.x Python; every op
 * below runs inside un-vendored TensorFlow).  Each entry point therefore cites the reference call
 * site whose TensorFlow op it replaces (file:line relative to the reference root).
 *
 * Conventions
 *   - extern "C", plain pointers and sizes; no C++ / torch types.
 *   - every compute entry returns 0 on success or a negative rn_status; rn_last_error() gives a
 *     thread-local message.  No exceptions cross the boundary.
 *   - the CALLER owns every buffer (device pointers); kernels are stateless and stream-ordered on
 *     `stream` (a cudaStream_t passed as void*; NULL = legacy default stream).  Scratch memory is
 *     passed in with its size; query it with the matching *_scratch_bytes().
 *   - activations are NHWC; conv kernels HWIO [k,k,Cin,Cout] exactly as TF stores them.
 *   - `dtype` selects the storage type of activations: RN_F32 (verification mode) or RN_BF16
 *     (bf16 in HBM, fp32 arithmetic/accumulation).  Parameters, statistics, gradients of
 *     parameters and optimizer state are always fp32.
 */
#ifndef RN_B200_H_
#define RN_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum { RN_F32 = 0, RN_BF16 = 1 } rn_dtype;

typedef enum {
  RN_OK = 0,
  RN_ERR_INVALID = -1,     /* bad argument (shape, alignment, dtype, NULL pointer)            */
  RN_ERR_SCRATCH = -2,     /* scratch buffer too small                                        */
  RN_ERR_CUDA = -3,        /* a CUDA runtime / driver call failed; see rn_last_error()        */
  RN_ERR_UNSUPPORTED = -4  /* valid request that this build does not implement                */
} rn_status;

const char* rn_last_error(void);
/* ABI version; bumped whenever a signature changes.  The Python bindings (_lib.py) refuse a library
 * whose rn_abi_version() differs from the version they were written against. */
#define RN_ABI_VERSION 5
int rn_abi_version(void);
/* Number of kernel launches issued through this library by the calling process (all threads). */
uint64_t rn_launch_count(void);

/* ------------------------------------------------------------------------------------------
 * Convolution (Conv2D / Conv2DBackpropInput / Conv2DBackpropFilter)
 *   replaces tf.layers.conv2d at resnet_model_official.py:87-91 (+ fixed_padding :53-77 folded
 *   into `pad`) and its autodiff ops created by optimizer.minimize, models/resnet_model.py:125.
 *   y[n,ho,wo,co] = sum_{r,s,ci} x[n, ho*stride - pad + r, wo*stride - pad + s, ci] * w[r,s,ci,co]
 *   with zero fill outside the image.  `pad` is the number of zero rows/cols BEFORE the image:
 *   (k-1)/2 for the official model (SAME at stride 1; symmetric fixed_padding + VALID at stride 2)
 *   and TF-SAME's pad_before for the upstream `_conv`.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  int32_t N, H, W, Cin;     /* input  [N,H,W,Cin]   */
  int32_t Ho, Wo, Cout;     /* output [N,Ho,Wo,Cout] */
  int32_t k, stride, pad;
  int32_t dtype;            /* rn_dtype of x / y / dy / dx */
} rn_conv_desc;

/* CUDA-core (FFMA) implicit-GEMM path: the fp32 verification mode, and the fallback for shapes
 * the tensor-core path does not take.  In RN_BF16 mode `w` is a bf16 HWIO copy of the weights; in
 * RN_F32 mode it is the fp32 master. */
int rn_conv2d_fprop(const rn_conv_desc* d, const void* x, const void* w, void* y, void* stream);
/* dx = beta*dx + dgrad(dy, w); beta is 0 or 1 (gradient fan-in at residual forks, AddN). */
int rn_conv2d_dgrad(const rn_conv_desc* d, const void* dy, const void* w, void* dx, float beta, void* stream);
/* dw (fp32 HWIO) = wgrad(x, dy).  Deterministic split-K: partials in `scratch`, fixed-order sum. */
size_t rn_conv2d_wgrad_scratch_bytes(const rn_conv_desc* d);
int rn_conv2d_wgrad(const rn_conv_desc* d, const void* x, const void* dy, float* dw,
                    void* scratch, size_t scratch_bytes, void* stream);

/* tcgen05 / TMA / TMEM implicit-GEMM path (bf16 operands, fp32 accumulation in tensor memory).
 * A plan owns the CUtensorMap descriptors (which embed the buffer addresses) for one conv layer
 * and one of its three GEMMs; it is created once at graph-build time and run every step.
 *   kind 0: fprop  a = x  [N,H,W,Cin] bf16,    b = w bf16 HWIO [k*k][Cin][Cout], out = y  bf16
 *   kind 1: dgrad  a = dy [N,Ho,Wo,Cout] bf16, b = w bf16 HWIO [k*k][Cin][Cout], out = dx bf16
 *           (the same bf16 copy of the HWIO kernel serves both: an MN-major UMMA operand for fprop,
 *            a K-major one for dgrad — no transposed copy is ever made)
 *   kind 2: wgrad  a = x, b = dy, out = dw fp32 HWIO (split-K partials in caller scratch)
 * rn_convtc_supported() says whether the shape is handled (else use the CUDA-core path). */
typedef struct rn_convtc_plan rn_convtc_plan;
int rn_convtc_supported(const rn_conv_desc* d, int kind);
size_t rn_convtc_scratch_bytes(const rn_conv_desc* d, int kind);
int rn_convtc_plan_create(const rn_conv_desc* d, int kind, const void* a, const void* b, void* out,
                          void* scratch, size_t scratch_bytes, rn_convtc_plan** plan);
/* beta (dgrad only): 0 overwrite, 1 accumulate into `out`. */
/* fprop with a fused epilogue (Cout a multiple of 16, store groups of 64 / 32 / 16 channels; beta must be 0 at run time):
 *   residual != NULL      : out = bf16(bf16(conv) + residual) — the bottleneck's `inputs + shortcut`
 *                           (resnet_model_official.py:130 / :175) without materialising the conv output;
 *                           residual has the shape and layout of out.
 *   stat_partials != NULL : per-channel (sum, sum of squares) of the STORED bf16 output, one row
 *                           [Cout][2] fp32 per CTA; rn_convtc_stat_rows() bounds the rows to allocate,
 *                           *stat_rows receives the number this plan writes; rn_bn_finalize turns them
 *                           into the batch-norm statistics (FusedBatchNorm, :41-50) without a pass over
 *                           the tensor. */
int rn_convtc_fused_supported(const rn_conv_desc* d);
int rn_convtc_stat_rows(const rn_conv_desc* d);
int rn_convtc_plan_create_fused(const rn_conv_desc* d, const void* x, const void* w_bf16, void* out,
                                const void* residual, float* stat_partials, int32_t* stat_rows,
                                rn_convtc_plan** plan);
/* dgrad whose output dx is the incoming gradient of a batch_norm_relu (resnet_model_official.py:41-50) that is
 * read by this convolution only (stride 1, Cin a multiple of 16, <= 1024; beta must be 0 at run time): the
 * epilogue also accumulates that BN's backward reductions — sum(g) and sum(g * xhat) with g = dx masked by the
 * ReLU, xhat from bn_x / mean / rstd — one row [Cin][2] per CTA in sum_partials (rn_convtc_stat_rows bounds the
 * rows); rn_bn_bwd_finalize turns them into dbeta / dgamma, replacing rn_bn_act_bwd_reduce's pass over dy and x. */
int rn_convtc_fused_dgrad_supported(const rn_conv_desc* d);
int rn_convtc_plan_create_fused_dgrad(const rn_conv_desc* d, const void* dy, const void* w_bf16, void* dx,
                                      const void* bn_x, const float* gamma, const float* beta, const float* mean,
                                      const float* rstd, float leak, float* sum_partials, int32_t* rows,
                                      rn_convtc_plan** plan);
/* Fold the finalisation of a fused epilogue's partial rows into the convolution launch: the CTA that finishes last
 * (atomic ticket) sums the rows in fp64 in a fixed order and writes what rn_bn_finalize (plan from
 * rn_convtc_plan_create_fused / _bnin with stat_partials) or rn_bn_bwd_finalize (plan from
 * rn_convtc_plan_create_fused_dgrad) would have written, so that launch disappears from the step
 * (FusedBatchNorm / FusedBatchNormGrad, resnet_model_official.py:41-50).  One CTA sums the table, which pays while
 * the BN has few channels (the executor folds up to 256); `ticket` is one zero-initialised int32 in device memory
 * owned by the caller and private to the plan.  Call after plan creation, before rn_convtc_run. */
int rn_convtc_plan_fold_bn_finalize(rn_convtc_plan* plan, int64_t M, float eps, float decay, int32_t unbiased,
                                    float* mean, float* rstd, float* moving_mean, float* moving_var, int32_t* ticket);
int rn_convtc_plan_fold_bn_bwd_finalize(rn_convtc_plan* plan, float* dgamma, float* dbeta, int32_t* ticket);
/* 1x1 convolutions (any stride, Cin a multiple of 64) whose input is the output of a training-mode
 * batch_norm_relu (resnet_model_official.py:41-50 followed by conv2d_fixed_padding :80-91 — conv1 / conv3 of a
 * bottleneck block and the projection shortcut): the plan reads the BN's INPUT `bn_x` and applies
 * act(gamma * (x - mean) * rstd + beta), rounded to bf16 exactly as rn_bn_apply_act_fwd stores it, to every operand
 * tile in shared memory before the tensor cores read it.  The normalised tensor is never written to or read from
 * HBM: the rn_bn_apply_act_fwd pass of that BN disappears.  kind 0 (fprop; b = bf16 HWIO weights; optional fused
 * residual / statistics epilogue as rn_convtc_plan_create_fused) or kind 2 (wgrad; b = dy, scratch as
 * rn_convtc_plan_create).  mean / rstd are read at RUN time (they are produced by the BN's statistics kernel
 * earlier in the same stream). */
int rn_convtc_bnin_supported(const rn_conv_desc* d, int kind);
int rn_convtc_plan_create_bnin(const rn_conv_desc* d, int kind, const void* bn_x, const float* gamma, const float* beta,
                               const float* mean, const float* rstd, float leak, const void* b, void* out,
                               const void* residual, float* stat_partials, int32_t* stat_rows, void* scratch,
                               size_t scratch_bytes, rn_convtc_plan** plan);
int rn_convtc_run(rn_convtc_plan* plan, float beta, void* stream);
int rn_convtc_plan_destroy(rn_convtc_plan* plan);
/* One line of text describing how a plan will run (tile shape, pipeline depth, CTA pairs, fusions): for logs and tests. */
int rn_convtc_plan_describe(const rn_convtc_plan* plan, char* buf, size_t n);
/* Re-layout of the fp32 HWIO master weights into the two bf16 operand layouts above
 * ([k*k][Cout][Cin] for fprop, [k*k][Cin][Cout] for dgrad). Either output may be NULL. */
int rn_weights_to_bf16(const float* w_hwio, int k, int Cin, int Cout, void* w_fprop, void* w_dgrad, void* stream);

/* The same re-layouts for every conv kernel of a net in ONE launch.  `jobs_device` is a device array
 * of njobs records { const float* w; bf16* w_fprop; bf16* w_dgrad; int32 taps, Cin, Cout, kpad; }
 * (rn_relayout_job_bytes() each); kpad > 0 selects the stem layout of rn_stem_weights. */
size_t rn_relayout_job_bytes(void);
int rn_weights_to_bf16_batched(const void* jobs_device, int32_t njobs, void* stream);

/* Stem convolution (the 3-channel image, resnet_model_official.py:250 / :310): TMA rows and UMMA K
 * steps need >= 16 bytes of channels, so the image is unfolded into bf16 rows of
 * Kpad = roundup(k*k*Cin, 64) columns ordered (r, s, ci) and the tensor-core kernels run the layer as
 * a 1x1 convolution over Kpad channels (rn_convtc_* with k=1).  rn_stem_weights builds the matching
 * bf16 [Cout][Kpad] operand from the fp32 HWIO master. */
int rn_im2col(const void* x, void* cols, int32_t N, int32_t H, int32_t W, int32_t Cin, int32_t Ho, int32_t Wo,
              int32_t k, int32_t stride, int32_t pad, int32_t Kpad, int32_t src_dtype, void* stream);
int rn_stem_weights(const float* w_hwio, int32_t rows, int32_t Cout, int32_t Kpad, void* w_fprop, void* stream);

/* Stride-2 stem without the unfolded matrix (the ImageNet 7x7/2 `conv2d_fixed_padding` over the image,
 * resnet_model_official.py:310; Cin <= 4, k = 7 or 8).  The padded image is packed 2x2 -> channels once per
 * step:  U[n, h2, w2, (dy*2+dx)*Cin + c] = x[n, 2*h2+dy-pad, 2*w2+dx-pad, c]  (bf16, 16 channels per pixel,
 * dims from rn_stem_s2d_geometry: [N][H2][W2][16]).  A TMA view of U with a 32-byte W stride and 128-byte rows
 * presents 4 neighbouring packed pixels as one 64-channel row, so the layer runs on the tensor-core kernels as
 * `taps` (= 4) vertical taps over 64 channels:
 *   kind 0: u, b = packed weights bf16 [taps][64][Cout] (rn_stem_s2d_weights from the bf16 HWIO kernel), out = y
 *   kind 2: u, b = dy, out = packed gradient fp32 [taps][64][Cout]; rn_stem_s2d_wgrad_unpack scatters it back
 *           to fp32 HWIO.  Plans run with rn_convtc_run and are freed with rn_convtc_plan_destroy. */
int rn_stem_s2d_supported(const rn_conv_desc* d);
int rn_stem_s2d_geometry(const rn_conv_desc* d, int32_t* H2, int32_t* W2, int32_t* taps);
int rn_stem_s2d_pack(const rn_conv_desc* d, const void* x, int32_t src_dtype, void* u, void* stream);
int rn_stem_s2d_weights(const rn_conv_desc* d, const void* w_hwio_bf16, void* w_packed, void* stream);
int rn_stem_s2d_wgrad_unpack(const rn_conv_desc* d, const float* dw_packed, float* dw_hwio, void* stream);
size_t rn_stem_s2d_scratch_bytes(const rn_conv_desc* d, int kind);
int rn_stem_s2d_plan_create(const rn_conv_desc* d, int kind, const void* u, const void* b, void* out,
                            void* scratch, size_t scratch_bytes, rn_convtc_plan** plan);

/* ------------------------------------------------------------------------------------------
 * Batch normalisation + (leaky) ReLU (+ residual add)
 *   replaces tf.layers.batch_normalization(fused=True) + tf.nn.relu, resnet_model_official.py:41-50
 *   (FusedBatchNorm epsilon 1.001e-5, decay 0.997, moving variance from the UNBIASED batch
 *   variance), the residual `inputs + shortcut` at :130 / :175, and FusedBatchNormGrad + ReluGrad
 *   + AddN from autodiff.  Also serves the upstream `_batch_norm` / `_relu(leakiness)` (biased
 *   moving variance, eps 1e-3, decay 0.9).  x viewed as [M = N*H*W, C]; C must be a multiple of 4 (f32)
 *   or 8 (bf16) — every access is a 128-bit vector, other C return RN_ERR_UNSUPPORTED — and the
 *   activation pointers (x, addend, sum_out, y, dy, dx, dx_addend) must be 16-byte aligned
 *   (RN_ERR_INVALID otherwise).
 * ------------------------------------------------------------------------------------------ */
size_t rn_bn_scratch_bytes(int64_t M, int32_t C);
/* Statistics from per-CTA partial sums partials[nparts][C][2] written by a fused convolution epilogue
 * (rn_convtc_plan_create_fused): same outputs and moving-average update as rn_bn_stats_fwd. */
int rn_bn_finalize(const float* partials, int32_t nparts, int64_t M, int32_t C, float eps, float decay,
                   int32_t unbiased, float* mean, float* rstd, float* moving_mean, float* moving_var, void* stream);
int rn_bn_bwd_finalize(const float* partials, int32_t nparts, int32_t C, float* dgamma, float* dbeta, void* stream);
/* The apply passes with that finalisation folded in: every CTA sums the partial rows for its own <= 64 channels
 * (fp64, one fixed order, so all CTAs derive identical constants) and one CTA per channel chunk writes
 * mean / rstd / moving averages (forward) or dgamma / dbeta (backward) for later use.  Removes the
 * rn_bn_finalize / rn_bn_bwd_finalize launch between the convolution and the apply pass; outputs equal those
 * of rn_bn_finalize + rn_bn_apply_act_fwd, resp. rn_bn_bwd_finalize + rn_bn_act_bwd_apply. */
int rn_bn_apply_act_fwd_partials(const void* x, void* y, int64_t M, int32_t C, int32_t dtype, const float* gamma,
                                 const float* beta, const float* partials, int32_t nparts, float eps, float decay,
                                 int32_t unbiased, float* mean, float* rstd, float* moving_mean, float* moving_var,
                                 float leak, void* stream);
int rn_bn_act_bwd_apply_partials(const void* dy, const void* x, void* dx, const void* dx_addend, int64_t M, int32_t C,
                                 int32_t dtype, const float* gamma, const float* beta, const float* mean,
                                 const float* rstd, const float* partials, int32_t nparts, float* dgamma, float* dbeta,
                                 float leak, void* stream);
/* Training statistics.  If `addend` != NULL the statistics are of z = x + addend and z is written
 * to `sum_out` (the residual add fused into the statistics pass).  Writes mean[C], rstd[C]
 * (= 1/sqrt(biased_var + eps)) and, if moving_mean != NULL, updates
 * mov -= (mov - batch) * (1 - decay) with the unbiased (unbiased != 0) or biased variance. */
int rn_bn_stats_fwd(const void* x, const void* addend, void* sum_out, int64_t M, int32_t C, int32_t dtype,
                    float eps, float decay, int32_t unbiased, float* mean, float* rstd,
                    float* moving_mean, float* moving_var, void* scratch, size_t scratch_bytes, void* stream);
/* Same statistics pass for hosts that cannot supply scratch memory — no counterpart in the
 * reference; it exists for embedders, e.g. a plugin living inside a TensorFlow-1.x process whose
 * allocator has already taken (almost) all device memory so that cudaMalloc fails, or a loop that
 * must never hit cudaMalloc's device-wide synchronisation.  The partial sums and ticket counters
 * live in a 512 KiB device array that is part of the library image (held by every GPU that loads the
 * library until the process exits).  The entry point performs no allocation of any kind.
 * Its first call on a device looks the array's address up (cudaGetSymbolAddress, that device
 * current): make that call during set-up, before starting any stream capture.  On every call the
 * device that owns `stream` must be the current device (the array used is the current device's).  All calls on a
 * device share the array; counters are re-zeroed at the start of each call, but calls that overlap
 * on the GPU race with each other — keep them ordered.  C is limited to 4096 channels (any M);
 * beyond that the call fails with RN_ERR_INVALID and rn_bn_stats_fwd must be used.
 * Arguments and results are those of rn_bn_stats_fwd without the scratch pair. */
int rn_bn_stats_fwd_noalloc(const void* x, const void* addend, void* sum_out, int64_t M, int32_t C, int32_t dtype,
                            float eps, float decay, int32_t unbiased, float* mean, float* rstd,
                            float* moving_mean, float* moving_var, void* stream);
/* y = act(gamma * (x - mean) * rstd + beta), act(v) = v < 0 ? leak * v : v.
 * If stats_are_moving != 0, `rstd` holds a VARIANCE and rstd = 1/sqrt(var + eps) is formed on the
 * fly (inference mode: mean/var are the moving statistics). */
int rn_bn_apply_act_fwd(const void* x, void* y, int64_t M, int32_t C, int32_t dtype, const float* gamma,
                        const float* beta, const float* mean, const float* rstd, int32_t stats_are_moving,
                        float eps, float leak, void* stream);
/* g = dy * act'(pre);  dbeta = sum g;  dgamma = sum g * xhat.  (pre is recomputed from x.) */
int rn_bn_act_bwd_reduce(const void* dy, const void* x, int64_t M, int32_t C, int32_t dtype,
                         const float* gamma, const float* beta, const float* mean, const float* rstd, float leak,
                         float* dgamma, float* dbeta, void* scratch, size_t scratch_bytes, void* stream);
/* dx = gamma * rstd * (g - dbeta/M - xhat * dgamma/M) [+ dx_addend].  dx may alias dx_addend. */
int rn_bn_act_bwd_apply(const void* dy, const void* x, void* dx, const void* dx_addend, int64_t M, int32_t C,
                        int32_t dtype, const float* gamma, const float* beta, const float* mean, const float* rstd,
                        const float* dgamma, const float* dbeta, float leak, void* stream);

/* ------------------------------------------------------------------------------------------
 * Pooling
 *   max: tf.layers.max_pooling2d(3, 2, 'SAME'), resnet_model_official.py:314-316 (+ MaxPoolGrad)
 *   avg: tf.layers.average_pooling2d(8|7, 1, 'VALID') :269-271 / :337-339 (+ AvgPoolGrad); also the
 *        upstream `_residual` shortcut avg_pool(stride, stride, 'VALID') + channel zero-pad and
 *        `_global_avg_pool`.
 * ------------------------------------------------------------------------------------------ */
int rn_maxpool3x3s2_fwd(const void* x, void* y, uint8_t* argmax, int32_t N, int32_t H, int32_t W, int32_t C,
                        int32_t Ho, int32_t Wo, int32_t pad, int32_t dtype, void* stream);
int rn_maxpool3x3s2_bwd(const void* dy, const uint8_t* argmax, void* dx, int32_t N, int32_t H, int32_t W, int32_t C,
                        int32_t Ho, int32_t Wo, int32_t pad, int32_t dtype, float beta, void* stream);
/* VALID k x k stride s average; output channel c_before + c (Cout >= C + c_before, rest zero). */
int rn_avgpool_fwd(const void* x, void* y, int32_t N, int32_t H, int32_t W, int32_t C, int32_t k, int32_t s,
                   int32_t Cout, int32_t c_before, int32_t in_dtype, int32_t out_dtype, void* stream);
int rn_avgpool_bwd(const void* dy, void* dx, int32_t N, int32_t H, int32_t W, int32_t C, int32_t k, int32_t s,
                   int32_t Cout, int32_t c_before, int32_t in_dtype, int32_t out_dtype, float beta, void* stream);

/* ------------------------------------------------------------------------------------------
 * Classifier head and loss
 *   dense:  tf.layers.dense, resnet_model_official.py:274 / :343 (MatMul + BiasAdd and their grads)
 *   xent:   tf.nn.softmax + tf.losses.softmax_cross_entropy, models/resnet_model.py:79-83
 *           (Sum / num_present reduction), d logits = (softmax * sum(y) - y) / B; plus the
 *           trainers' precision metric mean(argmax(pred) == argmax(label)),
 *           resnet_cifar_train.py:271-273.  All fp32.
 * ------------------------------------------------------------------------------------------ */
/* `scratch` (optional, rn_dense_scratch_bytes) enables a deterministic split-K pass for long reductions. */
size_t rn_dense_scratch_bytes(int32_t B, int32_t D, int32_t K);
int rn_dense_fwd(const float* x, const float* w, const float* b, float* y, int32_t B, int32_t D, int32_t K,
                 void* scratch, size_t scratch_bytes, void* stream);
/* dw / db / dx may each be NULL (skipped).  `scratch` (optional, rn_dense_bwd_scratch_bytes) enables the
 * deterministic split-K pass for dx, whose reduction runs over the K classes. */
size_t rn_dense_bwd_scratch_bytes(int32_t B, int32_t D, int32_t K);
int rn_dense_bwd(const float* x, const float* w, const float* dy, float* dw, float* db, float* dx,
                 int32_t B, int32_t D, int32_t K, void* scratch, size_t scratch_bytes, void* stream);
/* out_scalars[0] = mean xent, out_scalars[1] = precision. probs / dlogits may be NULL.  B <= 4096.
 * Per-row results pass through a static device buffer: calls must not run CONCURRENTLY on two streams of
 * one device (same-stream or event-ordered calls are fine). */
int rn_softmax_xent(const float* logits, const float* onehot, int32_t B, int32_t K, float* probs, float* dlogits,
                    float* out_scalars, void* stream);

/* ------------------------------------------------------------------------------------------
 * Gradient aggregation epilogue + optimizer over the flat fp32 buffers
 *   replaces SyncReplicasOptimizer's ConditionalAccumulator.take_grad(R) mean
 *   (models/resnet_model.py:104-116) / hvd.DistributedOptimizer's average (:118-120), the L2
 *   weight-decay gradient (:88-89; AddN(wgrad, wd*v)) and ApplyMomentum(use_nesterov=False) /
 *   ApplyGradientDescent (:98-101), in ONE pass:
 *     g' = g * inv_R + wd * mask * w ;  m = momentum * m + g' ;  w -= lr * m   (sgd: w -= lr * g')
 *   `g` holds the SUM over R replicas (ncclAllReduce sum).  lr is read from device memory
 *   (hyper[0]) so the step can live in a CUDA graph while the host LR hook
 *   (resnet_cifar_train.py:291-311) changes it.  granule_mask (nullable) has one byte per 64
 *   parameters: 1 = weight decay applies (all ones for the reference; 'DW' only upstream).
 *   If w_bf16 != NULL a bf16 copy of the updated weights is written (same offsets).
 *   If sumsq_out != NULL, sumsq_out[0] = sum(mask * w_old^2) (for cost = xent + wd * sumsq / 2),
 *   reduced deterministically through `scratch`.
 * ------------------------------------------------------------------------------------------ */
size_t rn_optimizer_scratch_bytes(int64_t n);
int rn_sgd_momentum_fused(float* w, const float* g, float* m, void* w_bf16, int64_t n, const uint8_t* granule_mask,
                          const float* hyper, float inv_R, float wd, float momentum, int32_t use_momentum,
                          float* sumsq_out, void* scratch, size_t scratch_bytes, void* stream);
/* out[0] = sum(mask * w^2) without updating anything (eval / reporting). */
int rn_l2_sumsq(const float* w, int64_t n, const uint8_t* granule_mask, float* out, void* scratch,
                size_t scratch_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * CIFAR binary-record decode + augmentation
 *   replaces cifar_input.build_input (cifar_input.py:25-119; pad 2 -> 36x36) and
 *   parse_record / preprocess_image (resnet_cifar_train.py:158-202; pad 4 -> 40x40):
 *   label byte -> one-hot; CHW uint8 -> HWC float; zero-pad; random crop 32x32; random
 *   left-right flip; per-image standardisation (x - mean) / max(stddev, 1/sqrt(3072)).
 *   records: uint8 [n_records][record_bytes] resident in HBM; indices[B] selects the batch.
 *   Random draws: Philox4x32-10, key (seed_lo, seed_hi), counter (record index, epoch, 0, 0);
 *   words 0,1 -> crop offsets (u32 % (2^31-1)) % (2*pad+1), word 3 -> flip iff
 *   bitcast((u & 0x7fffff)|0x3f800000) - 1 < 0.5.  If draws_in != NULL its int32 [B][3]
 *   (oy, ox, flip) are used instead.  draws_out (nullable) receives the (oy, ox, flip) used.
 *   train == 0: no pad/crop/flip (eval path, standardise only).
 *   staged != 0: `records` holds only this batch — row b is the record of image b, e.g. a buffer
 *   just copied from (pinned) host memory when the record file is not kept in HBM — while
 *   indices[b] still names the record for the random draws, so a staged batch decodes to exactly
 *   the same images as the resident path.
 * ------------------------------------------------------------------------------------------ */
int rn_cifar_decode(const uint8_t* records, int64_t n_records, int32_t record_bytes, int32_t label_offset,
                    int32_t classes, const int32_t* indices, int32_t B, int32_t train, int32_t pad,
                    uint32_t seed_lo, uint32_t seed_hi, uint32_t epoch, const int32_t* draws_in,
                    int32_t staged, void* out_img, int32_t out_dtype, float* out_onehot, int32_t* out_labels,
                    int32_t* draws_out, void* stream);

/* ------------------------------------------------------------------------------------------
 * ImageNet input preprocessing (decoded uint8 images of arbitrary sizes -> network input)
 *   replaces vgg_preprocessing.preprocess_image (vgg_preprocessing.py:284-353) as called from
 *   record_parser (resnet_imagenet_train.py:137-147): image/255; aspect-preserving bilinear resize
 *   (align_corners=False) to a smallest side drawn from [side_min, side_max] (train; the reference uses
 *   256..512) or to side_min (eval; 256); random / central crop out_h x out_w; random left-right flip
 *   (train); subtraction of the per-channel means 123.68/255, 116.78/255, 103.94/255.
 *   images: the B decoded HWC uint8 images packed back to back; offsets[b] = byte offset of image b;
 *   hw[b] = (height, width).  Random draws: Philox4x32-10, key (seed_lo, seed_hi), counter
 *   (index[b] or b, epoch, 1, 0): word 0 -> side = side_min + u % (side_max - side_min + 1),
 *   words 1, 2 -> crop offsets u % (resized size - crop + 1), word 3 -> flip iff the 23-bit uniform < 0.5.
 *   draws_in (nullable) int32 [B][4] = (side, offset_y, offset_x, flip) overrides them; draws_out (nullable)
 *   receives the values used.  error_flag (nullable device int, caller zeroes it): set to b + 1 if image b
 *   is smaller than the crop after resizing (the reference's "Crop size greater than the image size.").
 *   Output NHWC [B][out_h][out_w][3], RN_F32 or RN_BF16; if out_onehot != NULL also the dense one-hot
 *   float [B][classes] of labels[b] (tf.one_hot, resnet_imagenet_train.py:144-147).  JPEG decoding is not
 *   part of this library.
 * ------------------------------------------------------------------------------------------ */
int rn_vgg_preprocess(const uint8_t* images, const int64_t* offsets, const int32_t* hw, int32_t B, int32_t out_h,
                      int32_t out_w, int32_t side_min, int32_t side_max, int32_t train, uint32_t seed_lo,
                      uint32_t seed_hi, uint32_t epoch, const int32_t* index, const int32_t* draws_in, void* out,
                      int32_t out_dtype, int32_t* draws_out, int32_t* error_flag, const int32_t* labels,
                      int32_t classes, float* out_onehot, void* stream);

/* ------------------------------------------------------------------------------------------
 * Small utilities used by the host graph executor (no reference counterpart beyond tf.cast /
 * AddN / tf.pad at the call sites above).
 * ------------------------------------------------------------------------------------------ */
int rn_cast(const void* src, void* dst, int64_t n, int32_t src_dtype, int32_t dst_dtype, void* stream);
/* dst += src  (AddN of gradients when no fused path applies) */
int rn_add_inplace(void* dst, const void* src, int64_t n, int32_t dtype, void* stream);
/* y = a + b (forward residual add when it cannot be fused into a BN statistics pass) */
int rn_add(const void* a, const void* b, void* y, int64_t n, int32_t dtype, void* stream);
/* NHWC channel pad/copy: dst[..., 0:Cdst] = src[..., 0:Csrc] zero-extended (stem 3 -> 8 channels). */
int rn_pad_channels(const void* src, void* dst, int64_t pixels, int32_t Csrc, int32_t Cdst, int32_t src_dtype,
                    int32_t dst_dtype, void* stream);

/* CRC-32C of a HOST buffer, continuing from `crc` (0 to start).  Host-only helper for reading and
 * writing the reference's TF-V2 checkpoint bundles (what tf.train.Saver / MonitoredTrainingSession
 * produce at resnet_cifar_train.py:315-319 and saver.restore consumes at resnet_cifar_eval.py:85):
 * both the tensor data and the index-table blocks carry masked CRC-32C checksums. */
uint32_t rn_crc32c(const void* data, size_t n, uint32_t crc);

#ifdef __cplusplus
}
#endif
#endif /* RN_B200_H_ */
