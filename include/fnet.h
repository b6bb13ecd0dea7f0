/*
 * fnet.h -- C ABI of libfnet.so: the sm_100a kernels behind FeatureNet's per-candidate
 * train+evaluate path.
 *
 * The reference (yqtianust/FeatureNet) has no FFI layer of its own: every arithmetic step is
 * a Keras 2.2.4 layer call executed by TensorFlow 1.12 inside `session.run`.  Each entry
 * point below therefore names the reference call site whose device work it replaces
 * (paths relative to the reference tree).  INTEGRATION.md shows the ctypes binding that the
 * Python host side (featurenet_b200/lib.py) uses.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name says host; tensors are fp32, NHWC,
 *     dense, allocated and owned by the caller; nothing here allocates or frees memory;
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, nothing
 *     synchronises, so every call may be captured into a CUDA graph;
 *   - functions return 0 on success and a negative code otherwise; fnet_last_error() returns
 *     the thread-local message of the last failure; no exception crosses the boundary;
 *   - `accumulate != 0` means "add into the output" (a tensor with several consumers
 *     receives its gradient in several pieces), 0 means "overwrite";
 *   - workspaces are caller-provided: ask fnet_*_workspace_bytes() first.
 */
#ifndef FNET_H
#define FNET_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FNET_VERSION 100

enum { FNET_OK = 0, FNET_ERR_ARG = -1, FNET_ERR_CUDA = -2, FNET_ERR_WORKSPACE = -3, FNET_ERR_UNSUPPORTED = -4 };

/* keras.activations used by model/input.py:190,294-298 and model/operation.py:165 */
enum { FNET_ACT_NONE = 0, FNET_ACT_RELU = 1, FNET_ACT_TANH = 2, FNET_ACT_SIGMOID = 3, FNET_ACT_SOFTMAX = 4 };

enum { FNET_POOL_MAX = 0, FNET_POOL_AVG = 1 };

/* arithmetic used by the GEMM-shaped kernels */
enum { FNET_MATH_FP32 = 0,   /* FFMA, fp32 end to end: the parity mode                  */
       FNET_MATH_TF32 = 1,   /* tcgen05.mma kind::tf32, fp32 accumulate in TMEM         */
       FNET_MATH_BF16 = 2,   /* reserved: bf16 activation storage is NOT built; every entry point refuses it with
                              * FNET_ERR_UNSUPPORTED instead of running another mode under its name      */
       FNET_MATH_TF32X3 = 3,   /* fp32 accuracy on the tensor cores: every operand is split into a tf32 head and
                                * a tf32 tail, three tcgen05 MMAs per product (hi*hi + lo*hi + hi*lo)          */
       FNET_MATH_AUTO = 4 };   /* fp32 accuracy, fastest kernel: FNET_MATH_FP32 or FNET_MATH_TF32X3 per product,
                                * chosen from the GEMM shape alone (same shape -> same kernel -> same bits)    */

int fnet_version(void);
const char *fnet_last_error(void);

/* ------------------------------------------------------------------------------------------
 * Convolution.  Replaces keras.layers.Conv2D (model/input.py:294; the channel-aligning
 * "Reg_" convolution of model/operation.py:186) and, with kh = kw = 1, the pointwise half of
 * SeparableConv2D (model/input.py:296).  x[B,H,W,C], w[kh,kw,C,K] (HWIO), y[B,Ho,Wo,K].
 * Padding is explicit (TF SAME puts the smaller half first): out = (in + pads - k_eff)/stride + 1.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
    int32_t batch, in_h, in_w, in_c;
    int32_t out_h, out_w, out_c;
    int32_t kh, kw, stride_h, stride_w, pad_top, pad_left, dil_h, dil_w;
    int32_t act;  /* FNET_ACT_*: fused into the forward epilogue (softmax runs as a second kernel) */
    int32_t math; /* FNET_MATH_* */
} fnet_conv2d_desc;

size_t fnet_conv2d_workspace_bytes(const fnet_conv2d_desc *d);
/* y = act(conv(x, w) + bias); bias may be NULL */
int fnet_conv2d_fwd(const fnet_conv2d_desc *d, const float *x, const float *w, const float *bias, float *y,
                    void *ws, size_t ws_bytes, void *stream);
/* dx (+)= conv_transpose(dz, w), dz = dL/d(pre-activation) */
int fnet_conv2d_bwd_data(const fnet_conv2d_desc *d, const float *dz, const float *w, float *dx, int accumulate,
                         void *ws, size_t ws_bytes, void *stream);
/* the same, followed by the derivative of the activation of the layer that PRODUCED x:
 *   dx = (accumulate ? dx + conv_transpose(dz, w) : conv_transpose(dz, w)) * act'(y)
 * with y that layer's output (same shape as dx) and act relu / tanh / sigmoid.  This is Keras' backward through
 * `Conv2D(..., activation=act)` followed by another layer, minus one pass over the tensor (the separate
 * fnet_act_bwd).  With several gradient pieces, only the call that delivers the LAST one may carry the activation. */
int fnet_conv2d_bwd_data_act(const fnet_conv2d_desc *d, const float *dz, const float *w, float *dx, int accumulate,
                             int act, const float *y, void *ws, size_t ws_bytes, void *stream);
/* dw = x (*) dz, dbias = column sums of dz (dbias may be NULL); deterministic split reduction */
int fnet_conv2d_bwd_filter(const fnet_conv2d_desc *d, const float *x, const float *dz, float *dw, float *dbias,
                           void *ws, size_t ws_bytes, void *stream);

/* ------------------------------------------------------------------------------------------
 * Dense.  Replaces keras.layers.Dense on the last axis of a rank-2 or rank-4 tensor
 * (model/input.py:190) and the classifier head (model/keras_model.py:124).
 * x[M,K], w[K,N], y[M,N]; M = product of the leading axes.
 * ------------------------------------------------------------------------------------------ */
/* 1 when the data gradient of this layer runs on a tensor-core kernel, 0 when on the FFMA / streaming kernels, -1 on a
 * bad argument.  A planner's question, not a computation (no reference counterpart): the activation derivative of
 * the producing layer is folded into fnet_*_bwd_data_act only where that epilogue is cheap. */
int fnet_conv2d_bwd_data_on_tensor_core(const fnet_conv2d_desc *d);
int fnet_dense_bwd_data_on_tensor_core(int64_t m, int64_t k, int64_t n, int math);
size_t fnet_dense_workspace_bytes(int64_t m, int64_t k, int64_t n);
int fnet_dense_fwd(int64_t m, int64_t k, int64_t n, int act, int math, const float *x, const float *w,
                   const float *bias, float *y, void *ws, size_t ws_bytes, void *stream);
int fnet_dense_bwd_data(int64_t m, int64_t k, int64_t n, int math, const float *dz, const float *w, float *dx,
                        int accumulate, void *ws, size_t ws_bytes, void *stream);
/* fnet_dense_bwd_data with the activation derivative of the producing layer folded in (see fnet_conv2d_bwd_data_act) */
int fnet_dense_bwd_data_act(int64_t m, int64_t k, int64_t n, int math, const float *dz, const float *w, float *dx,
                            int accumulate, int act, const float *y, void *ws, size_t ws_bytes, void *stream);
int fnet_dense_bwd_filter(int64_t m, int64_t k, int64_t n, int math, const float *x, const float *dz, float *dw,
                          float *dbias, void *ws, size_t ws_bytes, void *stream);

/* ------------------------------------------------------------------------------------------
 * Dense forward that allocates nothing and needs no workspace.  It has no counterpart in the
 * reference; it exists for hosts that cannot allocate device memory -- above all libfnet
 * loaded into a TensorFlow 1.x process (the reference's own runtime), where the session has
 * already reserved almost the whole GPU and a plug-in's cudaMalloc may fail, and more generally
 * a worker sharing the device with a framework whose pool owns the memory, or a loop that must
 * not pay the device-wide synchronisation of cudaMalloc/cudaFree.
 *
 * Same arithmetic as fnet_dense_fwd with FNET_MATH_FP32: fp32 FFMA tiles with a fixed reduction
 * order (fnet_dense_fwd may choose another fp32 kernel for a shape, so the two agree to rounding,
 * not to the bit).  The partial sums of the split reduction go to an array of
 * FNET_STATIC_SCRATCH_BYTES built into the library; every GPU that loads libfnet.so holds that
 * array until the process exits.  Shapes whose reduction needs more than that are rejected
 * with FNET_ERR_WORKSPACE and a message (use fnet_dense_fwd and a workspace of
 * fnet_dense_workspace_bytes() for those).
 *
 * The first call on each device looks up the array's device address (cudaGetSymbolAddress,
 * with that device current); make that first call before starting any stream capture.
 * All calls on a device share the one array: the array keeps whatever the last call wrote, and
 * no call depends on it (every partial is written before it is read, nothing is assumed zero);
 * but two calls that run on the GPU at the same time -- different streams, or different host
 * threads -- race on it, so keep calls to this entry point ordered on one stream per device.
 * ------------------------------------------------------------------------------------------ */
#define FNET_STATIC_SCRATCH_BYTES (4u << 20)
int fnet_dense_fwd_static(int64_t m, int64_t k, int64_t n, int act, const float *x, const float *w,
                          const float *bias, float *y, void *stream);

/* ------------------------------------------------------------------------------------------
 * Depthwise convolution, depth multiplier 1.  Replaces keras.layers.DepthwiseConv2D
 * (model/input.py:298) and the depthwise half of SeparableConv2D (model/input.py:296).
 * x[B,H,W,C], w[kh,kw,C], y[B,Ho,Wo,C].  Memory-bound: 2*kh*kw flops per element moved.
 * ------------------------------------------------------------------------------------------ */
size_t fnet_dwconv2d_workspace_bytes(const fnet_conv2d_desc *d);
int fnet_dwconv2d_fwd(const fnet_conv2d_desc *d, const float *x, const float *w, const float *bias, float *y,
                      void *stream);
int fnet_dwconv2d_bwd_data(const fnet_conv2d_desc *d, const float *dz, const float *w, float *dx, int accumulate,
                           void *stream);
int fnet_dwconv2d_bwd_filter(const fnet_conv2d_desc *d, const float *x, const float *dz, float *dw, float *dbias,
                             void *ws, size_t ws_bytes, void *stream);

/* ------------------------------------------------------------------------------------------
 * Pooling.  Replaces MaxPooling2D / AveragePooling2D / GlobalAveragePooling2D
 * (model/input.py:233,237,240).  Average divides by the number of in-bounds taps; the max
 * backward routes to the first maximum in row-major window order (argmax holds that tap).
 * ------------------------------------------------------------------------------------------ */
typedef struct {
    int32_t batch, in_h, in_w, channels;
    int32_t out_h, out_w;
    int32_t kh, kw, stride_h, stride_w, pad_top, pad_left;
    int32_t kind; /* FNET_POOL_* */
} fnet_pool2d_desc;

int fnet_pool2d_fwd(const fnet_pool2d_desc *d, const float *x, float *y, uint8_t *argmax /* max only, may be NULL */,
                    void *stream);
int fnet_pool2d_bwd(const fnet_pool2d_desc *d, const float *dy, const uint8_t *argmax, float *dx, int accumulate,
                    void *stream);
int fnet_gap_fwd(int32_t batch, int32_t hw, int32_t channels, const float *x, float *y, void *stream);
int fnet_gap_bwd(int32_t batch, int32_t hw, int32_t channels, const float *dy, float *dx, int accumulate,
                 void *stream);

/* ------------------------------------------------------------------------------------------
 * Batch normalisation over axis 1.  Replaces keras.layers.BatchNormalization(axis=1)
 * (model/operation.py:142,150): the tensor is viewed as [outer, A, inner] (NHWC rank 4:
 * outer=B, A=H, inner=W*C; rank 2: outer=B, A=features, inner=1); statistics are per a over
 * outer*inner elements.  Training normalises with the biased batch variance and feeds
 * var*n/(n-(1+eps)) to the moving variance (Keras 2.2.x).
 * ws: fnet_bn_workspace_bytes(A) bytes.
 * ------------------------------------------------------------------------------------------ */
size_t fnet_bn_workspace_bytes(int32_t a);
int fnet_bn_fwd_train(int32_t outer, int32_t a, int32_t inner, float eps, float momentum, const float *x,
                      const float *gamma, const float *beta, float *y, float *save_mean, float *save_rstd,
                      float *moving_mean, float *moving_var, void *ws, size_t ws_bytes, void *stream);
int fnet_bn_fwd_infer(int32_t outer, int32_t a, int32_t inner, float eps, const float *x, const float *gamma,
                      const float *beta, const float *moving_mean, const float *moving_var, float *y,
                      void *stream);
/* input gradient in inference mode (moving statistics are constants): dx (+)= dy * gamma / sqrt(moving_var + eps);
 * what an adversarial attack differentiates (tensorflow_generator.py:151-173 through ART's loss_gradient) */
int fnet_bn_bwd_infer(int32_t outer, int32_t a, int32_t inner, float eps, const float *dy, const float *gamma,
                      const float *moving_var, float *dx, int accumulate, void *stream);
int fnet_bn_bwd(int32_t outer, int32_t a, int32_t inner, const float *x, const float *dy, const float *gamma,
                const float *save_mean, const float *save_rstd, float *dx, float *dgamma, float *dbeta,
                int accumulate, void *ws, size_t ws_bytes, void *stream);
/* The same when x is the OUTPUT of an activated layer with no other consumer: dx is also multiplied by that
 * activation's derivative, taken from x itself (act = FNET_ACT_RELU / TANH / SIGMOID; NONE = fnet_bn_bwd). */
int fnet_bn_bwd_act(int32_t outer, int32_t a, int32_t inner, const float *x, const float *dy, const float *gamma,
                    const float *save_mean, const float *save_rstd, float *dx, float *dgamma, float *dbeta,
                    int accumulate, int act, void *ws, size_t ws_bytes, void *stream);

/* ------------------------------------------------------------------------------------------
 * Element-wise and data-movement ops.
 * act:      keras.layers.Activation (model/operation.py:165); softmax is over rows of `cols`.
 * dropout:  keras.layers.Dropout / tf.nn.dropout (model/operation.py:114): keep <=> philox word
 *           >= ceil(rate*2^32); element i uses word i%4 of philox4x32-10(counter=(i/4, *step,
 *           stream_id, 0), key=seed); kept values are scaled by 1/(1-rate).  The backward
 *           recomputes the mask.  `step` is a device int32 (the optimizer iteration).
 * add/mul:  keras.layers.Add / Multiply (model/operation.py:222,248).
 * rows:     copy `rows` rows of `inner` floats between [batch, *_rows, inner] tensors:
 *           Concatenate(axis=1) forward/backward (model/operation.py:238).
 * pad2d:    keras.layers.ZeroPadding2D (model/operation.py:136).
 * ------------------------------------------------------------------------------------------ */
int fnet_act_fwd(int act, int64_t rows, int64_t cols, const float *x, float *y, void *stream);
/* dx (+)= dy * act'(.) given the activation OUTPUT y */
int fnet_act_bwd(int act, int64_t rows, int64_t cols, const float *y, const float *dy, float *dx, int accumulate,
                 void *stream);
int fnet_dropout_fwd(int64_t n, float rate, uint64_t seed, uint32_t stream_id, const int32_t *step,
                     const float *x, float *y, void *stream);
int fnet_dropout_bwd(int64_t n, float rate, uint64_t seed, uint32_t stream_id, const int32_t *step,
                     const float *dy, float *dx, int accumulate, void *stream);
int fnet_add_fwd(int64_t n, const float *a, const float *b, float *y, void *stream);
int fnet_mul_fwd(int64_t n, const float *a, const float *b, float *y, void *stream);
/* dx (+)= dy * other */
int fnet_mul_bwd(int64_t n, const float *dy, const float *other, float *dx, int accumulate, void *stream);
/* dst (+)= src */
int fnet_accumulate(int64_t n, const float *src, float *dst, int accumulate, void *stream);
int fnet_copy_rows(int32_t batch, int32_t src_rows, int32_t src_row0, int32_t dst_rows, int32_t dst_row0,
                   int32_t rows, int32_t inner, const float *src, float *dst, int accumulate, void *stream);
/* Concatenate(axis=1) of a[outer, rows_a, inner] and b[outer, rows_b, inner] in one launch, and its backward
 * (da / db may be NULL when that input needs no gradient; a and b must be distinct buffers) */
int fnet_concat_rows_fwd(int32_t outer, int32_t rows_a, int32_t rows_b, int32_t inner, const float *a, const float *b,
                         float *y, void *stream);
int fnet_concat_rows_bwd(int32_t outer, int32_t rows_a, int32_t rows_b, int32_t inner, const float *dy, float *da,
                         float *db, int acc_a, int acc_b, void *stream);
/* The same, with the activation derivative of the layer(s) that produced the operands folded in:
 * da = (acc_a ? da + slice : slice) * act_a'(ya) (ya = that layer's forward output; act 0 = none), likewise b. */
int fnet_concat_rows_bwd_act(int32_t outer, int32_t rows_a, int32_t rows_b, int32_t inner, const float *dy, float *da, float *db,
                             int acc_a, int acc_b, int act_a, const float *ya, int act_b, const float *yb, void *stream);
int fnet_pad2d_fwd(int32_t batch, int32_t h, int32_t w, int32_t c, int32_t pad_h, int32_t pad_w, const float *x,
                   float *y, void *stream);
int fnet_pad2d_bwd(int32_t batch, int32_t h, int32_t w, int32_t c, int32_t pad_h, int32_t pad_w, const float *dy,
                   float *dx, int accumulate, void *stream);
int fnet_fill(int64_t n, float value, float *dst, void *stream);

/* ------------------------------------------------------------------------------------------
 * Loss and metric.  Replaces loss='categorical_crossentropy', metrics=['accuracy'] of
 * tensorflow_generator.py:233-235 on the softmax output of the "out" layer
 * (model/keras_model.py:124): per row, renormalise p, clip to [1e-7, 1-1e-7], -sum t*log p;
 * correct <=> argmax(t) == argmax(p) (first index on ties, integer compare).
 * Adds the batch loss to *loss_sum (double) and the number of correct rows to *correct.
 * If dlogits != NULL also writes scale * dL/d(logits) through softmax, renormalisation and clip.
 * ------------------------------------------------------------------------------------------ */
int fnet_softmax_cce(int32_t batch, int32_t classes, const float *input, const float *targets, float scale,
                     int from_logits, float *probs_out, float *dlogits, double *loss_sum, int64_t *correct,
                     void *stream);
/* from_logits = 0: `input` holds the probabilities of the "out" layer, probs_out is ignored.
 * from_logits = 1: `input` holds the logits; the kernel computes the softmax itself and writes the
 * probabilities to probs_out (which may be `input`: in place), saving the head's softmax launch. */

/* ------------------------------------------------------------------------------------------
 * Optimizer and loop state.  Replaces keras.optimizers.Adam(lr=1e-3) of
 * tensorflow_generator.py:232 (and "sgd" of lenet5.py:21) as ONE launch over the flat
 * parameter arena, plus the per-step bookkeeping that Keras does on the host.
 * state (device, int32[8] viewed loosely): [0] t = completed iterations, [1] cursor = next
 * sample of the epoch, [2] lr (float bits), [3] lr_t (float bits, written by step_begin).
 * ------------------------------------------------------------------------------------------ */
int fnet_step_begin(int32_t *state, float beta1, float beta2, void *stream);
/* state[2] = lr, in stream order.  The value travels as a kernel argument, so callbacks that change the learning
 * rate (LearningRateScheduler per epoch, helpers.py:45-67; SGDRScheduler per batch, sgdr.py:60-66) cost one tiny
 * launch and no host wait. */
int fnet_state_set_lr(int32_t *state, float lr, void *stream);
int fnet_step_end(int32_t *state, int32_t batch, void *stream);
/* advance >= 0: the optimizer launch also does fnet_step_end(state, advance); advance < 0: state is only read */
int fnet_adam_step(int64_t n, float *params, const float *grads, float *m, float *v, int32_t *state, float beta1,
                   float beta2, float eps, int32_t advance, void *stream);
int fnet_sgd_step(int64_t n, float *params, const float *grads, float *velocity, int32_t *state, float momentum,
                  int nesterov, int32_t advance, void *stream);
/* fnet_gather_rows for the inputs AND the targets of a batch in one launch; begin_step != 0 also does the work of
 * fnet_step_begin(state, beta1, beta2).  x_dst[i, :] = x_src[perm[cursor+i], :], y_dst likewise. */
int fnet_gather_batch(int32_t batch, int64_t x_row, int64_t y_row, const int32_t *perm, int32_t *state,
                      const float *x_src, const float *y_src, float *x_dst, float *y_dst, int begin_step, float beta1,
                      float beta2, void *stream);
/* dst[i, :] = src[perm[cursor + i], :] for i < batch; replaces the host-side batch slicing of
 * keras Model.fit(shuffle=True) (helpers.py:105-111) on a device-resident dataset */
int fnet_gather_rows(int32_t batch, int64_t row_elems, const int32_t *perm, const int32_t *state,
                     const float *src, float *dst, void *stream);
/* The augmenting variant of fnet_gather_rows for images [*, h, w, c]: replaces
 * ImageDataGenerator(width_shift_range, height_shift_range, horizontal_flip, fill_mode='nearest').flow()
 * of helpers.py:115-167.  For position p = cursor + i of the epoch:
 * dst[i, r, q, :] = bilinear(src[perm[p]], r + tx[p], q' + ty[p]) with q' = w-1-q when flip[p] != 0,
 * coordinates clamped to the image (order-1 interpolation, 'nearest' fill).  tx, ty, flip are
 * per-epoch arrays (device), drawn by the host. */
int fnet_gather_shift_flip(int32_t batch, int32_t h, int32_t w, int32_t c, const int32_t *perm,
                           const int32_t *state, const float *tx, const float *ty, const uint8_t *flip,
                           const float *src, float *dst, void *stream);
/* dst = float(src) * scale : the `astype('float32') / 255` of tensorflow_generator.py:310-313 */
int fnet_u8_to_f32(int64_t n, const uint8_t *src, float scale, float *dst, void *stream);
/* First launch of a train step whose batch was uploaded as uint8 pixels (a quarter of the fp32 upload): the
 * `astype('float32') / 255` of tensorflow_generator.py:310-313 applied on the device to the whole batch
 * [batch][x_row], targets [batch][y_row] copied, and (begin_step != 0) the work of fnet_step_begin. */
int fnet_feed_batch_u8(int32_t batch, int64_t x_row, int64_t y_row, int32_t *state, const uint8_t *x_src, float scale,
                       const float *y_src, float *x_dst, float *y_dst, int begin_step, float beta1, float beta2,
                       void *stream);

/* ------------------------------------------------------------------------------------------
 * One step of an L2 gradient attack on a batch of inputs (the FGSM / PGD updates behind the robustness
 * fitness, tensorflow_generator.py:151-218 -> model/metrics.py:57-91 -> ART 0.9.0 FastGradientMethod /
 * ProjectedGradientDescent with norm 2).  Per row b with active[b] != 0 (active == NULL: all rows):
 *   delta = (adv_in[b] - x0[b]) + step * grad[b] / (||grad[b]||_2 + 1e-7)
 *   if radius > 0 and ||delta||_2 > radius: delta *= radius / (||delta||_2 + 1e-7)      (projection)
 *   adv_out[b] = clip(x0[b] + delta, clip_lo, clip_hi)
 * inactive rows of adv_out are left as they are.  adv_out may alias adv_in.  One CTA per row, fixed reduction order.
 * ------------------------------------------------------------------------------------------ */
int fnet_l2_attack_step(int32_t batch, int64_t row, const float *x0, const float *grad, const float *adv_in,
                        float *adv_out, float step, float radius, float clip_lo, float clip_hi,
                        const uint8_t *active, void *stream);

/* ------------------------------------------------------------------------------------------
 * CUDA-graph helpers: one train step (or eval step) is captured once and replayed, which
 * replaces one `session.run` per step.
 * ------------------------------------------------------------------------------------------ */
/* ------------------------------------------------------------------------------------------
 * Fused step.  A candidate with a handful of narrow layers (the products of main_1block_nas.xml: 1x1 / 2x2
 * convolutions between 3 and 6 channels, a concatenation, the softmax head -- 61 k parameters, 3 MB of activations
 * per batch) spends its step in launch latency: 13 launches of 4 us for 7 us of memory traffic.  Replaces the
 * whole `train_function` of one step (helpers.py:105 -> Model.fit) with ONE cooperative launch: a persistent
 * kernel, one CTA per SM, walks the list of operations below and separates dependent operations by a grid-wide
 * barrier instead of a kernel boundary; the activations stay in L2 from one operation to the next.
 * `ops` lives in DEVICE memory (the caller uploads it once per program); operations are executed in order, and
 * `barrier_after` says where later operations depend on earlier ones.  `sync` is two zero-initialised uint32 in
 * device memory owned by this program (never shared between programs that may run concurrently).
 * ------------------------------------------------------------------------------------------ */
enum {
    FNET_FK_BEGIN = 1,        /* p0 state; f0 beta1, f1 beta2: fnet_step_begin */
    FNET_FK_GATHER = 2,       /* fnet_gather_batch: d0 batch d1 x_row d2 y_row d3 begin_step; p0 perm p1 state p2 x_src p3 y_src p4 x_dst p5 y_dst; f0 f1 betas */
    FNET_FK_FEED_U8 = 3,      /* fnet_feed_batch_u8: d0 batch d1 x_row d2 y_row d3 begin_step; p0 state p1 x_src(u8) p2 y_src p3 x_dst p4 y_dst; f0 f1 betas f2 scale */
    FNET_FK_CONV_FWD = 4,     /* d0..d14 = B H W C Ho Wo K kh kw sh sw pt pl dh dw; act; p0 x p1 w p2 bias p3 y */
    FNET_FK_CONV_DGRAD = 5,   /* geometry as above; flags bit0 accumulate; act = activation of the producer (0 none); p0 dz p1 w p2 dx p3 y_of_producer */
    FNET_FK_CONV_WGRAD = 6,   /* geometry; flags bit0 = with bias row; p0 x p1 dz p2 partial[grid][rows*K] */
    FNET_FK_REDUCE = 7,       /* d0 parts d1 total d2 split (first d2 values -> p1, the rest -> p2); p0 partial p1 dw p2 dbias */
    FNET_FK_CONCAT_FWD = 8,   /* d0 outer d1 blk_a d2 blk_b (elements per outer index); p0 a p1 b p2 y */
    FNET_FK_CONCAT_BWD = 9,   /* same dims; flags bit0 acc_a bit1 acc_b; p0 dy p1 da (may be NULL) p2 db (may be NULL) */
    FNET_FK_HEAD_FWD_LOSS = 10, /* Dense(softmax) + categorical cross-entropy: d0 M d1 K d2 N; f0 scale; p0 x p1 w p2 bias p3 probs p4 targets p5 dz p6 row_loss[M] p7 row_hit[M] */
    FNET_FK_HEAD_DGRAD = 11,  /* d0 M d1 K d2 N; flags bit0 accumulate; act; p0 dz p1 w p2 dx p3 y_of_producer */
    FNET_FK_HEAD_WGRAD = 12,  /* d0 M d1 K d2 N; p0 x p1 dz p2 dw p3 dbias */
    FNET_FK_ACT_BWD = 13,     /* d0 n (low 31 bits) ; flags bit0 accumulate; act; p0 y p1 dy p2 dx */
    FNET_FK_ADAM = 14,        /* d0 n d1 advance d2 M (rows of the loss, 0 = none); f0 b1 f1 b2 f2 eps; p0 params p1 grads p2 m p3 v p4 state p5 row_loss p6 row_hit p7 metrics(double loss_sum, int64 correct) */
    FNET_FK_SGD = 15          /* d0 n d1 advance d2 M d3 nesterov; f0 momentum; p0 params p1 grads p2 velocity p4 state p5 row_loss p6 row_hit p7 metrics */
};
typedef struct fnet_fused_op {
    int32_t kind, barrier_after, flags, act;
    int32_t d[16];
    float f[4];
    void *p[8];
} fnet_fused_op;
/* The classifier head and its loss in ONE launch: Dense(n <= 16, softmax) of keras_model.py:124 on x[m][k], categorical
 * cross-entropy and accuracy of tensorflow_generator.py:232-235 against targets[m][n]: probs = softmax(x.w + bias),
 * dlogits (may be NULL) = scale * dL/dlogits, *loss_sum += batch loss, *correct += hits -- what fnet_dense_fwd followed
 * by fnet_softmax_cce(from_logits) computes, without the second launch.  `ticket`: one zero-initialised uint32 in
 * device memory owned by the caller (the kernel re-arms it); ws: fnet_head_fwd_loss_workspace_bytes(m) of scratch. */
size_t fnet_head_fwd_loss_workspace_bytes(int64_t m);
int fnet_head_fwd_loss(int64_t m, int64_t k, int64_t n, const float *x, const float *w, const float *bias, float *probs,
                       const float *targets, float scale, float *dlogits, double *loss_sum, int64_t *correct, uint32_t *ticket,
                       void *ws, size_t ws_bytes, void *stream);
/* CTAs the fused kernel must be launched with on the current device (its operation list sizes partials by it). */
int fnet_fused_step_grid(void);
/* 1 when a convolution / head of this shape has a fused implementation (small channel counts only), else 0. */
int fnet_fused_conv_supported(const fnet_conv2d_desc *d, int product /* 0 fwd, 1 dgrad, 2 wgrad */, int with_bias);
int fnet_fused_head_supported(int64_t m, int64_t k, int64_t n);
int fnet_fused_step(const fnet_fused_op *ops_dev, int32_t n_ops, uint32_t *sync, void *stream);
/* The same launch; CTA 0 also writes the GPU's nanosecond timer at kernel entry and after every operation into
 * stamps[0 .. n_ops] (device memory): where the time of a step goes, phase by phase (scripts/fused_phases.py). */
int fnet_fused_step_timed(const fnet_fused_op *ops_dev, int32_t n_ops, uint32_t *sync, uint64_t *stamps, void *stream);

/* A stream of the caller's own (cudaStreamNonBlocking).  Every candidate that is resident on a GPU captures and replays
 * its steps on streams nobody else uses: a stream handed out by a framework's pool (torch recycles 32 per device)
 * can be shared by two candidates, and launches that one host thread issues while another thread is capturing on
 * the same stream are recorded into the wrong graph instead of being run. */
int fnet_stream_create(void **stream_out);
int fnet_stream_destroy(void *stream);
int fnet_graph_begin(void *stream);
int fnet_graph_end(void *stream, void **graph_exec_out);
int fnet_graph_launch(void *graph_exec, void *stream);
/* n back-to-back replays in one call: an epoch of Model.fit (helpers.py:105) without per-batch callbacks is
 * floor(N / batch) identical steps -- one transition from the host language per epoch instead of one per step. */
int fnet_graph_launch_n(void *graph_exec, int32_t n, void *stream);
int fnet_graph_destroy(void *graph_exec);
/* One step fed from HOST memory, the shape of keras' train_on_batch / test_on_batch (one session.run with a
 * feed_dict, helpers.py:105 -> Model.fit -> train_function): copy the inputs and the targets to the device,
 * replay the captured step, copy `result_bytes` of results (loss sum, correct count) back and wait for them.
 * All on `stream`; host buffers may be pageable (pinned ones make the copies asynchronous until the final
 * wait).  Either host source may be NULL to skip that copy. */
int fnet_graph_step_fed(void *graph_exec, void *x_dev, const void *x_host, size_t x_bytes, void *y_dev,
                        const void *y_host, size_t y_bytes, const void *result_dev, void *result_host,
                        size_t result_bytes, void *stream);

/* One step of a PIPELINED host feed, the shape of keras' fit_generator (helpers.py:163: batches arrive from a
 * host-side queue while the previous step still runs).  The caller owns a ring of slots (two are enough); a slot is
 * a device buffer pair (x_slot, y_slot) the captured step reads its batch from, optional pinned staging buffers,
 * a pinned `result_host` and two events.  The call
 *   1. waits on the host for `slot_event`, i.e. for the step that last used this slot, and copies that step's
 *      results from `result_host` to `result_prev` (skipped when result_prev is NULL);
 *   2. if x_stage / y_stage are given, memcpy()s the (pageable) host batch into them; without staging buffers the
 *      host batch itself must be pinned and must stay untouched until the slot's next turn;
 *   3. on `copy_stream`: both host-to-device copies, then records `copied_event`;
 *   4. on `compute_stream`: waits for `copied_event`, replays the captured step, copies `result_bytes` from
 *      `result_dev` to `result_host`, records `slot_event`.
 * It does not wait for any of that: with two slots the upload of step i+1 overlaps the kernels of step i.
 * Events are cudaEvent_t handles the caller created (record each once before the first use). */
int fnet_graph_step_streamed(void *graph_exec, void *x_slot, const void *x_host, void *x_stage, size_t x_bytes,
                             void *y_slot, const void *y_host, void *y_stage, size_t y_bytes, const void *result_dev,
                             void *result_host, void *result_prev, size_t result_bytes, void *copy_stream,
                             void *compute_stream, void *slot_event, void *copied_event);

#ifdef __cplusplus
}
#endif
#endif /* FNET_H */
