/*
 * flowcuda.h — C ABI of libflowcuda.so: the B200 (sm_100a) implementation of the
 * normalizing-flow density-evaluation / sampling / training hot path of
 * kaijennissen/Normalizing_Flows.
 *
 * This is the drop-in boundary.  The reference has no FFI of its own
 * (it is pure Python on JAX); each entry point below replaces the XLA computation that
 * the cited reference function compiles to, and is what a `jax.ffi` / ctypes binding
 * on the reference side would bind.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer;
 *   - `stream` is a `cudaStream_t` passed as `void*`; all work is enqueued on it, the
 *     library never synchronises, allocates, frees or retains pointers;
 *   - arrays are dense row-major float32 unless stated; sample index is the slowest axis;
 *   - return value: 0 on success, negative `FLOWCUDA_E*` on error; the message is
 *     available (thread-local) from `flowcuda_last_error()`; no C++ exception crosses;
 *   - citations `file:line` are into the reference repository.
 */
#ifndef FLOWCUDA_H_
#define FLOWCUDA_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FLOWCUDA_VERSION 100 /* 0.1.0 */

#define FLOWCUDA_OK 0
#define FLOWCUDA_EINVAL (-1)    /* bad argument (shape, enum, alignment, NULL) */
#define FLOWCUDA_ECUDA (-2)     /* a CUDA runtime call failed */
#define FLOWCUDA_EWORKSPACE (-3)/* workspace too small */
#define FLOWCUDA_EUNSUPPORTED (-4)

int flowcuda_version(void);
const char* flowcuda_last_error(void);
/* Number of SMs of the current device (grid sizing is done in multiples of it). */
int flowcuda_sm_count(void);

/* ------------------------------------------------------------------------------------
 * Row #1 — hand-rolled Real NVP (src/real_nvp.py), event dim D = 2, split d = 1.
 *
 * `params`: n_layers * (4*hidden + 2) floats; per layer W1[1,hidden], b1[hidden],
 * W2[hidden,2] (row-major; column 0 = shift, column 1 = log_scale), b2[2]
 * — the reference's params[layer] = [[W1,b1],[W2,b2]] flattened (real_nvp.py:188-195).
 * `flip_mask`: bit l set <=> layer l is built with flip=True (real_nvp.py:209-218).
 * Conditioner: leaky_relu(0.01) MLP on the pass-through half (real_nvp.py:161-171).
 * n_layers <= 64, 1 <= hidden <= 32.
 * ---------------------------------------------------------------------------------- */

/* log_prob_nvp_chains(params, fns_config, log_prob_N01, y) (real_nvp.py:268-283):
 * the LAST layer's inverse is applied to the data first; base = N(0, I_2)
 * (real_nvp.py:57-68).  y[N,2] -> log_prob[N].
 * Optional outputs (may be NULL): z[N,2] = the base-space point. */
int flowcuda_realnvp_log_prob(const float* params, int n_layers, int hidden, uint64_t flip_mask,
                              const float* y, float* log_prob, float* z, int64_t n, void* stream);

/* sample_nvp_chain with the base noise given (real_nvp.py:228-248): layer 0's
 * forward_nvp first ... layer L-1 last.  z[N,2] -> x[N,2]. */
int flowcuda_realnvp_sample(const float* params, int n_layers, int hidden, uint64_t flip_mask,
                            const float* z, float* x, int64_t n, void* stream);

/* forward_nvp (real_nvp.py:71-101): one layer, x[N,2] -> y[N,2]. `layer_params` = that
 * layer's 4*hidden+2 floats. */
int flowcuda_realnvp_forward(const float* layer_params, int hidden, int flip,
                             const float* x, float* y, int64_t n, void* stream);

/* inverse_nvp (real_nvp.py:104-134): one layer, y[N,2] -> x[N,2], log_scale[N] (the
 * reference returns log_scale as [N,1]; same memory). */
int flowcuda_realnvp_inverse(const float* layer_params, int hidden, int flip,
                             const float* y, float* x, float* log_scale, int64_t n, void* stream);

/* value_and_grad of loss = -mean(log_prob_nvp_chains(...)) (real_nvp.py:292-297).
 * Deterministic two-stage reduction: per-block partials are written to `workspace`
 * (size from flowcuda_realnvp_nll_grad_workspace_bytes) and summed in a fixed order in
 * double precision.  loss[1], grad[n_layers*(4*hidden+2)] (same layout as params).
 * If lr != 0 the SGD update p <- p - lr*g (real_nvp.py:298) is fused into the second
 * stage and written to `params_out` (may alias params). */
size_t flowcuda_realnvp_nll_grad_workspace_bytes(int n_layers, int hidden, int64_t n);
int flowcuda_realnvp_nll_grad(const float* params, int n_layers, int hidden, uint64_t flip_mask,
                              const float* y, int64_t n, float* loss, float* grad,
                              float lr, float* params_out,
                              void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------
 * Row #5 — BernsteinBijector (src/BernsteinBijector.py).  2 <= order M <= 64.
 * ---------------------------------------------------------------------------------- */

/* constrain_thetas (BernsteinBijector.py:60-71): theta_raw[rows,M] -> theta[rows,M]:
 * theta_0 = raw_0, theta_k = theta_{k-1} + softplus(raw_k) + 1e-4. */
int flowcuda_bernstein_constrain_thetas(const float* theta_raw, float* theta, int64_t rows, int order,
                                        void* stream);

/* forward_and_log_det (BernsteinBijector.py:81-93,113-117) on CONSTRAINED thetas:
 *   xc = clip(x, 1e-7, 1-1e-7);  y = mean_k Beta(k, M-k+1).pdf(xc) * theta_k;
 *   log_det = log(sum_k Beta(k, M-k).pdf(xc) * (theta_{k+1}-theta_k)).
 * theta_stride = 0: one theta[M] shared by every element (bflow.py:24);
 * theta_stride = M: theta[n,M], one row per element (maf_mwe.py:103-107).
 * y or log_det may be NULL (forward / forward_log_det_jacobian alone). */
int flowcuda_bernstein_forward_and_log_det(const float* theta, int order, int64_t theta_stride,
                                           const float* x, float* y, float* log_det, int64_t n,
                                           void* stream);

/* inverse (BernsteinBijector.py:95-111): 200-point grid of forward() on
 * linspace(1e-7, 1-1e-7) per event column e, then linear interpolation (jnp.interp,
 * clamped).  theta[E,M] constrained, y[N,E] row-major -> x[E,N] (vmap(in_axes=-1)
 * stacks on axis 0: the reference's transposed output).  The log-det the reference
 * returns for this direction is the constant 0 (BernsteinBijector.py:119-123).
 * workspace: E*n_points floats (n_points = 200 in the reference). */
int flowcuda_bernstein_inverse(const float* theta, int order, int64_t n_events, int n_points,
                               const float* y, float* x, int64_t n, float* workspace, void* stream);

/* Transformed(Normal(0,1), Inverse(Chain([T4,T3,T2,T1]))).log_prob(x) of the Bernstein
 * flow demo (bflow.py:28-34,82-88), fused in one pass:
 *   T1: v = softplus(a)*(x+b);  T2: tfb.SoftClip(0,1,hinge_softness=1.5);
 *   T3: BernsteinBijector(theta) (theta as passed: already constrained as the caller
 *       wants it — the demo double-constrains);
 *   T4: v = softplus(alpha)*(v+beta);  base Normal(0,1).
 * scal[4] = {a, b, alpha, beta} (device). x[n] -> log_prob[n]. */
int flowcuda_bernstein_flow_log_prob(const float* theta, int order, const float* scal,
                                     const float* x, float* log_prob, int64_t n, void* stream);

/* Reverse mode of flowcuda_bernstein_forward_and_log_det — what jax.grad computes through
 * BernsteinBijector.forward_and_log_det (BernsteinBijector.py:82-96; maf_mwe.py:96-113,143).
 * Cotangents ybar[n], ldbar[n] (either may be NULL = zero) -> xbar[n] (d/dx, zero where the clip
 * is active) and thetabar (d/dtheta of the CONSTRAINED thetas): [M] summed over the elements when
 * theta_stride = 0 (bit-reproducible fixed-order reduction), [n, theta_stride] rows otherwise.
 * xbar or thetabar may be NULL.  workspace: flowcuda_bernstein_vjp_workspace_bytes (0 for rows). */
size_t flowcuda_bernstein_vjp_workspace_bytes(int order, int64_t theta_stride, int64_t n);
int flowcuda_bernstein_forward_and_log_det_vjp(const float* theta, int order, int64_t theta_stride,
                                               const float* x, const float* ybar, const float* ldbar,
                                               float* xbar, float* thetabar, int64_t n, void* workspace,
                                               size_t workspace_bytes, void* stream);

/* Reverse mode of flowcuda_bernstein_constrain_thetas: thetabar[rows,M] -> rawbar[rows,M],
 * rawbar_0 = sum_j thetabar_j, rawbar_k = sigmoid(raw_k) * sum_{j>=k} thetabar_j. */
int flowcuda_bernstein_constrain_thetas_vjp(const float* theta_raw, const float* thetabar, float* rawbar,
                                            int64_t rows, int order, void* stream);

/* ------------------------------------------------------------------------------------
 * Rows #2, #3, #4, #6 — distrax MaskedCoupling flows
 * (src/real_nvp_toy_densities.py:56-110, src/maf_toy_densities.py:55-111,
 *  src/maf_mnist.py:34-89).
 * ---------------------------------------------------------------------------------- */

#define FLOWCUDA_TRANSFORM_AFFINE_SCALAR 0 /* ScalarAffine, (shift,log_scale)[N,1] broadcast over D
                                              (real_nvp_toy_densities.py:84-86) */
#define FLOWCUDA_TRANSFORM_AFFINE 1        /* per-dimension (shift,log_scale): cond out [N,D,2] */
#define FLOWCUDA_TRANSFORM_RQS 2           /* RationalQuadraticSpline, cond out [N,D,3K+1]
                                              (maf_toy_densities.py:84-87, maf_mnist.py:63-66) */
#define FLOWCUDA_BASE_NORMAL 0             /* Independent(Normal(0,1)) */
#define FLOWCUDA_BASE_UNIFORM01 1          /* Independent(Uniform(0,1)) (maf_mnist.py:85-88) */
#define FLOWCUDA_MAX_HIDDEN_LAYERS 4

#define FLOWCUDA_PRECISION_FP32 0   /* CUDA-core FP32 FMA conditioner */
#define FLOWCUDA_PRECISION_3XTF32 1 /* tcgen05 error-compensated 3-pass split with 11-bit operand halves (the
                                     * accuracy of a tf32 split: ~2^-21 relative products, fp32 accumulate).
                                     * Since r01 the event_dim-2 kernel carries the halves as fp16 under
                                     * kind::f16 (same significand as tf32, half the TMEM/smem footprint);
                                     * the constant keeps its name for ABI stability.
                                     * RANGE: hidden activations are converted with saturation
                                     * (cvt.satfinite): a post-relu activation above 65504 x the layer's
                                     * power-of-two weight scale is clamped, where the FP32 path would carry
                                     * it on.  Conditioners of the reference's scale (|h| < 1e3) are far
                                     * inside; use FP32 for parameters that are not. */
#define FLOWCUDA_PRECISION_3PASS FLOWCUDA_PRECISION_3XTF32
#define FLOWCUDA_PRECISION_TF32 2   /* tcgen05 kind::tf32 single pass (reserved) */
#define FLOWCUDA_PRECISION_BF16X3 3 /* tcgen05 kind::f16 bf16, error-compensated 3-pass split (wide conditioners) */

typedef struct flowcuda_coupling_spec {
  int32_t event_dim;   /* D: flattened event size (2, or 784 for [28,28,1]) */
  int32_t num_layers;  /* flow layers L */
  int32_t num_hidden;  /* hidden Linear layers of the conditioner MLP (relu, activate_final) */
  int32_t hidden[FLOWCUDA_MAX_HIDDEN_LAYERS];
  int32_t transform;   /* FLOWCUDA_TRANSFORM_* */
  int32_t num_bins;    /* K (RQS only) */
  float range_min;     /* RQS only */
  float range_max;
  int32_t base;        /* FLOWCUDA_BASE_* */
  int32_t precision;   /* FLOWCUDA_PRECISION_* */
} flowcuda_coupling_spec;

/* Parameter layout (`params`): for each flow layer l = 0..L-1, for each Linear of the
 * conditioner in order (hidden..., final): W[in,out] row-major then b[out] — haiku's
 * `w`,`b` of mlp/~/linear_i and linear.  Layer l's mask is
 * (arange(D) % 2 == 1) for even l and its negation for odd l
 * (real_nvp_toy_densities.py:80-82,102); mask True = passed through and visible. */
int64_t flowcuda_coupling_param_count(const flowcuda_coupling_spec* spec);
size_t flowcuda_coupling_workspace_bytes(const flowcuda_coupling_spec* spec, int64_t n);

/* distrax.Transformed.log_prob: x[N,D] -> log_prob[N]; applies layer L-1's forward first
 * ... layer 0 last (Inverse(Chain(layers)), real_nvp_toy_densities.py:104), adds the base
 * log-density.  z (optional) receives the base-space point, log_det (optional) the
 * summed log-det. `layer_begin/layer_end` select a sub-chain (0, L for the full flow). */
int flowcuda_coupling_log_prob(const flowcuda_coupling_spec* spec, const float* params,
                               const float* x, float* log_prob, float* z, float* log_det, int64_t n,
                               int layer_begin, int layer_end, int with_base,
                               void* workspace, size_t workspace_bytes, void* stream);

/* distrax.Transformed.sample with the base noise given: z[N,D] -> x[N,D]; layer 0's
 * inverse first ... layer L-1 last (real_nvp_toy_densities.py:126-136).  log_det
 * (optional) receives the summed inverse log-det. */
int flowcuda_coupling_sample(const flowcuda_coupling_spec* spec, const float* params,
                             const float* z, float* x, float* log_det, int64_t n,
                             int layer_begin, int layer_end,
                             void* workspace, size_t workspace_bytes, void* stream);

/* Serving (precision BF16X3): repeated calls with UNCHANGED parameters need not re-pack the
 * weights into tile images.  pack_weights writes the images into `workspace` (they sit at
 * its start, independent of n; size the workspace for the largest batch to come); the
 * *_packed calls are log_prob / sample without that pass.  The caller promises that
 * `params`, `spec` and `workspace` are the ones given to pack_weights and that the
 * parameters have not changed since (biases are still read from `params`). */
int flowcuda_coupling_pack_weights(const flowcuda_coupling_spec* spec, const float* params,
                                   void* workspace, size_t workspace_bytes, void* stream);
int flowcuda_coupling_log_prob_packed(const flowcuda_coupling_spec* spec, const float* params,
                                      const float* x, float* log_prob, float* z, float* log_det,
                                      int64_t n, int layer_begin, int layer_end, int with_base,
                                      void* workspace, size_t workspace_bytes, void* stream);
int flowcuda_coupling_sample_packed(const flowcuda_coupling_spec* spec, const float* params,
                                    const float* z, float* x, float* log_det, int64_t n,
                                    int layer_begin, int layer_end,
                                    void* workspace, size_t workspace_bytes, void* stream);

/* value_and_grad of loss = -mean(Transformed.log_prob(x)) w.r.t. the flow parameters
 * (real_nvp_toy_densities.py:139-151,170; maf_toy_densities.py:140-152,171): hand-written
 * reverse mode, deterministic reductions (bit-identical run to run).  loss[1],
 * grad[param_count] (layout of `params`).
 *   event_dim 2, two hidden layers <= 64 (the toy flows): one fused FP32 kernel pair;
 *   any other shape with TRANSFORM_AFFINE / TRANSFORM_RQS (8 or 12 bins) — the
 *   maf_mnist.py:134-166 step: stored-activation reverse mode whose GEMMs run on FP32 CUDA
 *   cores (spec->precision = FP32) or on tcgen05 bf16x3 (spec->precision = BF16X3).  The bf16x3 engine runs
 *   the weight-gradient GEMMs on ONE internal non-blocking stream per device (created on first use, with four
 *   events, kept for the life of the process; joined back into `stream` before the call returns;
 *   FLOWCUDA_TRAIN_OVERLAP=0 disables it).  Concurrent calls from several host threads on one device are
 *   serialised while they enqueue. */
size_t flowcuda_coupling_nll_grad_workspace_bytes(const flowcuda_coupling_spec* spec, int64_t n);
int flowcuda_coupling_nll_grad(const flowcuda_coupling_spec* spec, const float* params, const float* x,
                               int64_t n, float* loss, float* grad, void* workspace,
                               size_t workspace_bytes, void* stream);

/* The same step with a per-flow-layer completion hook, for data-parallel training (the flat
 * gradient is all-reduced across ranks every step, 351 MB at the maf_mnist shape): the reverse mode finishes
 * the flow layers one after the other (layer 0 first), and a layer's gradient slice can be reduced while the
 * remaining layers are still being differentiated.  `on_layer` is called ON THE HOST, from inside this call,
 * one or more times per flow layer — with disjoint slices that together cover the layer's parameters (the
 * stored-activation engine reports a layer's last Linear, nine tenths of it, as soon as that Linear is done,
 * and the hidden Linears at the end of the layer) — each time right after the last piece of work that writes
 * grad[offset, offset + count) has been ENQUEUED; `stream` is the stream on which that slice is complete in stream order at that moment (the
 * library's internal side stream when the tensor-core engine overlaps the weight-gradient jobs, else the
 * caller's stream); at that point of `stream` the call has also finished READING the layer's parameters, so the
 * callee may update them in place (all-reduce, then the optimiser step of that layer, all under the remaining
 * layers' differentiation).  The callee typically records an event on `stream` and makes its own stream wait
 * for it; it must not synchronise or enqueue work that waits for the rest of this call.
 * on_layer == NULL: identical to flowcuda_coupling_nll_grad.  Results are the same bits either way. */
typedef void (*flowcuda_layer_grad_fn)(void* user, int layer, int64_t offset, int64_t count, void* stream);
int flowcuda_coupling_nll_grad_layers(const flowcuda_coupling_spec* spec, const float* params, const float* x,
                                      int64_t n, float* loss, float* grad, void* workspace,
                                      size_t workspace_bytes, void* stream, flowcuda_layer_grad_fn on_layer,
                                      void* user);

/* ------------------------------------------------------------------------------------
 * Row #8 — MADE-conditioned masked autoregressive affine flow (MAF), named by
 * BASELINE.json's north_star for the D=784 configs.  The reference has NO code for it
 * (src/maf_mwe.py:36-88 is commented out and is not MADE; the scripts import `tfb`,
 * whose MaskedAutoregressiveFlow/AutoregressiveNetwork this follows — SURVEY App. A.7):
 *   per flow layer a masked MLP [D -> hidden... -> 2D], relu; input degrees 1..D, hidden
 *   degrees max(1, ceil(k (D-1)/(H+1))) (k = 1..H), hidden mask deg_in <= deg_out, output
 *   mask deg_in < deg_out; output columns [shift_0..shift_{D-1}, log_scale_0..log_scale_{D-1}];
 *   odd layers see the dimensions in reversed order.
 * Parameter layout: per layer, per Linear: W[in,out] row-major then b[out] (UNMASKED
 * weights; the masks are applied by the library when it packs them).
 * ---------------------------------------------------------------------------------- */
typedef struct flowcuda_made_spec {
  int32_t event_dim;   /* D */
  int32_t num_layers;  /* flow layers L */
  int32_t num_hidden;  /* hidden Linears of each MADE */
  int32_t hidden[FLOWCUDA_MAX_HIDDEN_LAYERS];
  int32_t base;        /* FLOWCUDA_BASE_* */
  int32_t precision;   /* reverse mode (made_nll_grad): FLOWCUDA_PRECISION_FP32 or FLOWCUDA_PRECISION_BF16X3 */
} flowcuda_made_spec;

int64_t flowcuda_made_param_count(const flowcuda_made_spec* spec);
size_t flowcuda_made_workspace_bytes(const flowcuda_made_spec* spec, int64_t n);

/* Density pass x[N,D] -> log_prob[N]: layer L-1 first ... layer 0 last, per layer ONE
 * conditioner evaluation, u = (x - shift)·exp(-log_scale), log-det = -sum(log_scale); the
 * conditioner runs as tcgen05 bf16x3 GEMMs with fp32 accumulation (|d log_prob| <= 1e-3).
 * Optional outputs (may be NULL): u[N,D] the base-space point, log_det[N]. */
int flowcuda_made_log_prob(const flowcuda_made_spec* spec, const float* params, const float* x,
                           float* log_prob, float* u, float* log_det, int64_t n, int with_base,
                           void* workspace, size_t workspace_bytes, void* stream);

/* The same density pass with every product in FP32 on the CUDA cores (the engine of
 * flowcuda_made_nll_grad with precision FP32): the full-precision partner of the tensor-core pass
 * (|d log_prob| <= 1e-5 relative against the float64 restatement) — ~20x slower, for validation
 * and for callers whose parameters leave bf16x3's 1e-3 budget.  Normal base, two hidden layers (the
 * limits of made_nll_grad).  u (optional): base-space point. */
size_t flowcuda_made_log_prob_fp32_workspace_bytes(const flowcuda_made_spec* spec, int64_t n);
int flowcuda_made_log_prob_fp32(const flowcuda_made_spec* spec, const float* params, const float* x,
                                float* log_prob, float* u, int64_t n, void* workspace,
                                size_t workspace_bytes, void* stream);

/* Sampling with the base noise given, z[N,D] -> x[N,D]: layer 0 first ... layer L-1 last,
 * within a layer x_i = z_i·exp(log_scale_i(x_<i)) + shift_i(x_<i) sequentially over the D
 * positions — one persistent FP32 kernel with incrementally updated pre-activations.
 * Needs num_hidden == 2, hidden <= 512, event_dim <= 1024 (FLOWCUDA_EUNSUPPORTED otherwise). */
int flowcuda_made_sample(const flowcuda_made_spec* spec, const float* params, const float* z,
                         float* x, int64_t n, void* stream);

/* value_and_grad of loss = -mean(log_prob(x)) w.r.t. the MADE flow parameters (the training
 * step of src/maf_mnist.py:134-166 applied to the MADE flow): hand-written FP32 reverse mode over
 * stored activations, fixed summation order (bit-reproducible).  loss[1], grad[param_count]
 * (layout of `params`; masked-out weights get exactly 0).  Needs num_hidden == 2. */
/* Serving: as flowcuda_coupling_pack_weights / _log_prob_packed, for the MADE flow. */
int flowcuda_made_pack_weights(const flowcuda_made_spec* spec, const float* params, void* workspace,
                               size_t workspace_bytes, void* stream);
int flowcuda_made_log_prob_packed(const flowcuda_made_spec* spec, const float* params, const float* x,
                                  float* log_prob, float* u, float* log_det, int64_t n,
                                  int with_base, void* workspace, size_t workspace_bytes,
                                  void* stream);

size_t flowcuda_made_nll_grad_workspace_bytes(const flowcuda_made_spec* spec, int64_t n);
int flowcuda_made_nll_grad(const flowcuda_made_spec* spec, const float* params, const float* x,
                           int64_t n, float* loss, float* grad, void* workspace,
                           size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------
 * §8f-2/3 — the data step on the device.  Every output is a pure function of
 * (seed, element index): Philox4x32-10 with key = seed, counter = (index, stream tag), so a
 * batch is reproducible, stateless and shardable by index range (`offset`).  The
 * distributions are the reference's; the bit streams are not JAX's threefry.
 * ---------------------------------------------------------------------------------- */
#define FLOWCUDA_DENSITY_BANANA 0  /* src/densities/banana.py:31-41 */
#define FLOWCUDA_DENSITY_ENERGY1 1 /* src/densities/energies.py:47-66 */
#define FLOWCUDA_DENSITY_ENERGY2 2 /* src/densities/energies.py:69-82 */
#define FLOWCUDA_DENSITY_ENERGY3 3 /* src/densities/energies.py:85-104 */
#define FLOWCUDA_DENSITY_ENERGY4 4 /* src/densities/energies.py:107-127 */
#define FLOWCUDA_DENSITY_BLOBS 5   /* src/densities/gaussian_blobs.py:34-59 */

/* x[n,2] <- samples offset .. offset+n of the density's stream (replaces the host generators
 * make_dataset_<kind>(seed, batch_size, num_batches)).  num_blobs / r: gaussian blobs only. */
int flowcuda_density_sample(int kind, uint64_t seed, uint64_t offset, int num_blobs, float r,
                            float* x, int64_t n, void* stream);

/* out[i] = (bytes[i] + U[0,1)) / 256 — prepare_data of src/maf_mnist.py:101-106 without the
 * host uniform draw; `offset` = index of bytes[0] in the stream (multiple of 16). */
int flowcuda_dequantize_u8(const uint8_t* bytes, uint64_t seed, uint64_t offset, float* out,
                           int64_t count, void* stream);

/* Base noise for Transformed.sample(seed=...): N(0,1) (uniform = 0) or U[0,1) (uniform = 1);
 * offset a multiple of 4. */
int flowcuda_base_noise(int uniform, uint64_t seed, uint64_t offset, float* out, int64_t count,
                        void* stream);

/* §8f-4 — tutorial bijector OrthogonalProjection2D (src/rotation_flow.py:13-37, README.md:32-57):
 * y[n,2] = x[n,2] @ R, R = [[cos,-sin],[sin,cos]]^T (inverse != 0: x @ R^T).  The reference returns
 * the constant 1 as "log-det" (a bug; the true value is 0) — that choice lives in the host mirror. */
int flowcuda_rotate2d(const float* x, float cos_theta, float sin_theta, int inverse, float* y,
                      int64_t n, void* stream);

/* ------------------------------------------------------------------------------------
 * Rows #7, #9 — loss reduction and optimiser steps.
 * ---------------------------------------------------------------------------------- */

/* out[0] = -mean(log_prob[0..n)) scaled: out = scale * sum(log_prob), deterministic
 * (fixed-order two-stage, double accumulation). workspace: 1024 doubles. */
int flowcuda_sum_scaled(const float* v, int64_t n, double scale, float* out, void* workspace,
                        void* stream);

/* flowcuda_sum_scaled with NO workspace argument and NO allocation.  This entry point has no
 * counterpart in the reference.  Same kernels,
 * same grid rule and therefore the same bits as flowcuda_sum_scaled, but the 1024 stage-1 partials
 * (8 KB) live in a file-scope `static __device__ __align__(256)` array of the library, whose device
 * address is resolved once per device (cudaGetSymbolAddress, that device current) and cached.  It calls
 * no cudaMalloc / cudaFree / allocating library routine.  v, out: caller's device buffers.
 * Limit: n <= 2^26 elements (67 108 864 floats, 256 MB of input); larger n returns
 * FLOWCUDA_EINVAL with a message naming the limit — use flowcuda_sum_scaled there.
 * The array is shared by every call on a device and keeps the last call's partials (the slots a call
 * uses are cleared first): calls that overlap on one device (two streams, two host threads) RACE on it;
 * serialise them on one stream or with an event. */
int flowcuda_sum_scaled_static(const float* v, int64_t n, double scale, float* out, void* stream);

/* p <- p - lr*g (real_nvp.py:298). */
int flowcuda_sgd_step(float* p, const float* g, int64_t count, float lr, float grad_scale, void* stream);

/* optax.adam(lr) + apply_updates (real_nvp_toy_densities.py:162,171-172): b1,b2,eps as
 * given; step is 1-based; grad_scale multiplies g first (1/world_size after a sum
 * all-reduce). */
int flowcuda_adam_step(float* p, const float* g, float* m, float* v, int64_t count, int64_t step,
                       float lr, float b1, float b2, float eps, float grad_scale, void* stream);

/* ------------------------------------------------------------------------------------
 * Scalar bijectors and base densities as stand-alone elementwise kernels — what the host mirror's
 * `distrax.ScalarAffine`, `tfb.Scale`, `tfb.Shift`, `tfb.SoftClip`, `distrax.RationalQuadraticSpline`,
 * `Normal` / `Uniform` / `Independent.log_prob` run when a user composes them by hand
 * (real_nvp_toy_densities.py:84-86, maf_toy_densities.py:86-87, bflow.py:28-34,44,82-88,133-142).
 * Element i of the flattened input uses parameter row (i / inner) % rows: `inner` trailing elements
 * share a row ([N,1] parameters broadcast over D), rows repeat along the leading axes ([D] parameters).
 * Outputs y / ld may be NULL when not wanted.
 * ---------------------------------------------------------------------------------- */
/* y = x*scale + shift, ld = log|scale|; inverse: x = (y - shift)/scale, ld = -log|scale|.  `scale` holds
 * log-scales when scale_is_log != 0.  NULL shift = 0; NULL scale = 1. */
int flowcuda_affine_elementwise(const float* x, const float* shift, const float* scale, int scale_is_log, int inverse,
                                float* y, float* ld, int64_t n, int64_t rows, int64_t inner, void* stream);
/* tfb.SoftClip(low, high, hinge_softness). */
int flowcuda_softclip_elementwise(const float* x, float low, float high, float hinge_softness, int inverse, float* y,
                                  float* ld, int64_t n, void* stream);
/* distrax.RationalQuadraticSpline(params[rows, 3*num_bins+1], range_min, range_max). */
int flowcuda_rqs_elementwise(const float* x, const float* params, int num_bins, float range_min, float range_max,
                             int inverse, float* y, float* ld, int64_t n, int64_t rows, int64_t inner, void* stream);
/* kind 0: Normal(loc=a, scale=b).log_prob; kind 1: Uniform(low=a, high=b).log_prob (-inf outside [low, high]);
 * NULL a / b = 0 / 1. */
int flowcuda_base_log_prob(const float* x, const float* a, const float* b, int kind, float* out, int64_t n,
                           int64_t rows, int64_t inner, void* stream);
/* out[r] = sum_c x[r, c], fixed summation order (distrax.Independent / Block: sum over the event dims). */
int flowcuda_sum_last(const float* x, float* out, int64_t rows, int64_t cols, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FLOWCUDA_H_ */
