// coupling_api.cu — C entry points for the MaskedCoupling flows; dispatches to the fused
// event_dim=2 kernels (coupling_toy*.cu) or the generic dense path (coupling_dense*.cu).
#include "common.cuh"

namespace flowcuda {
int coupling_toy_dispatch(const flowcuda_coupling_spec* sp, const float* params, const float* in, float* out_v,
                          float* out_lp, int64_t n, int layer_begin, int layer_end, int with_base, bool inverse,
                          cudaStream_t s);
int coupling_toy_tc_dispatch(const flowcuda_coupling_spec* sp, const float* params, const float* in, float* out_v,
                             float* out_lp, int64_t n, int layer_begin, int layer_end, int with_base, bool inverse,
                             void* ws, size_t wsb, cudaStream_t s);
size_t coupling_toy_tc_workspace_bytes(const flowcuda_coupling_spec* sp);
bool coupling_dense_tc_supported(const flowcuda_coupling_spec* sp);
size_t coupling_dense_tc_workspace_bytes(const flowcuda_coupling_spec* sp, int64_t n);
int coupling_dense_tc_run(const flowcuda_coupling_spec* sp, const float* params, const float* in, float* out_v,
                          float* out_lp, float* out_ld, int64_t n, int layer_begin, int layer_end, int with_base,
                          bool inverse, void* workspace, size_t workspace_bytes, cudaStream_t s, int mode);
int64_t coupling_param_count(const flowcuda_coupling_spec* sp);
size_t coupling_dense_workspace_bytes(const flowcuda_coupling_spec* sp, int64_t n);
int coupling_dense_run(const flowcuda_coupling_spec* sp, const float* params, const float* in, float* out_v,
                       float* out_lp, float* out_ld, int64_t n, int layer_begin, int layer_end, int with_base,
                       bool inverse, void* workspace, size_t workspace_bytes, cudaStream_t s);

static int check_spec(const flowcuda_coupling_spec* sp) {
  FC_REQUIRE(sp, "coupling: NULL spec");
  FC_REQUIRE(sp->event_dim >= 2, "coupling: event_dim=%d must be >= 2", sp->event_dim);
  FC_REQUIRE(sp->num_layers >= 1 && sp->num_layers <= 1024, "coupling: num_layers=%d", sp->num_layers);
  FC_REQUIRE(sp->num_hidden >= 1 && sp->num_hidden <= FLOWCUDA_MAX_HIDDEN_LAYERS, "coupling: num_hidden=%d outside [1,%d]",
             sp->num_hidden, FLOWCUDA_MAX_HIDDEN_LAYERS);
  for (int i = 0; i < sp->num_hidden; ++i) FC_REQUIRE(sp->hidden[i] >= 1, "coupling: hidden[%d]=%d", i, sp->hidden[i]);
  FC_REQUIRE(sp->transform >= 0 && sp->transform <= 2, "coupling: unknown transform %d", sp->transform);
  if (sp->transform == FLOWCUDA_TRANSFORM_RQS) {
    FC_REQUIRE(sp->num_bins >= 1 && sp->num_bins <= 42, "coupling: num_bins=%d outside [1,42]", sp->num_bins);
    FC_REQUIRE(sp->range_max > sp->range_min, "coupling: empty spline range");
  }
  FC_REQUIRE(sp->base == FLOWCUDA_BASE_NORMAL || sp->base == FLOWCUDA_BASE_UNIFORM01, "coupling: unknown base %d", sp->base);
  FC_REQUIRE(sp->precision >= 0 && sp->precision <= 3 && sp->precision != FLOWCUDA_PRECISION_TF32, "coupling: unknown precision %d", sp->precision);
  return FLOWCUDA_OK;
}

static bool toy_eligible(const flowcuda_coupling_spec* sp) {
  return sp->event_dim == 2 && sp->num_hidden == 2 && sp->hidden[0] <= 64 && sp->hidden[1] <= 64 &&
         (sp->transform != FLOWCUDA_TRANSFORM_RQS || sp->num_bins == 12 || sp->num_bins == 8);
}

static int run(const flowcuda_coupling_spec* sp, const float* params, const float* in, float* out_v, float* out_lp,
               float* out_ld, int64_t n, int lb, int le, int with_base, bool inverse, void* ws, size_t wsb,
               cudaStream_t s, int mode = 0) {
  int rc = check_spec(sp);
  if (rc) return rc;
  FC_REQUIRE(n >= 0, "coupling: negative n");
  if (n == 0) return FLOWCUDA_OK;
  FC_REQUIRE(params && in, "coupling: NULL params or input");
  FC_REQUIRE(lb >= 0 && le <= sp->num_layers && lb <= le, "coupling: layer range [%d,%d) outside [0,%d]", lb, le, sp->num_layers);
  if (sp->precision == FLOWCUDA_PRECISION_BF16X3)
    return coupling_dense_tc_run(sp, params, in, out_v, out_lp, out_ld, n, lb, le, with_base, inverse, ws, wsb, s, mode);
  FC_REQUIRE(mode == 0, "coupling: packed weight images exist for precision BF16X3 only");
  if (sp->precision != FLOWCUDA_PRECISION_FP32) {
    FC_REQUIRE(sp->precision == FLOWCUDA_PRECISION_3XTF32 && toy_eligible(sp) && le > lb,
               "coupling: precision %d is not available for this flow shape (3xTF32 needs event_dim 2, two hidden layers <= 64)",
               sp->precision);
    FC_REQUIRE((((uintptr_t)in | (uintptr_t)out_v) & 7) == 0, "coupling: [N,2] buffers must be 8-byte aligned");
    if (out_lp && out_ld) {
      rc = coupling_toy_tc_dispatch(sp, params, in, nullptr, out_ld, n, lb, le, 0, inverse, ws, wsb, s);
      if (rc) return rc;
      return coupling_toy_tc_dispatch(sp, params, in, out_v, out_lp, n, lb, le, with_base, inverse, ws, wsb, s);
    }
    if (out_ld) return coupling_toy_tc_dispatch(sp, params, in, out_v, out_ld, n, lb, le, 0, inverse, ws, wsb, s);
    return coupling_toy_tc_dispatch(sp, params, in, out_v, out_lp, n, lb, le, with_base, inverse, ws, wsb, s);
  }
  if (toy_eligible(sp)) {
    FC_REQUIRE((((uintptr_t)in | (uintptr_t)out_v) & 7) == 0, "coupling: [N,2] buffers must be 8-byte aligned");
    if (out_lp && out_ld) {
      // two passes would be wasteful: log_det = log_prob without the base term
      rc = coupling_toy_dispatch(sp, params, in, nullptr, out_ld, n, lb, le, 0, inverse, s);
      if (rc) return rc;
      return coupling_toy_dispatch(sp, params, in, out_v, out_lp, n, lb, le, with_base, inverse, s);
    }
    if (out_ld) return coupling_toy_dispatch(sp, params, in, out_v, out_ld, n, lb, le, 0, inverse, s);
    return coupling_toy_dispatch(sp, params, in, out_v, out_lp, n, lb, le, with_base, inverse, s);
  }
  return coupling_dense_run(sp, params, in, out_v, out_lp, out_ld, n, lb, le, with_base, inverse, ws, wsb, s);
}

}  // namespace flowcuda

using namespace flowcuda;

extern "C" {

int64_t flowcuda_coupling_param_count(const flowcuda_coupling_spec* spec) {
  if (check_spec(spec)) return -1;
  return coupling_param_count(spec);
}

size_t flowcuda_coupling_workspace_bytes(const flowcuda_coupling_spec* spec, int64_t n) {
  if (check_spec(spec) || n < 0) return 0;
  if (toy_eligible(spec) && spec->precision == FLOWCUDA_PRECISION_FP32) return 0;
  if (toy_eligible(spec) && spec->precision == FLOWCUDA_PRECISION_3XTF32) return coupling_toy_tc_workspace_bytes(spec);
  if (spec->precision == FLOWCUDA_PRECISION_BF16X3) return coupling_dense_tc_workspace_bytes(spec, n);
  return coupling_dense_workspace_bytes(spec, n);
}

int flowcuda_coupling_log_prob(const flowcuda_coupling_spec* spec, const float* params, const float* x, float* log_prob,
                               float* z, float* log_det, int64_t n, int layer_begin, int layer_end, int with_base,
                               void* workspace, size_t workspace_bytes, void* stream) {
  if (n == 0) return FLOWCUDA_OK;
  FC_REQUIRE(log_prob || z || log_det, "coupling_log_prob: no output requested");
  return run(spec, params, x, z, log_prob, log_det, n, layer_begin, layer_end, with_base, false, workspace,
             workspace_bytes, (cudaStream_t)stream);
}

int flowcuda_coupling_sample(const flowcuda_coupling_spec* spec, const float* params, const float* z, float* x,
                             float* log_det, int64_t n, int layer_begin, int layer_end, void* workspace,
                             size_t workspace_bytes, void* stream) {
  if (n == 0) return FLOWCUDA_OK;
  FC_REQUIRE(x || log_det, "coupling_sample: no output requested");
  return run(spec, params, z, x, nullptr, log_det, n, layer_begin, layer_end, 0, true, workspace, workspace_bytes,
             (cudaStream_t)stream);
}

int flowcuda_coupling_pack_weights(const flowcuda_coupling_spec* spec, const float* params, void* workspace,
                                   size_t workspace_bytes, void* stream) {
  int rc = check_spec(spec);
  if (rc) return rc;
  FC_REQUIRE(params && workspace, "coupling_pack_weights: NULL params or workspace");
  FC_REQUIRE(spec->precision == FLOWCUDA_PRECISION_BF16X3, "coupling_pack_weights: precision BF16X3 only");
  return coupling_dense_tc_run(spec, params, nullptr, nullptr, nullptr, nullptr, 0, 0, spec->num_layers, 0, false, workspace,
                               workspace_bytes, (cudaStream_t)stream, 2);
}

int flowcuda_coupling_log_prob_packed(const flowcuda_coupling_spec* spec, const float* params, const float* x,
                                      float* log_prob, float* z, float* log_det, int64_t n, int layer_begin,
                                      int layer_end, int with_base, void* workspace, size_t workspace_bytes,
                                      void* stream) {
  if (n == 0) return FLOWCUDA_OK;
  FC_REQUIRE(log_prob || z || log_det, "coupling_log_prob_packed: no output requested");
  return run(spec, params, x, z, log_prob, log_det, n, layer_begin, layer_end, with_base, false, workspace,
             workspace_bytes, (cudaStream_t)stream, 1);
}

int flowcuda_coupling_sample_packed(const flowcuda_coupling_spec* spec, const float* params, const float* z, float* x,
                                    float* log_det, int64_t n, int layer_begin, int layer_end, void* workspace,
                                    size_t workspace_bytes, void* stream) {
  if (n == 0) return FLOWCUDA_OK;
  FC_REQUIRE(x || log_det, "coupling_sample_packed: no output requested");
  return run(spec, params, z, x, nullptr, log_det, n, layer_begin, layer_end, 0, true, workspace, workspace_bytes,
             (cudaStream_t)stream, 1);
}

}  // extern "C"
