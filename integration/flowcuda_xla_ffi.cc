// flowcuda_xla_ffi.cc — XLA FFI (jax.ffi) adapters over the C ABI of libflowcuda.so (include/flowcuda.h).
//
// UNVERIFIED: JAX, jaxlib and the XLA FFI headers exist on neither the build container nor the GPU box of this
// project, so this translation unit has never been compiled against xla/ffi/api/ffi.h or run.
// It is written from the published XLA FFI API (xla::ffi::Ffi::Bind, Buffer<F32>, PlatformStream<cudaStream_t>)
// and is what a maintainer of the reference adds on a machine that has JAX:
//
//   g++ -std=c++17 -shared -fPIC -I$(python -c "import jax.ffi; print(jax.ffi.include_dir())") -I include \
//       -I /usr/local/cuda/include integration/flowcuda_xla_ffi.cc -L normalizing_flows_b200 -lflowcuda \
//       -o integration/libflowcuda_xla_ffi.so
//
// `python -m normalizing_flows_b200.build` compiles it automatically when it finds the header (XLA_FFI_INCLUDE or an
// importable jax); without the header the file preprocesses to nothing, which is the only state testable here
// (tests/test_cabi.py::test_xla_ffi_adapter_is_inert_without_the_header).
//
// Handlers (each binds the entry point of the same name; attributes are the scalar arguments of the C call):
//   flowcuda_realnvp_log_prob_ffi      log_prob_nvp_chains               src/real_nvp.py:268-283
//   flowcuda_realnvp_nll_grad_ffi      loss + value_and_grad             src/real_nvp.py:292-297
//   flowcuda_coupling_log_prob_ffi     Transformed.log_prob              src/maf_mnist.py:52-118 and the toy scripts
//   flowcuda_coupling_nll_grad_ffi     jax.value_and_grad(loss_fn)       src/maf_mnist.py:134-166
//   flowcuda_bernstein_fwd_ffi         BernsteinBijector.forward_and_log_det   src/BernsteinBijector.py:81-93
//   flowcuda_adam_step_ffi             optax.adam + apply_updates        src/real_nvp_toy_densities.py:162-172
// Workspaces are extra RESULT buffers (XLA owns all memory): the Python side sizes them with
// flowcuda_*_workspace_bytes through ctypes and declares them as uint8 results of ffi_call.
#if defined(__has_include)
#if __has_include("xla/ffi/api/ffi.h")
#define FLOWCUDA_HAVE_XLA_FFI 1
#endif
#endif

#ifdef FLOWCUDA_HAVE_XLA_FFI
#include <cuda_runtime_api.h>

#include <cstdint>

#include "flowcuda.h"
#include "xla/ffi/api/ffi.h"

namespace ffi = xla::ffi;

static ffi::Error Status(int rc) {
  return rc == FLOWCUDA_OK ? ffi::Error::Success() : ffi::Error::Internal(flowcuda_last_error());
}

static flowcuda_coupling_spec MakeSpec(int64_t event_dim, int64_t num_layers, ffi::Span<const int64_t> hidden, int64_t transform,
                                       int64_t num_bins, float range_min, float range_max, int64_t base, int64_t precision) {
  flowcuda_coupling_spec sp = {};
  sp.event_dim = (int32_t)event_dim;
  sp.num_layers = (int32_t)num_layers;
  sp.num_hidden = (int32_t)hidden.size();
  for (size_t i = 0; i < hidden.size() && i < FLOWCUDA_MAX_HIDDEN_LAYERS; ++i) sp.hidden[i] = (int32_t)hidden[i];
  sp.transform = (int32_t)transform;
  sp.num_bins = (int32_t)num_bins;
  sp.range_min = range_min;
  sp.range_max = range_max;
  sp.base = (int32_t)base;
  sp.precision = (int32_t)precision;
  return sp;
}

// ---- row #1 ---------------------------------------------------------------------------------------------------
static ffi::Error RealNvpLogProb(cudaStream_t s, ffi::Buffer<ffi::F32> params, ffi::Buffer<ffi::F32> y,
                                 ffi::ResultBuffer<ffi::F32> log_prob, int64_t n_layers, int64_t hidden, int64_t flip_mask) {
  return Status(flowcuda_realnvp_log_prob(params.typed_data(), (int)n_layers, (int)hidden, (uint64_t)flip_mask, y.typed_data(),
                                          log_prob->typed_data(), nullptr, (int64_t)y.dimensions()[0], s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(flowcuda_realnvp_log_prob_ffi, RealNvpLogProb,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Attr<int64_t>("n_layers")
                                  .Attr<int64_t>("hidden")
                                  .Attr<int64_t>("flip_mask"));

static ffi::Error RealNvpNllGrad(cudaStream_t s, ffi::Buffer<ffi::F32> params, ffi::Buffer<ffi::F32> y,
                                 ffi::ResultBuffer<ffi::F32> loss, ffi::ResultBuffer<ffi::F32> grad,
                                 ffi::ResultBuffer<ffi::U8> workspace, int64_t n_layers, int64_t hidden, int64_t flip_mask) {
  return Status(flowcuda_realnvp_nll_grad(params.typed_data(), (int)n_layers, (int)hidden, (uint64_t)flip_mask, y.typed_data(),
                                          (int64_t)y.dimensions()[0], loss->typed_data(), grad->typed_data(), 0.f, nullptr,
                                          workspace->typed_data(), workspace->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(flowcuda_realnvp_nll_grad_ffi, RealNvpNllGrad,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::U8>>()
                                  .Attr<int64_t>("n_layers")
                                  .Attr<int64_t>("hidden")
                                  .Attr<int64_t>("flip_mask"));

// ---- rows #2-#4, #7 ---------------------------------------------------------------------------------------------
static ffi::Error CouplingLogProb(cudaStream_t s, ffi::Buffer<ffi::F32> params, ffi::Buffer<ffi::F32> x,
                                  ffi::ResultBuffer<ffi::F32> log_prob, ffi::ResultBuffer<ffi::U8> workspace, int64_t event_dim,
                                  int64_t num_layers, ffi::Span<const int64_t> hidden, int64_t transform, int64_t num_bins,
                                  float range_min, float range_max, int64_t base, int64_t precision) {
  const flowcuda_coupling_spec sp = MakeSpec(event_dim, num_layers, hidden, transform, num_bins, range_min, range_max, base, precision);
  return Status(flowcuda_coupling_log_prob(&sp, params.typed_data(), x.typed_data(), log_prob->typed_data(), nullptr, nullptr,
                                           (int64_t)x.dimensions()[0], 0, (int)num_layers, 1, workspace->typed_data(),
                                           workspace->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(flowcuda_coupling_log_prob_ffi, CouplingLogProb,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::U8>>()
                                  .Attr<int64_t>("event_dim")
                                  .Attr<int64_t>("num_layers")
                                  .Attr<ffi::Span<const int64_t>>("hidden")
                                  .Attr<int64_t>("transform")
                                  .Attr<int64_t>("num_bins")
                                  .Attr<float>("range_min")
                                  .Attr<float>("range_max")
                                  .Attr<int64_t>("base")
                                  .Attr<int64_t>("precision"));

static ffi::Error CouplingNllGrad(cudaStream_t s, ffi::Buffer<ffi::F32> params, ffi::Buffer<ffi::F32> x, ffi::ResultBuffer<ffi::F32> loss,
                                  ffi::ResultBuffer<ffi::F32> grad, ffi::ResultBuffer<ffi::U8> workspace, int64_t event_dim,
                                  int64_t num_layers, ffi::Span<const int64_t> hidden, int64_t transform, int64_t num_bins,
                                  float range_min, float range_max, int64_t base, int64_t precision) {
  const flowcuda_coupling_spec sp = MakeSpec(event_dim, num_layers, hidden, transform, num_bins, range_min, range_max, base, precision);
  return Status(flowcuda_coupling_nll_grad(&sp, params.typed_data(), x.typed_data(), (int64_t)x.dimensions()[0], loss->typed_data(),
                                           grad->typed_data(), workspace->typed_data(), workspace->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(flowcuda_coupling_nll_grad_ffi, CouplingNllGrad,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::U8>>()
                                  .Attr<int64_t>("event_dim")
                                  .Attr<int64_t>("num_layers")
                                  .Attr<ffi::Span<const int64_t>>("hidden")
                                  .Attr<int64_t>("transform")
                                  .Attr<int64_t>("num_bins")
                                  .Attr<float>("range_min")
                                  .Attr<float>("range_max")
                                  .Attr<int64_t>("base")
                                  .Attr<int64_t>("precision"));

// ---- row #5 ---------------------------------------------------------------------------------------------------
static ffi::Error BernsteinFwd(cudaStream_t s, ffi::Buffer<ffi::F32> theta, ffi::Buffer<ffi::F32> x, ffi::ResultBuffer<ffi::F32> y,
                               ffi::ResultBuffer<ffi::F32> log_det, int64_t order, int64_t theta_stride) {
  return Status(flowcuda_bernstein_forward_and_log_det(theta.typed_data(), (int)order, theta_stride, x.typed_data(), y->typed_data(),
                                                       log_det->typed_data(), (int64_t)x.element_count(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(flowcuda_bernstein_fwd_ffi, BernsteinFwd,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Attr<int64_t>("order")
                                  .Attr<int64_t>("theta_stride"));

// ---- row #9 ---------------------------------------------------------------------------------------------------
// In-place update: bind with input_output_aliases={0: 0, 2: 1, 3: 2} so that XLA hands the same buffers back.
static ffi::Error AdamStep(cudaStream_t s, ffi::Buffer<ffi::F32> p, ffi::Buffer<ffi::F32> g, ffi::Buffer<ffi::F32> m, ffi::Buffer<ffi::F32> v,
                           ffi::ResultBuffer<ffi::F32> p_out, ffi::ResultBuffer<ffi::F32> m_out, ffi::ResultBuffer<ffi::F32> v_out,
                           int64_t step, float lr, float b1, float b2, float eps, float grad_scale) {
  if (p_out->typed_data() != p.typed_data() || m_out->typed_data() != m.typed_data() || v_out->typed_data() != v.typed_data())
    return ffi::Error::InvalidArgument("flowcuda_adam_step_ffi: bind with input_output_aliases (the update is in place)");
  return Status(flowcuda_adam_step(p_out->typed_data(), g.typed_data(), m_out->typed_data(), v_out->typed_data(),
                                   (int64_t)p.element_count(), step, lr, b1, b2, eps, grad_scale, s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(flowcuda_adam_step_ffi, AdamStep,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Attr<int64_t>("step")
                                  .Attr<float>("lr")
                                  .Attr<float>("b1")
                                  .Attr<float>("b2")
                                  .Attr<float>("eps")
                                  .Attr<float>("grad_scale"));
#endif  // FLOWCUDA_HAVE_XLA_FFI
