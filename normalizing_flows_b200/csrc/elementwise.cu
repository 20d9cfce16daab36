// elementwise.cu — the scalar bijectors and base densities of the distrax / TFP surface as stand-alone
// elementwise kernels: distrax.ScalarAffine (real_nvp_toy_densities.py:84-86), tfb.Scale / tfb.Shift / tfb.SoftClip
// (bflow.py:28-31), distrax.RationalQuadraticSpline (maf_toy_densities.py:86-87, maf_mnist.py:65-66) and
// Normal / Uniform log_prob with distrax.Independent's event sum (bflow.py:44,87,108,134-142).
//
// The fused chain kernels are the hot path; these exist so that the host mirror's bijector CLASSES compose the way the
// reference composes them (a user-written Chain, a bijector_fn called on its own, Transformed over any bijector).
// Parameter broadcasting: element i of the flattened input uses parameter row (i / inner) % rows — `inner` trailing
// elements share a row (an [N,1] parameter broadcast over D), rows repeat along leading axes (a [D] parameter).
#include "transforms.cuh"

namespace flowcuda {

__device__ __forceinline__ int64_t ew_row(int64_t i, int64_t rows, int64_t inner) { return rows == 1 ? 0 : (i / inner) % rows; }

// y = x*scale + shift, ld = log|scale| (inverse: x = (y - shift)/scale, ld = -log|scale|); `sc` holds the log-scale
// (is_log) or the scale itself; NULL shift = 0, NULL sc = identity scale.
__global__ void __launch_bounds__(256) ew_affine_kernel(const float* __restrict__ x, const float* __restrict__ shift,
                                                        const float* __restrict__ sc, int is_log, int inverse,
                                                        float* __restrict__ y, float* __restrict__ ld, int64_t n, int64_t rows,
                                                        int64_t inner) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = ew_row(i, rows, inner);
    const float b = shift ? shift[r] : 0.f;
    float s = 1.f, l = 0.f;
    if (sc) {
      if (is_log) {
        l = sc[r];
        s = expf(l);
      } else {
        s = sc[r];
        l = logf(fabsf(s));
      }
    }
    const float v = x[i];
    if (y) y[i] = inverse ? (v - b) / s : fmaf(v, s, b);
    if (ld) ld[i] = inverse ? -l : l;
  }
}

// tfb.SoftClip(low, high, hinge_softness=c): y = high - k*sp(high - low - sp(x - low)), k = (high-low)/sp(high-low),
// sp(u) = c*softplus(u/c); dy/dx = k*sigmoid(u/c)*sigmoid((x-low)/c) with u = high - low - sp(x - low).
__device__ __forceinline__ float ew_log_sigmoid(float t) { return -softplus_f(-t); }
__device__ __forceinline__ float ew_softplus_inv(float v, float c) {
  // u with c*softplus(u/c) = v:  u = v + c*log(-expm1(-v/c))
  return v + c * logf(-expm1f(-v / c));
}
__global__ void __launch_bounds__(256) ew_softclip_kernel(const float* __restrict__ x, float low, float high, float c, int inverse,
                                                          float* __restrict__ y, float* __restrict__ ld, int64_t n) {
  const float r = high - low, rc = 1.f / c;
  const float k = r / (c * softplus_f(r * rc));
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float v = x[i];
    if (inverse) {
      const float u = ew_softplus_inv((high - v) / k, c);
      const float xi = low + ew_softplus_inv(r - u, c);
      if (y) y[i] = xi;
      v = xi;
    }
    const float inner = c * softplus_f((v - low) * rc);
    const float u = r - inner;
    if (!inverse && y) y[i] = high - k * (c * softplus_f(u * rc));
    if (ld) {
      const float f = logf(k) + ew_log_sigmoid(u * rc) + ew_log_sigmoid((v - low) * rc);
      ld[i] = inverse ? -f : f;
    }
  }
}

// distrax.RationalQuadraticSpline(params[rows, 3K+1], range_min, range_max) applied elementwise
template <bool INV>
__global__ void __launch_bounds__(128) ew_rqs_kernel(const float* __restrict__ x, const float* __restrict__ params, int K, float lo,
                                                     float hi, float* __restrict__ y, float* __restrict__ ld, int64_t n, int64_t rows,
                                                     int64_t inner) {
  const int P = 3 * K + 1;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float* p = params + ew_row(i, rows, inner) * P;
    float o, l;
    rqs_apply<INV, 0>([&](int j) { return p[j]; }, K, lo, hi, x[i], o, l);
    if (y) y[i] = o;
    if (ld) ld[i] = l;
  }
}

// Normal(loc, scale).log_prob / Uniform(low, high).log_prob, elementwise (kind 0 / 1; a = loc | low, b = scale | high)
__global__ void __launch_bounds__(256) ew_base_log_prob_kernel(const float* __restrict__ x, const float* __restrict__ a,
                                                               const float* __restrict__ b, int kind, float* __restrict__ out,
                                                               int64_t n, int64_t rows, int64_t inner) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = ew_row(i, rows, inner);
    const float v = x[i];
    if (kind == 0) {
      const float mu = a ? a[r] : 0.f, sg = b ? b[r] : 1.f;
      const float z = (v - mu) / sg;
      out[i] = fmaf(-0.5f * z, z, -kHalfLog2Pi) - (b ? logf(sg) : 0.f);
    } else {
      const float lo = a ? a[r] : 0.f, hi = b ? b[r] : 1.f;
      out[i] = (v < lo || v > hi) ? -INFINITY : -logf(hi - lo);
    }
  }
}

// out[r] = sum_c x[r, c] in a fixed order (one warp per row: lane-strided partial sums, then the shuffle tree)
__global__ void __launch_bounds__(256) ew_sum_last_kernel(const float* __restrict__ x, float* __restrict__ out, int64_t rows, int64_t cols) {
  const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (r >= rows) return;
  float s = 0.f;
  for (int64_t c = lane; c < cols; c += 32) s += x[r * cols + c];
  s = warp_sum(s);
  if (lane == 0) out[r] = s;
}

static inline int ew_grid(int64_t n, int threads) {
  int64_t b = (n + threads - 1) / threads;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (b > cap) b = cap;
  return (int)(b < 1 ? 1 : b);
}

}  // namespace flowcuda

using namespace flowcuda;

extern "C" {

int flowcuda_affine_elementwise(const float* x, const float* shift, const float* scale, int scale_is_log, int inverse, float* y,
                                float* ld, int64_t n, int64_t rows, int64_t inner, void* stream) {
  FC_REQUIRE(n >= 0 && rows >= 1 && inner >= 1, "affine_elementwise: n=%lld rows=%lld inner=%lld", (long long)n, (long long)rows,
             (long long)inner);
  FC_REQUIRE(n == 0 || (x && (y || ld)), "affine_elementwise: NULL pointer");
  if (n == 0) return FLOWCUDA_OK;
  ew_affine_kernel<<<ew_grid(n, 256), 256, 0, (cudaStream_t)stream>>>(x, shift, scale, scale_is_log, inverse, y, ld, n, rows, inner);
  FC_LAUNCH_CHECK("ew_affine_kernel");
  return FLOWCUDA_OK;
}

int flowcuda_softclip_elementwise(const float* x, float low, float high, float hinge_softness, int inverse, float* y, float* ld,
                                  int64_t n, void* stream) {
  FC_REQUIRE(n >= 0 && high > low && hinge_softness > 0.f, "softclip_elementwise: need high > low and hinge_softness > 0");
  FC_REQUIRE(n == 0 || (x && (y || ld)), "softclip_elementwise: NULL pointer");
  if (n == 0) return FLOWCUDA_OK;
  ew_softclip_kernel<<<ew_grid(n, 256), 256, 0, (cudaStream_t)stream>>>(x, low, high, hinge_softness, inverse, y, ld, n);
  FC_LAUNCH_CHECK("ew_softclip_kernel");
  return FLOWCUDA_OK;
}

int flowcuda_rqs_elementwise(const float* x, const float* params, int num_bins, float range_min, float range_max, int inverse,
                             float* y, float* ld, int64_t n, int64_t rows, int64_t inner, void* stream) {
  FC_REQUIRE(n >= 0 && rows >= 1 && inner >= 1, "rqs_elementwise: bad sizes");
  FC_REQUIRE(num_bins >= 1 && num_bins <= 64, "rqs_elementwise: num_bins=%d outside [1,64]", num_bins);
  FC_REQUIRE(range_min < range_max, "rqs_elementwise: `range_min` must be less than `range_max`");
  FC_REQUIRE(n == 0 || (x && params && (y || ld)), "rqs_elementwise: NULL pointer");
  if (n == 0) return FLOWCUDA_OK;
  if (inverse)
    ew_rqs_kernel<true><<<ew_grid(n, 128), 128, 0, (cudaStream_t)stream>>>(x, params, num_bins, range_min, range_max, y, ld, n, rows, inner);
  else
    ew_rqs_kernel<false><<<ew_grid(n, 128), 128, 0, (cudaStream_t)stream>>>(x, params, num_bins, range_min, range_max, y, ld, n, rows, inner);
  FC_LAUNCH_CHECK("ew_rqs_kernel");
  return FLOWCUDA_OK;
}

int flowcuda_base_log_prob(const float* x, const float* a, const float* b, int kind, float* out, int64_t n, int64_t rows,
                           int64_t inner, void* stream) {
  FC_REQUIRE(n >= 0 && rows >= 1 && inner >= 1, "base_log_prob: bad sizes");
  FC_REQUIRE(kind == 0 || kind == 1, "base_log_prob: kind=%d (0 Normal, 1 Uniform)", kind);
  FC_REQUIRE(n == 0 || (x && out), "base_log_prob: NULL pointer");
  if (n == 0) return FLOWCUDA_OK;
  ew_base_log_prob_kernel<<<ew_grid(n, 256), 256, 0, (cudaStream_t)stream>>>(x, a, b, kind, out, n, rows, inner);
  FC_LAUNCH_CHECK("ew_base_log_prob_kernel");
  return FLOWCUDA_OK;
}

int flowcuda_sum_last(const float* x, float* out, int64_t rows, int64_t cols, void* stream) {
  FC_REQUIRE(rows >= 0 && cols >= 1, "sum_last: bad sizes");
  FC_REQUIRE(rows == 0 || (x && out), "sum_last: NULL pointer");
  if (rows == 0) return FLOWCUDA_OK;
  ew_sum_last_kernel<<<(int)((rows * 32 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(x, out, rows, cols);
  FC_LAUNCH_CHECK("ew_sum_last_kernel");
  return FLOWCUDA_OK;
}

}  // extern "C"
