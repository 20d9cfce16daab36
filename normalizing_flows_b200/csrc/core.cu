// core.cu — error plumbing, device queries, loss reduction and optimiser steps (rows #7, #9).
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace flowcuda {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return FLOWCUDA_ECUDA;
}

bool pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("FLOWCUDA_PDL");
    on = (e && e[0] == '0') ? 0 : 1;
  }
  return on != 0;
}

std::recursive_mutex& side_mutex() {
  static std::recursive_mutex m;
  return m;
}

bool side_env_enabled(const char* name) {
  const char* e = getenv(name);
  return !(e && e[0] == '0');
}

int side_ctx_get(SideCtx** out) {
  static SideCtx tab[64] = {};
  int dev = -1;
  FC_CUDA(cudaGetDevice(&dev));
  FC_REQUIRE(dev >= 0 && dev < 64, "device ordinal %d outside [0, 64)", dev);
  std::lock_guard<std::recursive_mutex> lock(side_mutex());
  SideCtx& sd = tab[dev];
  if (!sd.q) {
    cudaStream_t q;
    FC_CUDA(cudaStreamCreateWithFlags(&q, cudaStreamNonBlocking));
    FC_CUDA(cudaEventCreateWithFlags(&sd.ev_m, cudaEventDisableTiming));
    FC_CUDA(cudaEventCreateWithFlags(&sd.ev_last, cudaEventDisableTiming));
    FC_CUDA(cudaEventCreateWithFlags(&sd.ev_s[0], cudaEventDisableTiming));
    FC_CUDA(cudaEventCreateWithFlags(&sd.ev_s[1], cudaEventDisableTiming));
    sd.q = q;
  }
  *out = &sd;
  return FLOWCUDA_OK;
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// ---- deterministic scaled sum: out = scale * sum(v) -------------------------------------
constexpr int kSumBlocks = 1024;  // == workspace doubles
constexpr int kSumThreads = 256;

__global__ void __launch_bounds__(kSumThreads) sum_stage1(const float* __restrict__ v, int64_t n,
                                                         double* __restrict__ partial) {
  __shared__ double sm[kSumThreads / 32];
  // fixed assignment: block b owns the contiguous chunk [b*chunk, (b+1)*chunk)
  int64_t chunk = (n + gridDim.x - 1) / gridDim.x;
  int64_t lo = (int64_t)blockIdx.x * chunk;
  int64_t hi = lo + chunk < n ? lo + chunk : n;
  double acc = 0.0;
  for (int64_t i = lo + threadIdx.x; i < hi; i += kSumThreads) acc += (double)v[i];
  acc = warp_sum_d(acc);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < kSumThreads / 32; ++w) t += sm[w];
    partial[blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(kSumThreads) sum_stage2(const double* __restrict__ partial, int nb,
                                                         double scale, float* __restrict__ out) {
  __shared__ double sm[kSumThreads / 32];
  double acc = 0.0;
  for (int i = threadIdx.x; i < nb; i += kSumThreads) acc += partial[i];
  acc = warp_sum_d(acc);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < kSumThreads / 32; ++w) t += sm[w];
    out[0] = (float)(t * scale);
  }
}

// ---- optimiser steps ---------------------------------------------------------------------
__global__ void __launch_bounds__(256) sgd_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                  int64_t n, float lr, float gs) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) p[i] = p[i] - lr * (g[i] * gs);
}

// optax.adam: 7 streams of 4 bytes per parameter (read p, g, m, v; write p, m, v) and nothing else — HBM-bound.
// 128-bit accesses, four parameters per thread and iteration (the scalar form ran at 35 % of the HBM roof: 0.55 ms for
// the 88 M parameters of the maf_mnist flow); rsqrt-free on purpose: sqrt and the division are IEEE, as in optax.
__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float lr, float b1, float b2, float eps, float gs,
                                         float inv_bc1, float inv_bc2) {
  const float gi = g * gs;
  m = b1 * m + (1.f - b1) * gi;
  v = b2 * v + (1.f - b2) * gi * gi;
  p = p - lr * (m * inv_bc1) / (sqrtf(v * inv_bc2) + eps);
}
__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                   float* __restrict__ m, float* __restrict__ v, int64_t n,
                                                   float lr, float b1, float b2, float eps, float gs,
                                                   float inv_bc1, float inv_bc2) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool vec = ((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0);
  const int64_t nq = vec ? n >> 2 : 0;
  for (int64_t q = tid; q < nq; q += stride) {
    float4 pp = reinterpret_cast<float4*>(p)[q], mm = reinterpret_cast<float4*>(m)[q], vv = reinterpret_cast<float4*>(v)[q];
    const float4 gg = ld_stream_f4(reinterpret_cast<const float4*>(g) + q);
    adam_one(pp.x, gg.x, mm.x, vv.x, lr, b1, b2, eps, gs, inv_bc1, inv_bc2);
    adam_one(pp.y, gg.y, mm.y, vv.y, lr, b1, b2, eps, gs, inv_bc1, inv_bc2);
    adam_one(pp.z, gg.z, mm.z, vv.z, lr, b1, b2, eps, gs, inv_bc1, inv_bc2);
    adam_one(pp.w, gg.w, mm.w, vv.w, lr, b1, b2, eps, gs, inv_bc1, inv_bc2);
    reinterpret_cast<float4*>(p)[q] = pp;
    reinterpret_cast<float4*>(m)[q] = mm;
    reinterpret_cast<float4*>(v)[q] = vv;
  }
  for (int64_t i = (nq << 2) + tid; i < n; i += stride) {
    float pi = p[i], mi = m[i], vi = v[i];
    adam_one(pi, g[i], mi, vi, lr, b1, b2, eps, gs, inv_bc1, inv_bc2);
    p[i] = pi;
    m[i] = mi;
    v[i] = vi;
  }
}

// ---- tutorial bijector: 2-D rotation (src/rotation_flow.py:13-37) -----------------------------
// forward y = x @ R with R = [[c,-s],[s,c]]^T; inverse y = x @ R^T.  Two samples per thread (one 16-byte access).
__global__ void __launch_bounds__(256) rotate2d_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t n,
                                                       float c, float s) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t pairs = n >> 1;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < pairs; i += stride) {
    const float4 v = reinterpret_cast<const float4*>(x)[i];
    reinterpret_cast<float4*>(y)[i] = make_float4(v.x * c - v.y * s, v.x * s + v.y * c, v.z * c - v.w * s, v.z * s + v.w * c);
  }
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
    const float a = x[2 * (n - 1)], b = x[2 * (n - 1) + 1];
    y[2 * (n - 1)] = a * c - b * s;
    y[2 * (n - 1) + 1] = a * s + b * c;
  }
}

}  // namespace flowcuda

using namespace flowcuda;

extern "C" {

int flowcuda_version(void) { return FLOWCUDA_VERSION; }
const char* flowcuda_last_error(void) { return g_err; }
int flowcuda_sm_count(void) { return sm_count(); }

int flowcuda_sum_scaled(const float* v, int64_t n, double scale, float* out, void* workspace, void* stream) {
  FC_REQUIRE(v && out && workspace && n >= 0, "sum_scaled: NULL pointer or negative n");
  cudaStream_t s = (cudaStream_t)stream;
  int nb = (int)((n + 4095) / 4096);
  if (nb < 1) nb = 1;
  if (nb > kSumBlocks) nb = kSumBlocks;
  sum_stage1<<<nb, kSumThreads, 0, s>>>(v, n, (double*)workspace);
  FC_LAUNCH_CHECK("sum_stage1");
  sum_stage2<<<1, kSumThreads, 0, s>>>((const double*)workspace, nb, scale, out);
  FC_LAUNCH_CHECK("sum_stage2");
  return FLOWCUDA_OK;
}

// ---- flowcuda_sum_scaled without a caller workspace and without any allocation ----------------------------
// The stage-1 partials live in a file-scope device array instead of caller memory.  kSumBlocks doubles = 8 KB per
// device, part of the module image: nothing here (or below it) calls cudaMalloc/cudaFree or an allocating library.
// The array is ONE per device and per loaded library: it keeps the previous call's partials, and two calls that
// overlap on the same device (different streams, or different host threads) race on it — callers serialise them
// (same stream, or an event between them) or use flowcuda_sum_scaled with their own workspaces.
static __device__ __align__(256) double g_sum_static_scratch[kSumBlocks];
constexpr int64_t kSumStaticMaxN = (int64_t)1 << 26;  // 64 Mi floats (256 MB of input): stated limit of the entry point
constexpr int kSumStaticMaxDevices = 64;

// Device address of the scratch array on the CURRENT device.  The host-side symbol is a handle, not a device
// address: it is resolved once per device with that device current, and the result is cached.
static int sum_static_scratch(double** out) {
  static double* cached[kSumStaticMaxDevices] = {nullptr};
  int dev = -1;
  FC_CUDA(cudaGetDevice(&dev));
  FC_REQUIRE(dev >= 0 && dev < kSumStaticMaxDevices, "sum_scaled_static: device ordinal %d outside [0, %d)", dev,
             kSumStaticMaxDevices);
  double* p = __atomic_load_n(&cached[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    void* sym = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&sym, g_sum_static_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_sum_static_scratch)");
    FC_REQUIRE(sym != nullptr && ((uintptr_t)sym & 255) == 0, "sum_scaled_static: scratch symbol resolved to %p", sym);
    p = (double*)sym;
    __atomic_store_n(&cached[dev], p, __ATOMIC_RELEASE);  // every thread resolves the same address: benign if repeated
  }
  *out = p;
  return FLOWCUDA_OK;
}

int flowcuda_sum_scaled_static(const float* v, int64_t n, double scale, float* out, void* stream) {
  FC_REQUIRE(out && n >= 0 && (v || n == 0), "sum_scaled_static: NULL pointer or negative n");
  FC_REQUIRE(n <= kSumStaticMaxN, "sum_scaled_static: n = %lld exceeds the limit of %lld elements of this entry point "
             "(use flowcuda_sum_scaled with a caller workspace)", (long long)n, (long long)kSumStaticMaxN);
  double* scratch = nullptr;
  int rc = sum_static_scratch(&scratch);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  int nb = (int)((n + 4095) / 4096);  // same grid rule as flowcuda_sum_scaled -> same summation order, same bits
  if (nb < 1) nb = 1;
  if (nb > kSumBlocks) nb = kSumBlocks;
  // Stage 1 overwrites every partial stage 2 reads, so nothing has to start at zero for a correct result; the slots
  // are cleared all the same so that a stale partial of an earlier (larger) call can never be summed by mistake.
  FC_CUDA(cudaMemsetAsync(scratch, 0, (size_t)nb * sizeof(double), s));
  sum_stage1<<<nb, kSumThreads, 0, s>>>(v, n, scratch);
  FC_LAUNCH_CHECK("sum_stage1 (static scratch)");
  sum_stage2<<<1, kSumThreads, 0, s>>>(scratch, nb, scale, out);
  FC_LAUNCH_CHECK("sum_stage2 (static scratch)");
  return FLOWCUDA_OK;
}

int flowcuda_rotate2d(const float* x, float cos_theta, float sin_theta, int inverse, float* y, int64_t n, void* stream) {
  FC_REQUIRE(n >= 0 && ((x && y) || n == 0), "rotate2d: NULL pointer or negative n");
  FC_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)y & 15) == 0, "rotate2d: buffers must be 16-byte aligned");
  if (n == 0) return FLOWCUDA_OK;
  rotate2d_kernel<<<grid_for(n > 1 ? n >> 1 : 1, 256), 256, 0, (cudaStream_t)stream>>>(x, y, n, cos_theta,
                                                                                      inverse ? -sin_theta : sin_theta);
  FC_LAUNCH_CHECK("rotate2d_kernel");
  return FLOWCUDA_OK;
}

int flowcuda_sgd_step(float* p, const float* g, int64_t count, float lr, float grad_scale, void* stream) {
  FC_REQUIRE(p && g && count >= 0, "sgd_step: NULL pointer or negative count");
  if (count == 0) return FLOWCUDA_OK;
  sgd_kernel<<<grid_for(count, 256), 256, 0, (cudaStream_t)stream>>>(p, g, count, lr, grad_scale);
  FC_LAUNCH_CHECK("sgd_kernel");
  return FLOWCUDA_OK;
}

int flowcuda_adam_step(float* p, const float* g, float* m, float* v, int64_t count, int64_t step, float lr,
                       float b1, float b2, float eps, float grad_scale, void* stream) {
  FC_REQUIRE(p && g && m && v && count >= 0 && step >= 1, "adam_step: bad arguments");
  if (count == 0) return FLOWCUDA_OK;
  float inv_bc1 = (float)(1.0 / (1.0 - pow((double)b1, (double)step)));
  float inv_bc2 = (float)(1.0 / (1.0 - pow((double)b2, (double)step)));
  adam_kernel<<<grid_for((count + 3) / 4, 256, 16), 256, 0, (cudaStream_t)stream>>>(p, g, m, v, count, lr, b1, b2, eps,
                                                                      grad_scale, inv_bc1, inv_bc2);
  FC_LAUNCH_CHECK("adam_kernel");
  return FLOWCUDA_OK;
}

}  // extern "C"
