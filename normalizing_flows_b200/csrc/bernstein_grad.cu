// bernstein_grad.cu — reverse mode of row #5 (row #7 of SURVEY §8a for the BernsteinBijector):
// the vector-Jacobian product of  (y, log_det) = BernsteinBijector.forward_and_log_det(x)
// (src/BernsteinBijector.py:82-96) and of constrain_thetas (src/BernsteinBijector.py:60-71) — what
// jax.grad walks through in src/maf_mwe.py:96-113,143.
//
// With u = clip(x, 1e-7, 1-1e-7), v = 1-u, n = M-1 and b_{j,d}(u) = C(d,j) u^j v^(d-j):
//     y  = Σ_j θ_j b_{j,n}            J = n Σ_j Δθ_j b_{j,n-1}            ld = log J
//     dJ/du = n(n-1) Σ_j Δ²θ_j b_{j,n-2}
// so, for cotangents (ȳ, l̄d) and g = l̄d / J:
//     x̄   = 1[lo < x < hi] · (ȳ·J + g·dJ/du)
//     θ̄_j = Σ_elements  ȳ·b_{j,n} + g·n·(b_{j-1,n-1} - b_{j,n-1})
// The three bases come from ONE direct evaluation, the lowest degree e_j = b_{j,n-2} (products of powers of
// u and v: all factors in [0,1], no division), raised twice by the convex recurrence
// b_{j,d+1} = v·b_{j,d} + u·b_{j-1,d} — the numerically benign direction.  Because
// b_{j,n} = v·d_j + u·d_{j-1} (d_j = b_{j,n-1}), the θ̄ update folds into two FMAs per coefficient:
//     θ̄_j += (ȳ·v - g·n)·d_j + (ȳ·u + g·n)·d_{j-1}.
// Everything is register resident: the order is padded to a template width MP (4..64) so all arrays are
// statically indexed; per-thread θ̄ accumulators are reduced warp -> CTA -> (fixed order, FP64) grid, which
// makes the shared-θ gradient bit-reproducible.  HBM traffic: x, ȳ, l̄d in, x̄ out (16 B per element).
#include "common.cuh"

namespace flowcuda {

constexpr int kBgThreads = 128;
constexpr int kBgMaxOrder = 64;
constexpr int kBgMaxGrid = 2048;

// C(n, j) in double, multiplicative form
__device__ __forceinline__ double bg_binom(int n, int j) {
  if (j < 0 || j > n) return 0.0;
  const int k = j < n - j ? j : n - j;
  double num = 1.0, den = 1.0;
  for (int i = 1; i <= k; ++i) {
    num *= (double)(n - k + i);
    den *= (double)i;
  }
  return num / den;
}

// jnp.clip(x, 1e-7, 1 - 1e-7) in float32 and its derivative as jax differentiates it: clip is
// minimum(maximum(x, lo), hi) and lax.max / lax.min split the cotangent evenly on a tie, so the derivative is
// 1 inside, 0 outside and 1/2 for x exactly on a bound.
__device__ __forceinline__ float bg_clip(float x, float& dclip) {
  const float lo = 1e-7f, hi = 0.99999988079071044921875f;
  dclip = (x > lo && x < hi) ? 1.f : ((x == lo || x == hi) ? 0.5f : 0.f);
  return fminf(fmaxf(x, lo), hi);
}

// d[j] = b_{j,n-1}(u) for j < MP (zero for j >= n), given the binomials ce[j] = C(n-2, j) (zero padded).
// Returns through `dd`; callers fold J and dJ/du with their own coefficients while the values are hot.
template <int MP, typename F>
__device__ __forceinline__ void bg_basis(int n, float u, float v, const float* __restrict__ ce, float (&dd)[MP], F&& visit) {
  // vq[j] = v^(n-2-j) for j <= n-2 (1 beyond): descending running product, uniform predicates
  float t = 1.f;
#pragma unroll
  for (int j = MP - 1; j >= 0; --j) {
    if (j < n - 2) t *= v;
    dd[j] = t;
  }
  float pu = 1.f, eprev = 0.f;
#pragma unroll
  for (int j = 0; j < MP; ++j) {
    const float e = ce[j] * (pu * dd[j]);  // b_{j,n-2}
    float d = fmaf(v, e, u * eprev);       // b_{j,n-1}
    if (j == 0 && n == 1) d = 1.f;         // degree-0 basis: there is no b_{.,-1} to raise
    dd[j] = d;
    visit(j, e, d);
    eprev = e;
    pu *= u;
  }
}

// ---- shared θ: theta[M], x[n] -> xbar[n], per-CTA partial thetabar ------------------------------------
struct BgTab {            // per coefficient j: {C(n-2,j), n·Δθ_j, n(n-1)·Δ²θ_j}
  float ce[kBgMaxOrder];
  float dn[kBgMaxOrder];
  float d2[kBgMaxOrder];
};

template <int MP>
__global__ void __launch_bounds__(kBgThreads)
bg_shared_kernel(const float* __restrict__ theta, int M, const float* __restrict__ x, const float* __restrict__ ybar,
                 const float* __restrict__ ldbar, float* __restrict__ xbar, float* __restrict__ partial, int64_t n_el) {
  __shared__ BgTab tab;
  __shared__ float red[kBgThreads / 32][MP];
  const int n = M - 1;
  for (int j = threadIdx.x; j < kBgMaxOrder; j += blockDim.x) {
    double c = 0.0, a = 0.0, b = 0.0;
    if (j <= n - 2) c = bg_binom(n - 2, j);
    if (j <= n - 1) a = (double)n * ((double)theta[j + 1] - (double)theta[j]);
    if (j <= n - 2) b = (double)n * (double)(n - 1) * ((double)theta[j + 2] - 2.0 * (double)theta[j + 1] + (double)theta[j]);
    tab.ce[j] = (float)c;
    tab.dn[j] = (float)a;
    tab.d2[j] = (float)b;
  }
  __syncthreads();

  float acc[MP];
#pragma unroll
  for (int j = 0; j < MP; ++j) acc[j] = 0.f;
  const float fn = (float)n;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_el; i += stride) {
    float dclip;
    const float u = bg_clip(ld_stream_f1(x + i), dclip), v = 1.f - u;
    const float yb = ybar ? ld_stream_f1(ybar + i) : 0.f;
    const float lb = ldbar ? ld_stream_f1(ldbar + i) : 0.f;
    float dd[MP];
    float J = 0.f, dJ = 0.f;
    bg_basis<MP>(n, u, v, tab.ce, dd, [&](int j, float e, float d) {
      J = fmaf(tab.dn[j], d, J);
      dJ = fmaf(tab.d2[j], e, dJ);
    });
    const float g = lb / J;
    if (xbar) st_stream_f1(xbar + i, dclip * fmaf(yb, J, g * dJ));
    const float gn = g * fn, c0 = fmaf(yb, v, -gn), c1 = fmaf(yb, u, gn);
    acc[0] = fmaf(c0, dd[0], acc[0]);
#pragma unroll
    for (int j = 1; j < MP; ++j) acc[j] = fmaf(c0, dd[j], fmaf(c1, dd[j - 1], acc[j]));
  }
  if (partial == nullptr) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int j = 0; j < MP; ++j) {
    const float s = warp_sum(acc[j]);
    if (lane == 0) red[warp][j] = s;
  }
  __syncthreads();
  for (int j = threadIdx.x; j < MP; j += blockDim.x) {
    float s = 0.f;
    for (int w = 0; w < kBgThreads / 32; ++w) s += red[w][j];
    partial[(int64_t)blockIdx.x * MP + j] = s;
  }
}

// The same computation for TWO elements per thread on the packed fp32x2 pipe (FFMA2): every operation of the
// basis construction and of the θ̄ update is elementwise in the element, so a pair costs the instruction count
// of one.  Used for padded orders <= 32 (the pair doubles the register arrays).  Tables are stored splatted
// ({c, c}) so a 64-bit shared load is already a packed operand.
struct BgTab2 {
  float2 ce[kBgMaxOrder];
  float2 dn[kBgMaxOrder];
  float2 d2[kBgMaxOrder];
};

template <int MP>
__global__ void __launch_bounds__(kBgThreads)
bg_shared2_kernel(const float* __restrict__ theta, int M, const float* __restrict__ x, const float* __restrict__ ybar,
                  const float* __restrict__ ldbar, float* __restrict__ xbar, float* __restrict__ partial, int64_t n_el) {
  __shared__ __align__(16) BgTab2 tab;
  __shared__ float red[kBgThreads / 32][MP];
  const int n = M - 1;
  for (int j = threadIdx.x; j < kBgMaxOrder; j += blockDim.x) {
    double c = 0.0, a = 0.0, b = 0.0;
    if (j <= n - 2) c = bg_binom(n - 2, j);
    if (j <= n - 1) a = (double)n * ((double)theta[j + 1] - (double)theta[j]);
    if (j <= n - 2) b = (double)n * (double)(n - 1) * ((double)theta[j + 2] - 2.0 * (double)theta[j + 1] + (double)theta[j]);
    tab.ce[j] = splat2((float)c);
    tab.dn[j] = splat2((float)a);
    tab.d2[j] = splat2((float)b);
  }
  __syncthreads();
  const f2_t* ce = reinterpret_cast<const f2_t*>(tab.ce);
  const f2_t* dn = reinterpret_cast<const f2_t*>(tab.dn);
  const f2_t* d2 = reinterpret_cast<const f2_t*>(tab.d2);

  f2_t acc[MP];
#pragma unroll
  for (int j = 0; j < MP; ++j) acc[j] = f2_splat(0.f);
  const float fn = (float)n;
  const int64_t npairs = (n_el + 1) >> 1;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < npairs; p += stride) {
    const int64_t i = p << 1;
    const bool two = i + 1 < n_el;
    float x0, x1 = 0.5f, yb0 = 0.f, yb1 = 0.f, lb0 = 0.f, lb1 = 0.f;
    if (two) {
      const float2 xv = ld_stream_f2(reinterpret_cast<const float2*>(x + i));
      x0 = xv.x; x1 = xv.y;
      if (ybar) { const float2 t = ld_stream_f2(reinterpret_cast<const float2*>(ybar + i)); yb0 = t.x; yb1 = t.y; }
      if (ldbar) { const float2 t = ld_stream_f2(reinterpret_cast<const float2*>(ldbar + i)); lb0 = t.x; lb1 = t.y; }
    } else {
      x0 = x[i];
      if (ybar) yb0 = ybar[i];
      if (ldbar) lb0 = ldbar[i];
    }
    float dc0, dc1;
    const float u0 = bg_clip(x0, dc0), u1 = bg_clip(x1, dc1), v0 = 1.f - u0, v1 = 1.f - u1;
    const f2_t u = f2_pack(u0, u1), v = f2_pack(v0, v1);
    f2_t dd[MP];
    f2_t t = f2_splat(1.f);
#pragma unroll
    for (int j = MP - 1; j >= 0; --j) {
      if (j < n - 2) t = f2_mul(t, v);
      dd[j] = t;
    }
    f2_t pu = f2_splat(1.f), eprev = f2_splat(0.f), J = f2_splat(0.f), dJ = f2_splat(0.f);
#pragma unroll
    for (int j = 0; j < MP; ++j) {
      const f2_t e = f2_mul(ce[j], f2_mul(pu, dd[j]));
      f2_t d = f2_fma(v, e, f2_mul(u, eprev));
      if (j == 0 && n == 1) d = f2_splat(1.f);
      dd[j] = d;
      J = f2_fma(dn[j], d, J);
      dJ = f2_fma(d2[j], e, dJ);
      eprev = e;
      pu = f2_mul(pu, u);
    }
    float J0, J1, dJ0, dJ1;
    f2_unpack(J, J0, J1);
    f2_unpack(dJ, dJ0, dJ1);
    const float g0 = lb0 / J0, g1 = lb1 / J1;
    if (xbar) {
      const float r0 = dc0 * fmaf(yb0, J0, g0 * dJ0), r1 = dc1 * fmaf(yb1, J1, g1 * dJ1);
      if (two) st_stream_f2(reinterpret_cast<float2*>(xbar + i), make_float2(r0, r1));
      else xbar[i] = r0;
    }
    const float gn0 = g0 * fn, gn1 = g1 * fn;
    const f2_t c0 = f2_pack(fmaf(yb0, v0, -gn0), fmaf(yb1, v1, -gn1));
    const f2_t c1 = f2_pack(fmaf(yb0, u0, gn0), fmaf(yb1, u1, gn1));
    acc[0] = f2_fma(c0, dd[0], acc[0]);
#pragma unroll
    for (int j = 1; j < MP; ++j) acc[j] = f2_fma(c0, dd[j], f2_fma(c1, dd[j - 1], acc[j]));
  }
  if (partial == nullptr) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int j = 0; j < MP; ++j) {
    float a, b;
    f2_unpack(acc[j], a, b);
    const float s = warp_sum(a + b);
    if (lane == 0) red[warp][j] = s;
  }
  __syncthreads();
  for (int j = threadIdx.x; j < MP; j += blockDim.x) {
    float s = 0.f;
    for (int w = 0; w < kBgThreads / 32; ++w) s += red[w][j];
    partial[(int64_t)blockIdx.x * MP + j] = s;
  }
}

// thetabar[j] = Σ_blocks partial[b][j], ascending b, FP64 accumulate
__global__ void bg_finalize_kernel(const float* __restrict__ partial, int blocks, int MP, int M, float* __restrict__ thetabar) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= M) return;
  double s = 0.0;
  for (int b = 0; b < blocks; ++b) s += (double)partial[(int64_t)b * MP + j];
  thetabar[j] = (float)s;
}

// ---- per-element θ rows (src/maf_mwe.py:103-107): theta[n, stride], x[n] -> xbar[n], thetabar[n, stride] ----
// One thread per element; a warp's 32 θ rows (and its 32 θ̄ rows) go through shared memory so the global
// accesses are coalesced runs instead of 32 row-strided ones.
template <int MP>
__global__ void __launch_bounds__(kBgThreads)
bg_rows_kernel(const float* __restrict__ theta, int M, int64_t tstride, const float* __restrict__ x,
               const float* __restrict__ ybar, const float* __restrict__ ldbar, float* __restrict__ xbar,
               float* __restrict__ thetabar, int64_t n_el) {
  __shared__ float ce_s[kBgMaxOrder];
  __shared__ float tile[kBgThreads / 32][32][MP + 1];
  const int n = M - 1;
  for (int j = threadIdx.x; j < kBgMaxOrder; j += blockDim.x) ce_s[j] = j <= n - 2 ? (float)bg_binom(n - 2, j) : 0.f;
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float(*tw)[MP + 1] = tile[warp];
  const float fn = (float)n, fnn = (float)n * (float)(n - 1);
  const int64_t nwarps = (int64_t)gridDim.x * (kBgThreads / 32);
  const int64_t ngroups = (n_el + 31) >> 5;
  for (int64_t grp = (int64_t)blockIdx.x * (kBgThreads / 32) + warp; grp < ngroups; grp += nwarps) {
    const int64_t base = grp << 5;
    const int rows = (int)min((int64_t)32, n_el - base);
    for (int f = lane; f < rows * M; f += 32) {
      const int r = f / M, j = f - r * M;
      tw[r][j] = theta[(base + r) * tstride + j];
    }
    __syncwarp();
    const int64_t i = base + lane;
    const bool live = lane < rows;
    float th[MP];
#pragma unroll
    for (int j = 0; j < MP; ++j) th[j] = (live && j < M) ? tw[lane][j] : 0.f;
    __syncwarp();
    float dclip = 0.f, u = 0.5f, yb = 0.f, lb = 0.f;
    if (live) {
      u = bg_clip(ld_stream_f1(x + i), dclip);
      if (ybar) yb = ld_stream_f1(ybar + i);
      if (ldbar) lb = ld_stream_f1(ldbar + i);
    }
    const float v = 1.f - u;
    float dd[MP];
    float J = 0.f, dJ = 0.f;
    bg_basis<MP>(n, u, v, ce_s, dd, [&](int j, float e, float d) {
      // Δθ_j and Δ²θ_j of this element's own row (zero beyond the order: th is zero padded, and the
      // basis values they would multiply are zero there as well — guard the differences that straddle M)
      const float t0 = th[j], t1 = j + 1 < MP ? th[j + 1] : 0.f, t2 = j + 2 < MP ? th[j + 2] : 0.f;
      if (j <= n - 1) J = fmaf(t1 - t0, d, J);
      if (j <= n - 2) dJ = fmaf((t2 - t1) - (t1 - t0), e, dJ);
    });
    J *= fn;
    dJ *= fnn;
    const float g = live ? lb / J : 0.f;
    if (live && xbar) st_stream_f1(xbar + i, dclip * fmaf(yb, J, g * dJ));
    if (thetabar) {
      const float gn = g * fn, c0 = fmaf(yb, v, -gn), c1 = fmaf(yb, u, gn);
#pragma unroll
      for (int j = 0; j < MP; ++j)
        if (j < M) tw[lane][j] = j == 0 ? c0 * dd[0] : fmaf(c0, dd[j], c1 * dd[j - 1]);
      __syncwarp();
      for (int f = lane; f < rows * M; f += 32) {
        const int r = f / M, j = f - r * M;
        thetabar[(base + r) * tstride + j] = tw[r][j];
      }
      __syncwarp();
    }
  }
}

// ---- constrain_thetas VJP: θ = cumsum([raw_0, softplus(raw_k)+1e-4 ...]) (BernsteinBijector.py:60-71) ----
//     raw̄_k = σ(raw_k)·Σ_{j>=k} θ̄_j  (k >= 1),   raw̄_0 = Σ_j θ̄_j
__global__ void __launch_bounds__(256)
bg_constrain_vjp_kernel(const float* __restrict__ raw, const float* __restrict__ thetabar, float* __restrict__ rawbar,
                        int64_t rows, int M) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  const float* p = raw + r * M;
  const float* tb = thetabar + r * M;
  float* o = rawbar + r * M;
  float suffix = 0.f;
  for (int k = M - 1; k >= 1; --k) {
    suffix += tb[k];
    o[k] = suffix / (1.f + expf(-p[k]));
  }
  o[0] = suffix + tb[0];
}

template <int MP>
static int bg_launch(const float* theta, int M, int64_t tstride, const float* x, const float* ybar, const float* ldbar,
                     float* xbar, float* thetabar, int64_t n, float* partial, cudaStream_t s) {
  if (tstride == 0) {
    int grid;
    const bool pair_ok = ((((uintptr_t)x | (uintptr_t)ybar | (uintptr_t)ldbar | (uintptr_t)xbar) & 7) == 0);
    if constexpr (MP <= 32) {
      if (pair_ok) {
        grid = grid_resident(bg_shared2_kernel<MP>, kBgThreads, 0, (n + 1) / 2, kBgThreads);
        if (grid > kBgMaxGrid) grid = kBgMaxGrid;
        bg_shared2_kernel<MP><<<grid, kBgThreads, 0, s>>>(theta, M, x, ybar, ldbar, xbar, thetabar ? partial : nullptr, n);
      } else {
        grid = grid_resident(bg_shared_kernel<MP>, kBgThreads, 0, n, kBgThreads);
        if (grid > kBgMaxGrid) grid = kBgMaxGrid;
        bg_shared_kernel<MP><<<grid, kBgThreads, 0, s>>>(theta, M, x, ybar, ldbar, xbar, thetabar ? partial : nullptr, n);
      }
    } else {
      grid = grid_resident(bg_shared_kernel<MP>, kBgThreads, 0, n, kBgThreads);
      if (grid > kBgMaxGrid) grid = kBgMaxGrid;
      bg_shared_kernel<MP><<<grid, kBgThreads, 0, s>>>(theta, M, x, ybar, ldbar, xbar, thetabar ? partial : nullptr, n);
    }
    FC_LAUNCH_CHECK("bernstein vjp kernel");
    if (thetabar) {
      bg_finalize_kernel<<<1, 64, 0, s>>>(partial, grid, MP, M, thetabar);
      FC_LAUNCH_CHECK("bernstein vjp finalize kernel");
    }
  } else {
    const int grid = grid_resident(bg_rows_kernel<MP>, kBgThreads, 0, n, kBgThreads);
    bg_rows_kernel<MP><<<grid, kBgThreads, 0, s>>>(theta, M, tstride, x, ybar, ldbar, xbar, thetabar, n);
    FC_LAUNCH_CHECK("bernstein rows vjp kernel");
  }
  return FLOWCUDA_OK;
}

}  // namespace flowcuda

using namespace flowcuda;

extern "C" {

size_t flowcuda_bernstein_vjp_workspace_bytes(int order, int64_t theta_stride, int64_t n) {
  (void)order;
  (void)n;
  return theta_stride == 0 ? (size_t)kBgMaxGrid * kBgMaxOrder * sizeof(float) : 0;
}

int flowcuda_bernstein_forward_and_log_det_vjp(const float* theta, int order, int64_t theta_stride, const float* x,
                                               const float* ybar, const float* ldbar, float* xbar, float* thetabar,
                                               int64_t n, void* workspace, size_t workspace_bytes, void* stream) {
  FC_REQUIRE(theta && (n <= 0 || (x && (ybar || ldbar))),
             "bernstein_vjp: NULL pointer (theta, x and one of ybar/ldbar are required)");
  FC_REQUIRE(xbar || thetabar, "bernstein_vjp: neither xbar nor thetabar requested");
  FC_REQUIRE(order >= 2 && order <= kBgMaxOrder, "bernstein_vjp: order=%d outside [2,%d]", order, kBgMaxOrder);
  FC_REQUIRE(theta_stride == 0 || theta_stride >= order, "bernstein_vjp: theta_stride must be 0 or >= order");
  FC_REQUIRE(n >= 0, "bernstein_vjp: negative n");
  cudaStream_t s = (cudaStream_t)stream;
  if (theta_stride == 0 && thetabar) {
    FC_REQUIRE(workspace && workspace_bytes >= flowcuda_bernstein_vjp_workspace_bytes(order, 0, n),
               "bernstein_vjp: workspace too small (%zu bytes)", workspace_bytes);
  }
  if (n == 0) {
    if (theta_stride == 0 && thetabar) FC_CUDA(cudaMemsetAsync(thetabar, 0, (size_t)order * sizeof(float), s));
    return FLOWCUDA_OK;
  }
  float* partial = reinterpret_cast<float*>(workspace);
#define BG_CASE(W) \
  if (order <= W) return bg_launch<W>(theta, order, theta_stride, x, ybar, ldbar, xbar, thetabar, n, partial, s)
  BG_CASE(4);
  BG_CASE(8);
  BG_CASE(12);
  BG_CASE(16);
  BG_CASE(24);
  BG_CASE(32);
  BG_CASE(48);
  BG_CASE(64);
#undef BG_CASE
  return FLOWCUDA_EINVAL;
}

int flowcuda_bernstein_constrain_thetas_vjp(const float* theta_raw, const float* thetabar, float* rawbar, int64_t rows,
                                            int order, void* stream) {
  FC_REQUIRE(theta_raw && thetabar && rawbar, "bernstein_constrain_vjp: NULL pointer");
  FC_REQUIRE(order >= 1 && rows >= 0, "bernstein_constrain_vjp: bad sizes");
  if (rows == 0) return FLOWCUDA_OK;
  bg_constrain_vjp_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, (cudaStream_t)stream>>>(theta_raw, thetabar, rawbar,
                                                                                            rows, order);
  FC_LAUNCH_CHECK("bernstein constrain vjp kernel");
  return FLOWCUDA_OK;
}

}  // extern "C"
