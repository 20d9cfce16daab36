// bernstein.cu — row #5: BernsteinBijector (src/BernsteinBijector.py) and the fused Bernstein
// flow density of src/bflow.py.
//
// The reference evaluates  y = mean_k Beta(k, M-k+1).pdf(x)·θ_k  and
// J = Σ_k Beta(k, M-k).pdf(x)·(θ_{k+1}-θ_k)  by materialising [N, M] Beta-pdf temporaries.
// With integer Beta parameters these are Bernstein forms of degree n=M-1 and n-1
// (SURVEY §4 K5):   y = Σ_j θ_j C(n,j) u^j v^(n-j),   J = n Σ_j Δθ_j C(n-1,j) u^j v^(n-1-j),
// u = clip(x), v = 1-u.  Both are evaluated in one pass with the uniform-coefficient
// recurrence  r <- r·v + c_j·u^j  (stable on the whole of [0,1]; all lanes read the same
// coefficient), on the packed fp32x2 pipe: 5 packed ops per coefficient per two elements.
// HBM traffic: x in, (y, log_det) out — nothing else.
#include "common.cuh"

namespace flowcuda {

constexpr int kBernThreads = 256;
constexpr int kBernMaxOrder = 64;

__device__ __forceinline__ float bern_clip(float x) {
  // jnp.clip(x, 1e-7, 1 - 1e-7) in float32: the upper bound rounds to 1 - 2^-23
  return fminf(fmaxf(x, 1e-7f), 0.99999988079071044921875f);
}

// Coefficient tables for a shared theta, every coefficient duplicated ({c, c}: a ready-made packed
// operand, so the inner loop needs no splat moves), zero padded:
//   a[j] = θ_j C(n,j) (j <= n),   d[j] = n (θ_{j+1}-θ_j) C(n-1,j) (j <= n-1), d[n] = 0.
struct BernTab {
  float2 a[kBernMaxOrder + 8];
  float2 d[kBernMaxOrder + 8];
};

// C(n, j) in double by the multiplicative formula over min(j, n-j) factors (one division at the end)
__device__ __forceinline__ double bern_binom(int n, int j) {
  if (j < 0 || j > n) return 0.0;
  const int k = j < n - j ? j : n - j;
  double num = 1.0, den = 1.0;
  for (int i = 1; i <= k; ++i) {
    num *= (double)(n - k + i);
    den *= (double)i;
  }
  return num / den;
}

// one thread per coefficient (a serial build by thread 0 costs ~10 us of FP64 latency per CTA)
__device__ void bern_build_table(BernTab* tab, const float* __restrict__ theta, int M) {
  const int n = M - 1;
  for (int j = threadIdx.x; j < kBernMaxOrder + 8; j += blockDim.x) {
    float a = 0.f, d = 0.f;
    if (j <= n) {
      a = (float)((double)theta[j] * bern_binom(n, j));
      if (j < n) d = (float)((double)n * ((double)theta[j + 1] - (double)theta[j]) * bern_binom(n - 1, j));
    }
    tab->a[j] = make_float2(a, a);
    tab->d[j] = make_float2(d, d);
  }
  __syncthreads();
}

// Evaluate y = Σ_j a_j u^j v^(n-j) and vJ = Σ_j d_j u^j v^(n-j) (= v·J: the J form has degree n-1, its
// padding coefficient d_n = 0 costs one factor v that the caller divides out) for FOUR clipped elements,
// two packed pairs.  The plain recurrence r <- r·v + c_j·u^j costs 2.5 FP32 ops per coefficient and
// polynomial and is FMA-pipe bound well below the HBM roof; here eight coefficients are folded per
// step against the precomputed monomials m_i = u^i v^(7-i):
//     r <- r·v^8 + u^(8b) · Σ_{i<8} c_{8b+i} m_i            (1.3 ops per coefficient and polynomial)
// which is the same sum of non-negative Bernstein terms, only associated differently (stable on [0,1]).
__device__ __forceinline__ void bern_eval_quad(const BernTab* __restrict__ tab, int M, f2_t u0, f2_t u1, f2_t& y0,
                                               f2_t& J0, f2_t& y1, f2_t& J1) {
  const f2_t one = f2_splat(1.f), neg1 = f2_splat(-1.f);
  const f2_t v0 = f2_fma(u0, neg1, one), v1 = f2_fma(u1, neg1, one);
  f2_t ry0 = f2_splat(0.f), ry1 = ry0, rj0 = ry0, rj1 = ry0, pw0 = one, pw1 = one;
  const int nb = M >> 3;
  if (nb > 0) {
    f2_t m0[8], m1[8];
    {
      // powers u^k, v^k (k <= 7), then m[i] = u^i v^(7-i)
      f2_t up0[8], vp0[8], up1[8], vp1[8];
      up0[0] = one; vp0[0] = one; up1[0] = one; vp1[0] = one;
#pragma unroll
      for (int k = 1; k < 8; ++k) {
        up0[k] = f2_mul(up0[k - 1], u0); vp0[k] = f2_mul(vp0[k - 1], v0);
        up1[k] = f2_mul(up1[k - 1], u1); vp1[k] = f2_mul(vp1[k - 1], v1);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        m0[i] = f2_mul(up0[i], vp0[7 - i]);
        m1[i] = f2_mul(up1[i], vp1[7 - i]);
      }
    }
    const f2_t u80 = f2_mul(m0[7], u0), v80 = f2_mul(m0[0], v0), u81 = f2_mul(m1[7], u1), v81 = f2_mul(m1[0], v1);
#pragma unroll 1
    for (int b = 0; b < nb; ++b) {
      const ulonglong2* ta = reinterpret_cast<const ulonglong2*>(tab->a + 8 * b);
      const ulonglong2* td = reinterpret_cast<const ulonglong2*>(tab->d + 8 * b);
      f2_t iy0, iy1, ij0, ij1;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const ulonglong2 c = ta[i], e = td[i];  // {a_{2i}, a_{2i+1}} and {d_{2i}, d_{2i+1}}, each splatted
        if (i == 0) {
          iy0 = f2_mul(c.x, m0[0]); iy1 = f2_mul(c.x, m1[0]);
          ij0 = f2_mul(e.x, m0[0]); ij1 = f2_mul(e.x, m1[0]);
        } else {
          iy0 = f2_fma(c.x, m0[2 * i], iy0); iy1 = f2_fma(c.x, m1[2 * i], iy1);
          ij0 = f2_fma(e.x, m0[2 * i], ij0); ij1 = f2_fma(e.x, m1[2 * i], ij1);
        }
        iy0 = f2_fma(c.y, m0[2 * i + 1], iy0); iy1 = f2_fma(c.y, m1[2 * i + 1], iy1);
        ij0 = f2_fma(e.y, m0[2 * i + 1], ij0); ij1 = f2_fma(e.y, m1[2 * i + 1], ij1);
      }
      ry0 = f2_fma(ry0, v80, f2_mul(pw0, iy0));
      ry1 = f2_fma(ry1, v81, f2_mul(pw1, iy1));
      rj0 = f2_fma(rj0, v80, f2_mul(pw0, ij0));
      rj1 = f2_fma(rj1, v81, f2_mul(pw1, ij1));
      pw0 = f2_mul(pw0, u80);
      pw1 = f2_mul(pw1, u81);
    }
  }
  for (int j = nb << 3; j < M; ++j) {  // remaining coefficients, one at a time
    const f2_t c = *reinterpret_cast<const f2_t*>(tab->a + j), e = *reinterpret_cast<const f2_t*>(tab->d + j);
    ry0 = f2_fma(ry0, v0, f2_mul(c, pw0));
    ry1 = f2_fma(ry1, v1, f2_mul(c, pw1));
    rj0 = f2_fma(rj0, v0, f2_mul(e, pw0));
    rj1 = f2_fma(rj1, v1, f2_mul(e, pw1));
    pw0 = f2_mul(pw0, u0);
    pw1 = f2_mul(pw1, u1);
  }
  y0 = ry0; y1 = ry1;
  // divide the padding factor v out of the Jacobian form
  float a, b, c, d, va, vb, vc, vd;
  f2_unpack(rj0, a, b); f2_unpack(rj1, c, d);
  f2_unpack(v0, va, vb); f2_unpack(v1, vc, vd);
  J0 = f2_pack(__fdividef(a, va), __fdividef(b, vb));
  J1 = f2_pack(__fdividef(c, vc), __fdividef(d, vd));
}

// scalar form for the ragged tail
__device__ __forceinline__ void bern_eval_pair(const BernTab* __restrict__ tab, int M, float2 u, float2& y, float2& J) {
  f2_t y0, J0, y1, J1;
  bern_eval_quad(tab, M, f2_pack(u.x, u.y), f2_pack(u.x, u.y), y0, J0, y1, J1);
  f2_unpack(y0, y.x, y.y);
  f2_unpack(J0, J.x, J.y);
}

// tfb.SoftClip(0, 1, hinge_softness=c) forward and log-det (bflow.py:29).
__device__ __forceinline__ void softclip01(float x, float c, float k, float logk, float& y, float& ld) {
  float inv_c = 1.f / c;
  float spx = c * softplus_f(x * inv_c);
  float u = 1.f - spx;
  y = 1.f - k * (c * softplus_f(u * inv_c));
  ld = logk - softplus_f(-u * inv_c) - softplus_f(-x * inv_c);
}

// MODE 0: forward_and_log_det (y, ld);  MODE 1: fused bflow log_prob.
template <int MODE>
__global__ void __launch_bounds__(kBernThreads, 3)
bern_shared_kernel(const float* __restrict__ theta, int M, const float* __restrict__ scal,
                   const float* __restrict__ x, float* __restrict__ out_y, float* __restrict__ out_ld, int64_t n) {
  __shared__ __align__(16) BernTab tab_s;
  const BernTab* tab = &tab_s;
  bern_build_table(&tab_s, theta, M);
  float s1 = 1.f, b1 = 0.f, s4 = 1.f, b4 = 0.f, lconst = 0.f, kk = 1.f, logk = 0.f;
  if (MODE == 1) {
    // T1 = Chain([Scale(softplus(a)), Shift(b)]): v = softplus(a)·(x + b)  (bflow.py:28)
    s1 = softplus_f(scal[0]);
    b1 = scal[1];
    s4 = softplus_f(scal[2]);
    b4 = scal[3];
    kk = 1.f / (1.5f * softplus_f(1.f / 1.5f));
    logk = logf(kk);
    lconst = logf(s1) + logf(s4) - kHalfLog2Pi;
  }
  const int64_t nquads = n >> 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  float4 xnext = make_float4(0.5f, 0.5f, 0.5f, 0.5f);
  if (q < nquads) xnext = ld_stream_f4(reinterpret_cast<const float4*>(x) + q);
  for (; q < nquads; q += stride) {
    const float4 xv = xnext;
    if (q + stride < nquads) xnext = ld_stream_f4(reinterpret_cast<const float4*>(x) + q + stride);  // next iteration's load
    float xs[4] = {xv.x, xv.y, xv.z, xv.w};
    float pre_ld[4] = {0.f, 0.f, 0.f, 0.f};
    if (MODE == 1) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float t = s1 * (xs[i] + b1);
        softclip01(t, 1.5f, kk, logk, xs[i], pre_ld[i]);
      }
    }
    f2_t y01, J01, y23, J23;
    bern_eval_quad(tab, M, f2_pack(bern_clip(xs[0]), bern_clip(xs[1])), f2_pack(bern_clip(xs[2]), bern_clip(xs[3])),
                   y01, J01, y23, J23);
    float ys[4], jj[4];
    f2_unpack(y01, ys[0], ys[1]);
    f2_unpack(y23, ys[2], ys[3]);
    f2_unpack(J01, jj[0], jj[1]);
    f2_unpack(J23, jj[2], jj[3]);
    // MUFU.LG2 based log: relative error 2^-22 away from J = 1, absolute 1e-7 near it (inside the 1e-5 budget)
    float ld[4] = {__logf(jj[0]), __logf(jj[1]), __logf(jj[2]), __logf(jj[3])};
    if (MODE == 0) {
      if (out_y) st_stream_f4(reinterpret_cast<float4*>(out_y) + q, make_float4(ys[0], ys[1], ys[2], ys[3]));
      if (out_ld) st_stream_f4(reinterpret_cast<float4*>(out_ld) + q, make_float4(ld[0], ld[1], ld[2], ld[3]));
    } else {
      float lp[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float z = s4 * (ys[i] + b4);  // T4 (bflow.py:31-32)
        lp[i] = fmaf(-0.5f * z, z, lconst + pre_ld[i] + ld[i]);
      }
      st_stream_f4(reinterpret_cast<float4*>(out_ld) + q, make_float4(lp[0], lp[1], lp[2], lp[3]));
    }
  }
  // tail (n % 4 elements)
  int64_t t = (nquads << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) {
    float xi = x[t], pl = 0.f;
    if (MODE == 1) softclip01(s1 * (xi + b1), 1.5f, kk, logk, xi, pl);
    float2 y, J;
    bern_eval_pair(tab, M, splat2(bern_clip(xi)), y, J);
    if (MODE == 0) {
      if (out_y) out_y[t] = y.x;
      if (out_ld) out_ld[t] = logf(J.x);
    } else {
      float z = s4 * (y.x + b4);
      out_ld[t] = fmaf(-0.5f * z, z, lconst + pl + logf(J.x));
    }
  }
}

// Small orders (M <= 16: the reference's own M = 9, bflow.py:24, and M = 8, maf_mwe.py:103) — HBM-bound work that the
// general kernel above ran latency-bound: its eight-coefficient folding carries 94 registers (two CTAs per SM, 22 % of
// the warps) to save FMAs that at these orders are not the limit (ncu r01: M = 9 at 31 % of the HBM roof).  Here the
// order is a template parameter, the plain recurrences  y <- y·v + a_j·u^j  (j <= n) and  J <- J·v + d_j·u^j  (j <= n-1:
// no padding factor to divide out) are fully unrolled on packed pairs, and the kernel keeps < 52 registers so that
// five CTAs (40 warps) per SM hide the load latency; 37 packed operations per element pair at M = 9.
template <int M, int MODE>
__global__ void __launch_bounds__(kBernThreads, 5)
bern_small_kernel(const float* __restrict__ theta, const float* __restrict__ scal, const float* __restrict__ x,
                  float* __restrict__ out_y, float* __restrict__ out_ld, int64_t n) {
  constexpr int deg = M - 1;
  __shared__ __align__(16) float2 ta[M], td[M];
  if (threadIdx.x < M) {
    const int j = threadIdx.x;
    const float a = (float)((double)theta[j] * bern_binom(deg, j));
    float d = 0.f;
    if (j < deg) d = (float)((double)deg * ((double)theta[j + 1] - (double)theta[j]) * bern_binom(deg - 1, j));
    ta[j] = make_float2(a, a);
    td[j] = make_float2(d, d);
  }
  __syncthreads();
  float s1 = 1.f, b1 = 0.f, s4 = 1.f, b4 = 0.f, lconst = 0.f, kk = 1.f, logk = 0.f;
  if (MODE == 1) {
    s1 = softplus_f(scal[0]);
    b1 = scal[1];
    s4 = softplus_f(scal[2]);
    b4 = scal[3];
    kk = 1.f / (1.5f * softplus_f(1.f / 1.5f));
    logk = logf(kk);
    lconst = logf(s1) + logf(s4) - kHalfLog2Pi;
  }
  const f2_t one = f2_splat(1.f), neg1 = f2_splat(-1.f);
  const int64_t nquads = n >> 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < nquads; q += stride) {
    const float4 xv = ld_stream_f4(reinterpret_cast<const float4*>(x) + q);
    float xs[4] = {xv.x, xv.y, xv.z, xv.w};
    float pre_ld[4] = {0.f, 0.f, 0.f, 0.f};
    if (MODE == 1) {
#pragma unroll
      for (int i = 0; i < 4; ++i) softclip01(s1 * (xs[i] + b1), 1.5f, kk, logk, xs[i], pre_ld[i]);
    }
    f2_t u[2] = {f2_pack(bern_clip(xs[0]), bern_clip(xs[1])), f2_pack(bern_clip(xs[2]), bern_clip(xs[3]))};
    f2_t v[2], ry[2], rj[2], pw[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      v[h] = f2_fma(u[h], neg1, one);
      ry[h] = *reinterpret_cast<const f2_t*>(&ta[0]);
      rj[h] = *reinterpret_cast<const f2_t*>(&td[0]);
      pw[h] = u[h];
    }
#pragma unroll
    for (int j = 1; j <= deg; ++j) {
      const f2_t a = *reinterpret_cast<const f2_t*>(&ta[j]), d = *reinterpret_cast<const f2_t*>(&td[j]);
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        ry[h] = f2_fma(ry[h], v[h], f2_mul(a, pw[h]));
        if (j < deg) {
          rj[h] = f2_fma(rj[h], v[h], f2_mul(d, pw[h]));
          pw[h] = f2_mul(pw[h], u[h]);
        }
      }
    }
    float ys[4], jj[4];
    f2_unpack(ry[0], ys[0], ys[1]);
    f2_unpack(ry[1], ys[2], ys[3]);
    f2_unpack(rj[0], jj[0], jj[1]);
    f2_unpack(rj[1], jj[2], jj[3]);
    const float ld[4] = {__logf(jj[0]), __logf(jj[1]), __logf(jj[2]), __logf(jj[3])};
    if (MODE == 0) {
      if (out_y) st_stream_f4(reinterpret_cast<float4*>(out_y) + q, make_float4(ys[0], ys[1], ys[2], ys[3]));
      if (out_ld) st_stream_f4(reinterpret_cast<float4*>(out_ld) + q, make_float4(ld[0], ld[1], ld[2], ld[3]));
    } else {
      float lp[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float z = s4 * (ys[i] + b4);
        lp[i] = fmaf(-0.5f * z, z, lconst + pre_ld[i] + ld[i]);
      }
      st_stream_f4(reinterpret_cast<float4*>(out_ld) + q, make_float4(lp[0], lp[1], lp[2], lp[3]));
    }
  }
  // tail (n % 4 elements), scalar
  const int64_t t = (nquads << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) {
    float xi = x[t], pl = 0.f;
    if (MODE == 1) softclip01(s1 * (xi + b1), 1.5f, kk, logk, xi, pl);
    const float uu = bern_clip(xi), vv = 1.f - uu;
    float y = ta[0].x, J = td[0].x, p = uu;
#pragma unroll
    for (int j = 1; j <= deg; ++j) {
      y = fmaf(y, vv, ta[j].x * p);
      if (j < deg) {
        J = fmaf(J, vv, td[j].x * p);
        p *= uu;
      }
    }
    if (MODE == 0) {
      if (out_y) out_y[t] = y;
      if (out_ld) out_ld[t] = logf(J);
    } else {
      const float z = s4 * (y + b4);
      out_ld[t] = fmaf(-0.5f * z, z, lconst + pl + logf(J));
    }
  }
}

// (A four-monomial folding for 16 < M <= 32 — half the registers of bern_eval_quad, four CTAs per SM instead of two, 1.5
// instead of 1.25 operations per coefficient — was measured at 106.9 us against 105.1 us for the eight-fold kernel at
// M = 32, 2^24 elements: at these orders the kernel is bound by the packed-FMA rate, not by occupancy or HBM; removed.)
template <int MODE>
static int bern_small_launch(int M, const float* theta, const float* scal, const float* x, float* y, float* ld, int64_t n,
                             cudaStream_t s) {
#define FC_BERN_CASE(K)                                                                                        \
  case K: {                                                                                                    \
    const int grid = grid_resident(bern_small_kernel<K, MODE>, kBernThreads, 0, (n + 3) / 4, kBernThreads);    \
    bern_small_kernel<K, MODE><<<grid, kBernThreads, 0, s>>>(theta, scal, x, y, ld, n);                        \
    break;                                                                                                     \
  }
  switch (M) {
    FC_BERN_CASE(2) FC_BERN_CASE(3) FC_BERN_CASE(4) FC_BERN_CASE(5) FC_BERN_CASE(6) FC_BERN_CASE(7) FC_BERN_CASE(8)
    FC_BERN_CASE(9) FC_BERN_CASE(10) FC_BERN_CASE(11) FC_BERN_CASE(12) FC_BERN_CASE(13) FC_BERN_CASE(14) FC_BERN_CASE(15)
    FC_BERN_CASE(16)
    default:
      return 1;  // not a small order
  }
#undef FC_BERN_CASE
  FC_LAUNCH_CHECK("bern_small_kernel");
  return FLOWCUDA_OK;
}
constexpr int kBernSmallMax = 16;

// Per-element theta rows (maf_mwe.py:103-107): theta[n, M], x[n].  Binomials are uniform.
__global__ void __launch_bounds__(kBernThreads)
bern_rows_kernel(const float* __restrict__ theta, int M, int64_t theta_stride, const float* __restrict__ x,
                 float* __restrict__ out_y, float* __restrict__ out_ld, int64_t n) {
  __shared__ float binA[kBernMaxOrder], binD[kBernMaxOrder];
  const int deg = M - 1;
  if (threadIdx.x == 0) {
    double ca = 1.0, cd = 1.0;
    for (int j = 0; j <= deg; ++j) {
      binA[j] = (float)ca;
      binD[j] = j < deg ? (float)((double)deg * cd) : 0.f;
      ca = ca * (double)(deg - j) / (double)(j + 1);
      if (j < deg - 1) cd = cd * (double)(deg - 1 - j) / (double)(j + 1);
    }
  }
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float* th = theta + i * theta_stride;
    float u = bern_clip(x[i]), v = 1.f - u;
    float t0 = th[0], t1 = th[1];
    float y = t0 * binA[0], J = (t1 - t0) * binD[0], pw = u;
    for (int j = 1; j < deg; ++j) {
      t0 = t1;
      t1 = th[j + 1];
      y = fmaf(y, v, t0 * binA[j] * pw);
      J = fmaf(J, v, (t1 - t0) * binD[j] * pw);
      pw *= u;
    }
    y = fmaf(y, v, t1 * binA[deg] * pw);
    if (out_y) out_y[i] = y;
    if (out_ld) out_ld[i] = logf(J);
  }
}

__global__ void __launch_bounds__(256)
bern_constrain_kernel(const float* __restrict__ raw, float* __restrict__ out, int64_t rows, int M) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  const float* p = raw + r * M;
  float* o = out + r * M;
  float acc = p[0];  // cumsum of [θ0, softplus(θ_k)+1e-4 ...] (BernsteinBijector.py:60-71)
  o[0] = acc;
  for (int k = 1; k < M; ++k) {
    acc += softplus_f(p[k]) + 1e-4f;
    o[k] = acc;
  }
}

// inverse, step 1: y_fit[e][p] = forward(theta_e, x_fit[p]), x_fit = linspace(1e-7, 1-1e-7, P)
__global__ void __launch_bounds__(256)
bern_grid_kernel(const float* __restrict__ theta, int M, int64_t E, int P, float* __restrict__ yfit) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= E * P) return;
  int64_t e = i / P;
  int p = (int)(i - e * P);
  const float* th = theta + e * M;
  const double lo = 1e-7, hi = 1.0 - 1e-7;
  float xf = (float)(lo + (hi - lo) * (double)p / (double)(P - 1));
  float u = bern_clip(xf), v = 1.f - u;
  const int deg = M - 1;
  double c = 1.0;
  float y = th[0], pw = u;
  for (int j = 1; j <= deg; ++j) {
    c = c * (double)(deg - j + 1) / (double)j;
    y = fmaf(y, v, th[j] * (float)c * pw);
    pw *= u;
  }
  yfit[i] = y;
}

// inverse, step 2: x[e][n] = interp(y[n][e], y_fit[e], x_fit)   (BernsteinBijector.py:104-110)
__global__ void __launch_bounds__(256)
bern_interp_kernel(const float* __restrict__ yfit, int64_t E, int P, const float* __restrict__ y,
                   float* __restrict__ x, int64_t n) {
  const double lo = 1e-7, hi = 1.0 - 1e-7;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    for (int64_t e = 0; e < E; ++e) {
      const float* f = yfit + e * P;
      float yq = y[i * E + e];
      // np.interp: idx = searchsorted(xp, xq, side='right'); clamp outside
      int a = 0, b = P;  // first index with f[idx] > yq
      while (a < b) {
        int m = (a + b) >> 1;
        if (__ldg(f + m) <= yq) a = m + 1; else b = m;
      }
      float r;
      if (a == 0) r = (float)lo;
      else if (a >= P) r = (float)hi;
      else {
        float f0 = __ldg(f + a - 1), f1 = __ldg(f + a);
        float x0 = (float)(lo + (hi - lo) * (double)(a - 1) / (double)(P - 1));
        float x1 = (float)(lo + (hi - lo) * (double)a / (double)(P - 1));
        float df = f1 - f0;
        r = df > 0.f ? x0 + (yq - f0) / df * (x1 - x0) : x0;
      }
      x[e * n + i] = r;
    }
  }
}

}  // namespace flowcuda

using namespace flowcuda;

extern "C" {

int flowcuda_bernstein_constrain_thetas(const float* theta_raw, float* theta, int64_t rows, int order, void* stream) {
  FC_REQUIRE(theta_raw && theta && rows >= 0 && order >= 1, "constrain_thetas: bad arguments");
  if (rows == 0) return FLOWCUDA_OK;
  bern_constrain_kernel<<<(int)((rows + 255) / 256), 256, 0, (cudaStream_t)stream>>>(theta_raw, theta, rows, order);
  FC_LAUNCH_CHECK("bern_constrain_kernel");
  return FLOWCUDA_OK;
}

int flowcuda_bernstein_forward_and_log_det(const float* theta, int order, int64_t theta_stride, const float* x,
                                           float* y, float* log_det, int64_t n, void* stream) {
  if (n == 0) return FLOWCUDA_OK;
  FC_REQUIRE(theta && x && (y || log_det), "bernstein_forward: NULL pointer");
  FC_REQUIRE(order >= 2 && order <= kBernMaxOrder, "bernstein_forward: order=%d outside [2,%d]", order, kBernMaxOrder);
  FC_REQUIRE(theta_stride == 0 || theta_stride >= order, "bernstein_forward: theta_stride must be 0 or >= order");
  FC_REQUIRE(n >= 0, "bernstein_forward: negative n");
  if (n == 0) return FLOWCUDA_OK;
  cudaStream_t s = (cudaStream_t)stream;
  if (theta_stride == 0) {
    FC_REQUIRE((((uintptr_t)x | (uintptr_t)y | (uintptr_t)log_det) & 15) == 0, "bernstein_forward: buffers must be 16-byte aligned");
    if (order <= kBernSmallMax) return bern_small_launch<0>(order, theta, nullptr, x, y, log_det, n, s);
    int grid = grid_resident(bern_shared_kernel<0>, kBernThreads, 0, (n + 3) / 4, kBernThreads);
    bern_shared_kernel<0><<<grid, kBernThreads, 0, s>>>(theta, order, nullptr, x, y, log_det, n);
  } else {
    bern_rows_kernel<<<grid_for(n, kBernThreads, 8), kBernThreads, 0, s>>>(theta, order, theta_stride, x, y, log_det, n);
  }
  FC_LAUNCH_CHECK("bernstein forward kernel");
  return FLOWCUDA_OK;
}

int flowcuda_bernstein_flow_log_prob(const float* theta, int order, const float* scal, const float* x, float* log_prob,
                                     int64_t n, void* stream) {
  if (n == 0) return FLOWCUDA_OK;
  FC_REQUIRE(theta && scal && x && log_prob, "bernstein_flow_log_prob: NULL pointer");
  FC_REQUIRE(order >= 2 && order <= kBernMaxOrder, "bernstein_flow_log_prob: order=%d outside [2,%d]", order, kBernMaxOrder);
  FC_REQUIRE((((uintptr_t)x | (uintptr_t)log_prob) & 15) == 0, "bernstein_flow_log_prob: buffers must be 16-byte aligned");
  FC_REQUIRE(n >= 0, "bernstein_flow_log_prob: negative n");
  if (n == 0) return FLOWCUDA_OK;
  if (order <= kBernSmallMax) return bern_small_launch<1>(order, theta, scal, x, nullptr, log_prob, n, (cudaStream_t)stream);
  int grid = grid_resident(bern_shared_kernel<1>, kBernThreads, 0, (n + 3) / 4, kBernThreads);
  bern_shared_kernel<1><<<grid, kBernThreads, 0, (cudaStream_t)stream>>>(theta, order, scal, x, nullptr, log_prob, n);
  FC_LAUNCH_CHECK("bern_shared_kernel<1>");
  return FLOWCUDA_OK;
}

int flowcuda_bernstein_inverse(const float* theta, int order, int64_t n_events, int n_points, const float* y, float* x,
                               int64_t n, float* workspace, void* stream) {
  if (n == 0) return FLOWCUDA_OK;
  FC_REQUIRE(theta && y && x && workspace, "bernstein_inverse: NULL pointer");
  FC_REQUIRE(order >= 2 && order <= kBernMaxOrder, "bernstein_inverse: order=%d outside [2,%d]", order, kBernMaxOrder);
  FC_REQUIRE(n_events >= 1 && n_points >= 2 && n >= 0, "bernstein_inverse: bad sizes");
  if (n == 0) return FLOWCUDA_OK;
  cudaStream_t s = (cudaStream_t)stream;
  int64_t tot = n_events * n_points;
  bern_grid_kernel<<<(int)((tot + 255) / 256), 256, 0, s>>>(theta, order, n_events, n_points, workspace);
  FC_LAUNCH_CHECK("bern_grid_kernel");
  bern_interp_kernel<<<grid_for(n, 256, 8), 256, 0, s>>>(workspace, n_events, n_points, y, x, n);
  FC_LAUNCH_CHECK("bern_interp_kernel");
  return FLOWCUDA_OK;
}

}  // extern "C"
