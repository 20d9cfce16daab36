// rqs_grad.cuh — hand-written reverse mode of the inner bijectors (forward direction, the one the
// density / training path uses): given the conditioner outputs o[] of one (sample, dim), the input x,
// and the adjoints ybar = dL/dy, ldbar = dL/dlogdet, produce y, logdet, xbar = dL/dx and obar[] = dL/do.
// Follows transforms.cuh (distrax.RationalQuadraticSpline) operation by operation.
#pragma once
#include "transforms.cuh"

namespace flowcuda {

__device__ __forceinline__ float sigmoid_f(float x) { return 1.f / (1.f + expf(-x)); }

// FAST: exp / log / division / sigmoid on the MUFU approximations (ex2, lg2, rcp: ~1e-7 relative) instead of libm and IEEE
// division — the adjoint kernel of the D=784 training step is instruction-bound, and roughly 40 % of this function's
// instructions were range reduction and division slow-path code.  The gradient budget (1e-4 of the largest entry) is
// three orders of magnitude above the difference; the value outputs y / ld of that kernel are not used at all.
template <bool FAST> __device__ __forceinline__ float rg_exp(float x) { return FAST ? __expf(x) : expf(x); }
template <bool FAST> __device__ __forceinline__ float rg_log(float x) { return FAST ? __logf(x) : logf(x); }
template <bool FAST> __device__ __forceinline__ float rg_div(float a, float b) { return FAST ? __fdividef(a, b) : a / b; }
template <bool FAST> __device__ __forceinline__ float rg_sigmoid(float x) { return FAST ? __fdividef(1.f, 1.f + __expf(-x)) : sigmoid_f(x); }

template <int K, bool FAST = false>
__device__ __forceinline__ void rqs_fwd_bwd(const float* o, float lo, float hi, float x, float ybar, float ldbar,
                                            float& y, float& ld, float& xbar, float* obar) {
  const float scale = (hi - lo) - (float)K * kRqsMinBin;
  float mw = -INFINITY, mh = -INFINITY;
#pragma unroll
  for (int i = 0; i < K; ++i) {
    mw = fmaxf(mw, o[i]);
    mh = fmaxf(mh, o[K + i]);
  }
  float pw[K], ph[K], zw = 0.f, zh = 0.f;
#pragma unroll
  for (int i = 0; i < K; ++i) {
    pw[i] = rg_exp<FAST>(o[i] - mw);
    ph[i] = rg_exp<FAST>(o[K + i] - mh);
    zw += pw[i];
    zh += ph[i];
  }
  const float iw = rg_div<FAST>(1.f, zw), ih = rg_div<FAST>(1.f, zh);
  float xk = lo, yk = lo, xl = lo, xr = lo, yl = lo, yr = lo, ul = 0.f, ur = 0.f;
  int bin = 0;
#pragma unroll
  for (int i = 0; i < K; ++i) {
    pw[i] *= iw;  // softmax probabilities
    ph[i] *= ih;
    const float w = fmaf(pw[i], scale, kRqsMinBin), h = fmaf(ph[i], scale, kRqsMinBin);
    const float xn = (i == K - 1) ? hi : xk + w, yn = (i == K - 1) ? hi : yk + h;
    const bool take = (i == 0) || (x >= xk);
    bin = take ? i : bin;
    xl = take ? xk : xl; xr = take ? xn : xr;
    yl = take ? yk : yl; yr = take ? yn : yr;
    ul = take ? o[2 * K + i] : ul;
    ur = take ? o[2 * K + i + 1] : ur;
    xk = xn; yk = yn;
  }
#pragma unroll
  for (int i = 0; i < 3 * K + 1; ++i) obar[i] = 0.f;
  if (x <= lo || x >= hi) {  // linear tails: only the boundary slope matters
    const bool left = x <= lo;
    const float u = left ? o[2 * K] : o[3 * K];
    const float edge = left ? lo : hi;
    const float s0 = rqs_slope(u);
    y = fmaf(x - edge, s0, edge);
    ld = rg_log<FAST>(s0);
    xbar = ybar * s0;
    const float sbar = ybar * (x - edge) + rg_div<FAST>(ldbar, s0);
    const float g = sbar * rg_sigmoid<FAST>(u + kRqsSlopeOffset);
    if (left) obar[2 * K] = g; else obar[3 * K] = g;
    return;
  }
  const float sl = rqs_slope(ul), sr = rqs_slope(ur);
  const float w = xr - xl, h = yr - yl;
  const float rw = rg_div<FAST>(1.f, w);
  const float s = FAST ? h * rw : h / w;
  const float t = sr + sl - 2.f * s;
  const float z = fminf(fmaxf(FAST ? (x - xl) * rw : (x - xl) / w, 0.f), 1.f);
  const float q = z - z * z, omz = 1.f - z;
  const float A = s * z * z + sl * q;
  const float den = fmaf(t, q, s);
  const float B = sr * z * z + 2.f * s * q + sl * omz * omz;
  const float rden = rg_div<FAST>(1.f, den);
  y = FAST ? fmaf(h * A, rden, yl) : yl + h * A / den;
  ld = FAST ? 0.f : 2.f * logf(s) + logf(B) - 2.f * logf(den);  // FAST callers do not use the value outputs
  // ---- reverse
  const float Abar = FAST ? ybar * h * rden : ybar * h / den;
  float hbar = FAST ? ybar * A * rden : ybar * A / den;
  const float denbar = FAST ? (-ybar * h * A * rden - 2.f * ldbar) * rden : -ybar * h * A / (den * den) - 2.f * ldbar / den;
  const float ylbar = ybar;
  const float Bbar = rg_div<FAST>(ldbar, B);
  float sbar = rg_div<FAST>(2.f * ldbar, s) + Bbar * 2.f * q + Abar * z * z + denbar;
  float srbar = Bbar * z * z;
  float slbar = Bbar * omz * omz + Abar * q;
  const float tbar = denbar * q;
  srbar += tbar; slbar += tbar; sbar -= 2.f * tbar;
  const float qbar = Bbar * 2.f * s + Abar * sl + denbar * t;
  const float zbar = Abar * 2.f * s * z + Bbar * (2.f * sr * z - 2.f * sl * omz) + qbar * (1.f - 2.f * z);
  hbar += FAST ? sbar * rw : sbar / w;
  float wbar = FAST ? -sbar * s * rw : -sbar * s / w;
  xbar = FAST ? zbar * rw : zbar / w;
  const float xlbar = -xbar;
  wbar -= FAST ? zbar * z * rw : zbar * z / w;
  // knots -> normalised bins: xl = lo + sum_{j<bin} w_j, w = w_bin (sum_j w_j == hi - lo identically)
  float dotw = 0.f, doth = 0.f;
#pragma unroll
  for (int i = 0; i < K; ++i) {
    const float gw = (i < bin ? xlbar : 0.f) + (i == bin ? wbar : 0.f);
    const float gh = (i < bin ? ylbar : 0.f) + (i == bin ? hbar : 0.f);
    dotw += pw[i] * gw;
    doth += ph[i] * gh;
  }
#pragma unroll
  for (int i = 0; i < K; ++i) {
    const float gw = (i < bin ? xlbar : 0.f) + (i == bin ? wbar : 0.f);
    const float gh = (i < bin ? ylbar : 0.f) + (i == bin ? hbar : 0.f);
    obar[i] = scale * pw[i] * (gw - dotw);
    obar[K + i] = scale * ph[i] * (gh - doth);
  }
  const float gl = slbar * rg_sigmoid<FAST>(ul + kRqsSlopeOffset), gr = srbar * rg_sigmoid<FAST>(ur + kRqsSlopeOffset);
#pragma unroll
  for (int i = 0; i <= K; ++i) obar[2 * K + i] = (i == bin ? gl : 0.f) + (i == bin + 1 ? gr : 0.f);
}

// ScalarAffine forward: y = x e^{ls} + shift, ld = ls.
__device__ __forceinline__ void affine_fwd_bwd(const float* o, float x, float ybar, float ldbar, float& y, float& ld,
                                               float& xbar, float* obar) {
  const float e = expf(o[1]);
  y = fmaf(x, e, o[0]);
  ld = o[1];
  xbar = ybar * e;
  obar[0] = ybar;
  obar[1] = ybar * x * e + ldbar;
}

}  // namespace flowcuda
