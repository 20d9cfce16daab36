// transforms.cuh — inner bijectors of distrax.MaskedCoupling as per-(sample, dim) device
// functions: ScalarAffine and RationalQuadraticSpline, and the base
// densities (A.1).  `Get` is any callable int -> float returning conditioner output i of
// this (sample, dim); loops over the bin count are written with selects only, so that the
// parameters may live in registers (TMEM epilogue) or shared memory alike.
#pragma once
#include "common.cuh"

namespace flowcuda {

constexpr float kRqsMinBin = 1e-4f;
constexpr float kRqsMinSlope = 1e-4f;
// log(exp(1 - 1e-4) - 1): softplus offset so that an unconstrained slope of 0 maps to 1
constexpr float kRqsSlopeOffset = 0.54116539f;

__device__ __forceinline__ float rqs_slope(float u) { return softplus_f(u + kRqsSlopeOffset) + kRqsMinSlope; }

// ScalarAffine.  INV=false: y = x·exp(ls)+shift, ld = ls;  INV=true: x = (y-shift)·exp(-ls), ld = -ls.
template <bool INV>
__device__ __forceinline__ void affine_apply(float shift, float ls, float v, float& out, float& ld) {
  if (INV) {
    out = (v - shift) * expf(-ls);
    ld = -ls;
  } else {
    out = fmaf(v, expf(ls), shift);
    ld = ls;
  }
}

// distrax.RationalQuadraticSpline(params[3K+1], range_min=lo, range_max=hi), unconstrained
// boundary slopes, min_bin_size = min_knot_slope = 1e-4.  KT > 0 fixes K at compile time.
template <bool INV, int KT, typename Get>
__device__ __forceinline__ void rqs_apply(Get p, int Krt, float lo, float hi, float v, float& out, float& ld) {
  const int K = KT > 0 ? KT : Krt;
  const float scale = (hi - lo) - (float)K * kRqsMinBin;
  float mw = -INFINITY, mh = -INFINITY;
#pragma unroll
  for (int i = 0; i < K; ++i) {
    {
      mw = fmaxf(mw, p(i));
      mh = fmaxf(mh, p(K + i));
    }
  }
  float zw = 0.f, zh = 0.f;
#pragma unroll
  for (int i = 0; i < K; ++i) {
    {
      zw += expf(p(i) - mw);
      zh += expf(p(K + i) - mh);
    }
  }
  const float cw = scale / zw, ch = scale / zh;
  // walk the knots; bin = last i whose left knot is <= v (bin 0 when v lies left of the range)
  float xk = lo, yk = lo;
  float xl = lo, xr = lo, yl = lo, yr = lo, ul = 0.f, ur = 0.f;
#pragma unroll
  for (int i = 0; i < K; ++i) {
    {
      float w = fmaf(expf(p(i) - mw), cw, kRqsMinBin);
      float h = fmaf(expf(p(K + i) - mh), ch, kRqsMinBin);
      float xn = (i == K - 1) ? hi : xk + w;
      float yn = (i == K - 1) ? hi : yk + h;
      bool take = (i == 0) || (v >= (INV ? yk : xk));
      xl = take ? xk : xl;
      xr = take ? xn : xr;
      yl = take ? yk : yl;
      yr = take ? yn : yr;
      ul = take ? p(2 * K + i) : ul;
      ur = take ? p(2 * K + i + 1) : ur;
      xk = xn;
      yk = yn;
    }
  }
  const float sl_l = rqs_slope(ul), sl_r = rqs_slope(ur);
  const float w = xr - xl, h = yr - yl;
  const float s = h / w;
  const float t = sl_r + sl_l - 2.f * s;
  float z;
  if (INV) {
    float om = fminf(fmaxf((v - yl) / h, 0.f), 1.f);
    float c = -s * om;
    float b = sl_l - t * om;
    float a = s - b;
    z = fminf(fmaxf(-2.f * c / (b + sqrtf(b * b - 4.f * a * c)), 0.f), 1.f);
    out = fmaf(w, z, xl);
  } else {
    z = fminf(fmaxf((v - xl) / w, 0.f), 1.f);
  }
  const float sq_z = z * z;
  const float z1mz = z - sq_z;
  const float omz = 1.f - z;
  const float den = fmaf(t, z1mz, s);
  const float fld = 2.f * logf(s) + logf(sl_r * sq_z + 2.f * s * z1mz + sl_l * omz * omz) - 2.f * logf(den);
  if (INV) {
    ld = -fld;
  } else {
    out = yl + h * (s * sq_z + sl_l * z1mz) / den;
    ld = fld;
  }
  // linear tails outside [lo, hi]
  if (v <= lo) {
    float s0 = rqs_slope(p(2 * K));
    out = INV ? (v - lo) / s0 + lo : fmaf(v - lo, s0, lo);
    ld = INV ? -logf(s0) : logf(s0);
  } else if (v >= hi) {
    float sK = rqs_slope(p(3 * K));
    out = INV ? (v - hi) / sK + hi : fmaf(v - hi, sK, hi);
    ld = INV ? -logf(sK) : logf(sK);
  }
}

// Independent(Normal(0,1)) / Independent(Uniform(0,1)) contribution of one coordinate.
__device__ __forceinline__ float base_term(int base, float z) {
  if (base == FLOWCUDA_BASE_NORMAL) return fmaf(-0.5f * z, z, -kHalfLog2Pi);
  return (z >= 0.f && z <= 1.f) ? 0.f : -INFINITY;
}

}  // namespace flowcuda
