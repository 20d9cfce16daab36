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
// exp(±ls) is FMUL + bare MUFU.EX2 (relative error ~1e-7·max(1,|ls|), inside every parity budget for the
// log-scales a flow produces; libm's expf costs ~10 instructions per element in epilogues that are issue-bound).
template <bool INV>
__device__ __forceinline__ void affine_apply(float shift, float ls, float v, float& out, float& ld) {
  if (INV) {
    out = (v - shift) * ex2_ftz(-kLog2e * ls);
    ld = -ls;
  } else {
    out = fmaf(v, ex2_ftz(kLog2e * ls), shift);
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

// ---- MUFU-based variant for the tensor-core toy path (coupling_toy_tc.cu) -------------------------
// Same spline, restated to cut the instruction count of the epilogue (ncu: the RQS was 680 of the
// 1340 thread-instructions per sample and layer, mostly the range-reduction code around 24 expf,
// 3 logf and 5 IEEE divisions):
//   * p(i), i < 2K arrive PRE-MULTIPLIED by log2(e) (folded into the unscale FMA of the epilogue), so a
//     softmax term is FADD + MUFU.EX2;
//   * the softmax denominators come out of the knot prefix sums, every knot is one FFMA of its prefix sum
//     (independent, not a serial chain); koff[i] = lo + i*min_bin_size is read from the constant bank;
//   * divisions are MUFU.RCP + FMUL (1 ulp), the three logs of the log-det are one MUFU.LG2 of the ratio;
//   * the two knot slopes keep log1pf (their relative accuracy enters the log-det directly).
__device__ __forceinline__ float lg2_approx(float x) {
  float r;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float rqs_slope_mufu(float u) {
  const float x = u + kRqsSlopeOffset;
  return fmaxf(x, 0.f) + log1pf(ex2_ftz(-fabsf(x) * kLog2e)) + kRqsMinSlope;
}

template <bool INV, int K, typename Get>
__device__ __forceinline__ void rqs_apply_mufu(Get p, const float* __restrict__ koff, float lo, float hi, float v,
                                               float& out, float& ld) {
  const float scale = (hi - lo) - (float)K * kRqsMinBin;
  float mw = p(0), mh = p(K);
#pragma unroll
  for (int i = 1; i < K; ++i) {
    mw = fmaxf(mw, p(i));
    mh = fmaxf(mh, p(K + i));
  }
  float ew[K], eh[K];  // inclusive prefix sums of the unnormalised bin sizes
  float aw = 0.f, ah = 0.f;
#pragma unroll
  for (int i = 0; i < K; ++i) {
    aw += ex2_ftz(p(i) - mw);
    ah += ex2_ftz(p(K + i) - mh);
    ew[i] = aw;
    eh[i] = ah;
  }
  const float cw = scale * rcp_approx(aw), ch = scale * rcp_approx(ah);
  // bin = last i whose left knot is <= v (bin 0 when v lies left of the range)
  float xl = lo, yl = lo;
  float xr = (K == 1) ? hi : fmaf(ew[0], cw, koff[1]);
  float yr = (K == 1) ? hi : fmaf(eh[0], ch, koff[1]);
  float ul = p(2 * K), ur = p(2 * K + 1);
  float xi = xr, yi = yr;
#pragma unroll
  for (int i = 1; i < K; ++i) {
    const float xn = (i == K - 1) ? hi : fmaf(ew[i], cw, koff[i + 1]);
    const float yn = (i == K - 1) ? hi : fmaf(eh[i], ch, koff[i + 1]);
    const bool take = v >= (INV ? yi : xi);
    xl = take ? xi : xl;
    xr = take ? xn : xr;
    yl = take ? yi : yl;
    yr = take ? yn : yr;
    ul = take ? p(2 * K + i) : ul;
    ur = take ? p(2 * K + i + 1) : ur;
    xi = xn;
    yi = yn;
  }
  const float sl_l = rqs_slope_mufu(ul), sl_r = rqs_slope_mufu(ur);
  const float w = xr - xl, h = yr - yl;
  const float rw = rcp_approx(w);
  const float s = h * rw;
  const float t = sl_r + sl_l - 2.f * s;
  float z;
  if (INV) {
    float om = fminf(fmaxf((v - yl) * rcp_approx(h), 0.f), 1.f);
    float c = -s * om;
    float b = sl_l - t * om;
    float a = s - b;
    z = fminf(fmaxf(-2.f * c * rcp_approx(b + sqrtf(b * b - 4.f * a * c)), 0.f), 1.f);
    out = fmaf(w, z, xl);
  } else {
    z = fminf(fmaxf((v - xl) * rw, 0.f), 1.f);
  }
  const float sq_z = z * z;
  const float z1mz = z - sq_z;
  const float omz = 1.f - z;
  const float rden = rcp_approx(fmaf(t, z1mz, s));
  const float num = fmaf(sl_r, sq_z, fmaf(2.f * s, z1mz, sl_l * omz * omz));
  const float sr = s * rden;
  const float fld = kLn2 * lg2_approx(sr * sr * num);
  if (INV) {
    ld = -fld;
  } else {
    out = fmaf(h * fmaf(s, sq_z, sl_l * z1mz), rden, yl);
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

// Two independent splines evaluated in lockstep (same arithmetic, same order per spline as rqs_apply_mufu: results are
// bit-identical).  The D=784 spline epilogue is a dependent chain of ~360 instructions per (sample, dim) with one or two
// warps per scheduler: issue-active was 27 % and the epilogue, not the MMAs, bounded the kernel.  Writing every step for
// both splines side by side gives the scheduler two independent chains per thread; the rare linear tails are patched up
// after the common path so that they do not split it into basic blocks.  p0 / p1: the 3K+1 conditioner outputs of each.
template <bool INV, int K>
__device__ __forceinline__ void rqs_apply_mufu_x2(const float* __restrict__ p0, const float* __restrict__ p1,
                                                  const float* __restrict__ koff, float lo, float hi, const float (&v)[2],
                                                  float (&out)[2], float (&ld)[2]) {
  const float scale = (hi - lo) - (float)K * kRqsMinBin;
  const float* p[2] = {p0, p1};
  float mw[2], mh[2];
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    mw[u] = p[u][0];
    mh[u] = p[u][K];
  }
#pragma unroll
  for (int i = 1; i < K; ++i) {
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      mw[u] = fmaxf(mw[u], p[u][i]);
      mh[u] = fmaxf(mh[u], p[u][K + i]);
    }
  }
  float ew[2][K], eh[2][K];
  float aw[2] = {0.f, 0.f}, ah[2] = {0.f, 0.f};
#pragma unroll
  for (int i = 0; i < K; ++i) {
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      aw[u] += ex2_ftz(p[u][i] - mw[u]);
      ah[u] += ex2_ftz(p[u][K + i] - mh[u]);
      ew[u][i] = aw[u];
      eh[u][i] = ah[u];
    }
  }
  float cw[2], ch[2], xl[2], yl[2], xr[2], yr[2], ul[2], ur[2], xi[2], yi[2];
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    cw[u] = scale * rcp_approx(aw[u]);
    ch[u] = scale * rcp_approx(ah[u]);
    xl[u] = lo;
    yl[u] = lo;
    xr[u] = (K == 1) ? hi : fmaf(ew[u][0], cw[u], koff[1]);
    yr[u] = (K == 1) ? hi : fmaf(eh[u][0], ch[u], koff[1]);
    ul[u] = p[u][2 * K];
    ur[u] = p[u][2 * K + 1];
    xi[u] = xr[u];
    yi[u] = yr[u];
  }
#pragma unroll
  for (int i = 1; i < K; ++i) {
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const float xn = (i == K - 1) ? hi : fmaf(ew[u][i], cw[u], koff[i + 1]);
      const float yn = (i == K - 1) ? hi : fmaf(eh[u][i], ch[u], koff[i + 1]);
      const bool take = v[u] >= (INV ? yi[u] : xi[u]);
      xl[u] = take ? xi[u] : xl[u];
      xr[u] = take ? xn : xr[u];
      yl[u] = take ? yi[u] : yl[u];
      yr[u] = take ? yn : yr[u];
      ul[u] = take ? p[u][2 * K + i] : ul[u];
      ur[u] = take ? p[u][2 * K + i + 1] : ur[u];
      xi[u] = xn;
      yi[u] = yn;
    }
  }
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    const float sl_l = rqs_slope_mufu(ul[u]), sl_r = rqs_slope_mufu(ur[u]);
    const float w = xr[u] - xl[u], h = yr[u] - yl[u];
    const float rw = rcp_approx(w);
    const float s = h * rw;
    const float t = sl_r + sl_l - 2.f * s;
    float z;
    if (INV) {
      float om = fminf(fmaxf((v[u] - yl[u]) * rcp_approx(h), 0.f), 1.f);
      float c = -s * om;
      float b = sl_l - t * om;
      float a = s - b;
      z = fminf(fmaxf(-2.f * c * rcp_approx(b + sqrtf(b * b - 4.f * a * c)), 0.f), 1.f);
      out[u] = fmaf(w, z, xl[u]);
    } else {
      z = fminf(fmaxf((v[u] - xl[u]) * rw, 0.f), 1.f);
    }
    const float sq_z = z * z;
    const float z1mz = z - sq_z;
    const float omz = 1.f - z;
    const float rden = rcp_approx(fmaf(t, z1mz, s));
    const float num = fmaf(sl_r, sq_z, fmaf(2.f * s, z1mz, sl_l * omz * omz));
    const float sr = s * rden;
    const float fld = kLn2 * lg2_approx(sr * sr * num);
    if (INV) {
      ld[u] = -fld;
    } else {
      out[u] = fmaf(h * fmaf(s, sq_z, sl_l * z1mz), rden, yl[u]);
      ld[u] = fld;
    }
  }
  // linear tails outside [lo, hi] (rare: one branch for the pair)
  if (v[0] <= lo || v[0] >= hi || v[1] <= lo || v[1] >= hi) {
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (v[u] <= lo) {
        float s0 = rqs_slope(p[u][2 * K]);
        out[u] = INV ? (v[u] - lo) / s0 + lo : fmaf(v[u] - lo, s0, lo);
        ld[u] = INV ? -logf(s0) : logf(s0);
      } else if (v[u] >= hi) {
        float sK = rqs_slope(p[u][3 * K]);
        out[u] = INV ? (v[u] - hi) / sK + hi : fmaf(v[u] - hi, sK, hi);
        ld[u] = INV ? -logf(sK) : logf(sK);
      }
    }
  }
}

// Independent(Normal(0,1)) / Independent(Uniform(0,1)) contribution of one coordinate.
__device__ __forceinline__ float base_term(int base, float z) {
  if (base == FLOWCUDA_BASE_NORMAL) return fmaf(-0.5f * z, z, -kHalfLog2Pi);
  return (z >= 0.f && z <= 1.f) ? 0.f : -INFINITY;
}

}  // namespace flowcuda
