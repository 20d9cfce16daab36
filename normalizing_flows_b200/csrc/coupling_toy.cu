// coupling_toy.cu — rows #2/#3/#6 at event_dim = 2: the whole distrax MaskedCoupling stack
// (src/real_nvp_toy_densities.py:74-110, src/maf_toy_densities.py:73-111) in ONE kernel,
// FP32 CUDA-core conditioner (FLOWCUDA_PRECISION_FP32).
//
// A CTA owns 256*S samples whose state (x0, x1, log-det) stays in registers for the whole
// stack; per flow layer the CTA stages that layer's conditioner once in shared memory
// (zero-padded to 64-wide, only the output columns the unmasked dimension needs) and every
// thread runs the MLP for its own samples with weights broadcast from shared memory.
// HBM traffic is x in and log_prob out.  The tensor-core variant lives in coupling_toy_tc.cu.
#include "common.cuh"
#include "transforms.cuh"

namespace flowcuda {

constexpr int kToyThreads = 256;
constexpr int kToyS = 4;    // samples per thread kept in registers across the stack
constexpr int kToyH = 64;   // padded hidden width

struct ToyArgs {
  const float* params;
  const float* in;     // [N,2]
  float* out_v;        // [N,2] or NULL
  float* out_lp;       // [N]   or NULL  (log_prob incl. base, or plain log-det if !with_base)
  int64_t n;
  int L, h1, h2, out_full;   // conditioner 2 -> h1 -> h2 -> out_full
  int layer_begin, layer_end;
  int with_base, base;
  float lo, hi;
};

// smem image of one layer (floats): w1[2][64] b1[64] W2[64][64] b2[64] W3[64][OP] b3[OP]
template <int OP>
struct ToyImage {
  static constexpr int w1 = 0, b1 = 128, W2 = 192, b2 = 192 + 4096, W3 = b2 + 64, b3 = W3 + 64 * OP, total = b3 + OP;
};

template <int OP>
__device__ void toy_stage_layer(float* sw, const ToyArgs& a, int l, int col0, int ncols) {
  using I = ToyImage<OP>;
  const int h1 = a.h1, h2 = a.h2, of = a.out_full;
  const size_t per_layer = (size_t)2 * h1 + h1 + (size_t)h1 * h2 + h2 + (size_t)h2 * of + of;
  const float* p = a.params + per_layer * l;
  const float* pW1 = p;
  const float* pb1 = pW1 + 2 * h1;
  const float* pW2 = pb1 + h1;
  const float* pb2 = pW2 + (size_t)h1 * h2;
  const float* pW3 = pb2 + h2;
  const float* pb3 = pW3 + (size_t)h2 * of;
  for (int i = threadIdx.x; i < I::total; i += blockDim.x) {
    float v = 0.f;
    if (i < I::b1) {
      int r = i >> 6, c = i & 63;
      if (c < h1) v = pW1[r * h1 + c];
    } else if (i < I::W2) {
      int c = i - I::b1;
      if (c < h1) v = pb1[c];
    } else if (i < I::b2) {
      int r = (i - I::W2) >> 6, c = (i - I::W2) & 63;
      if (r < h1 && c < h2) v = pW2[r * h2 + c];
    } else if (i < I::W3) {
      int c = i - I::b2;
      if (c < h2) v = pb2[c];
    } else if (i < I::b3) {
      int r = (i - I::W3) / OP, c = (i - I::W3) % OP;
      if (r < h2 && c < ncols) v = pW3[r * of + col0 + c];
    } else {
      int c = i - I::b3;
      if (c < ncols) v = pb3[col0 + c];
    }
    sw[i] = v;
  }
}

// TR: FLOWCUDA_TRANSFORM_*; KT: RQS bins; OP: padded needed-output count; INV: inner inverse
template <int TR, int KT, int OP, bool INV>
__global__ void __launch_bounds__(kToyThreads, 2) coupling_toy_kernel(ToyArgs a) {
  extern __shared__ __align__(16) float smem[];
  using I = ToyImage<OP>;
  float* sw = smem;
  float* hs = smem + ((I::total + 3) & ~3);  // [64][256] thread-private columns
  const int tid = threadIdx.x;
  constexpr int NEED = TR == FLOWCUDA_TRANSFORM_RQS ? 3 * KT + 1 : 2;

  const int64_t span = (int64_t)gridDim.x * kToyThreads * kToyS;
  for (int64_t base = (int64_t)blockIdx.x * kToyThreads * kToyS; base < a.n; base += span) {
    float x0[kToyS], x1[kToyS], ld[kToyS];
#pragma unroll
    for (int s = 0; s < kToyS; ++s) {
      int64_t idx = base + (int64_t)s * kToyThreads + tid;
      float2 v = idx < a.n ? ld_stream_f2(reinterpret_cast<const float2*>(a.in) + idx) : make_float2(0.f, 0.f);
      x0[s] = v.x;
      x1[s] = v.y;
      ld[s] = 0.f;
    }
    const int nl = a.layer_end - a.layer_begin;
    for (int it = 0; it < nl; ++it) {
      // data->base applies the LAST layer first (Inverse(Chain)); sampling the first layer first
      const int l = INV ? a.layer_begin + it : a.layer_end - 1 - it;
      const int t = l & 1;  // transformed dim: mask = (arange(2)%2==1) on even layers, negated on odd
      const int col0 = TR == FLOWCUDA_TRANSFORM_AFFINE_SCALAR ? 0 : t * NEED;
      __syncthreads();
      toy_stage_layer<OP>(sw, a, l, col0, NEED);
      __syncthreads();
      const float* w1 = sw + I::w1 + (1 - t) * 64;  // row of W1 for the visible dim
#pragma unroll 1
      for (int s = 0; s < kToyS; ++s) {
        const float xm = t ? x0[s] : x1[s];  // visible (masked-in) value
        const float xt = t ? x1[s] : x0[s];  // value to transform
        // Linear(2->h1)+relu on where(mask, x, 0): only the visible row of W1 contributes
#pragma unroll
        for (int k = 0; k < kToyH; k += 4) {
          float4 w = *reinterpret_cast<const float4*>(w1 + k);
          float4 b = *reinterpret_cast<const float4*>(sw + I::b1 + k);
          hs[(k + 0) * kToyThreads + tid] = fmaxf(fmaf(xm, w.x, b.x), 0.f);
          hs[(k + 1) * kToyThreads + tid] = fmaxf(fmaf(xm, w.y, b.y), 0.f);
          hs[(k + 2) * kToyThreads + tid] = fmaxf(fmaf(xm, w.z, b.z), 0.f);
          hs[(k + 3) * kToyThreads + tid] = fmaxf(fmaf(xm, w.w, b.w), 0.f);
        }
        float acc[kToyH];
#pragma unroll
        for (int j = 0; j < kToyH; j += 4) {
          float4 b = *reinterpret_cast<const float4*>(sw + I::b2 + j);
          acc[j] = b.x; acc[j + 1] = b.y; acc[j + 2] = b.z; acc[j + 3] = b.w;
        }
#pragma unroll 2
        for (int k = 0; k < kToyH; ++k) {
          const float hk = hs[k * kToyThreads + tid];
          const float4* wr = reinterpret_cast<const float4*>(sw + I::W2 + k * 64);
#pragma unroll
          for (int j = 0; j < kToyH / 4; ++j) {
            float4 w = wr[j];
            acc[4 * j] = fmaf(hk, w.x, acc[4 * j]);
            acc[4 * j + 1] = fmaf(hk, w.y, acc[4 * j + 1]);
            acc[4 * j + 2] = fmaf(hk, w.z, acc[4 * j + 2]);
            acc[4 * j + 3] = fmaf(hk, w.w, acc[4 * j + 3]);
          }
        }
#pragma unroll
        for (int j = 0; j < kToyH; ++j) hs[j * kToyThreads + tid] = fmaxf(acc[j], 0.f);
        float o[OP];
#pragma unroll
        for (int j = 0; j < OP; j += 4) {
          float4 b = *reinterpret_cast<const float4*>(sw + I::b3 + j);
          o[j] = b.x; o[j + 1] = b.y; o[j + 2] = b.z; o[j + 3] = b.w;
        }
#pragma unroll 2
        for (int k = 0; k < kToyH; ++k) {
          const float hk = hs[k * kToyThreads + tid];
          const float4* wr = reinterpret_cast<const float4*>(sw + I::W3 + k * OP);
#pragma unroll
          for (int j = 0; j < OP / 4; ++j) {
            float4 w = wr[j];
            o[4 * j] = fmaf(hk, w.x, o[4 * j]);
            o[4 * j + 1] = fmaf(hk, w.y, o[4 * j + 1]);
            o[4 * j + 2] = fmaf(hk, w.z, o[4 * j + 2]);
            o[4 * j + 3] = fmaf(hk, w.w, o[4 * j + 3]);
          }
        }
        float nv, l1;
        if (TR == FLOWCUDA_TRANSFORM_RQS) {
          rqs_apply<INV, KT>([&](int i) { return o[i]; }, KT, a.lo, a.hi, xt, nv, l1);
        } else {
          affine_apply<INV>(o[0], o[1], xt, nv, l1);
        }
        if (t) x1[s] = nv; else x0[s] = nv;
        ld[s] += l1;
      }
    }
#pragma unroll
    for (int s = 0; s < kToyS; ++s) {
      int64_t idx = base + (int64_t)s * kToyThreads + tid;
      if (idx < a.n) {
        if (a.out_v) st_stream_f2(reinterpret_cast<float2*>(a.out_v) + idx, make_float2(x0[s], x1[s]));
        if (a.out_lp) {
          float lp = ld[s];
          if (a.with_base) lp += base_term(a.base, x0[s]) + base_term(a.base, x1[s]);
          a.out_lp[idx] = lp;
        }
      }
    }
  }
}

template <int TR, int KT, int OP, bool INV>
static int toy_launch(const ToyArgs& a, cudaStream_t s) {
  using I = ToyImage<OP>;
  size_t smem = (((size_t)I::total + 3) & ~(size_t)3) * 4 + (size_t)kToyH * kToyThreads * 4;
  auto kern = coupling_toy_kernel<TR, KT, OP, INV>;
  FC_ENSURE_SMEM(kern, smem);
  int64_t need = (a.n + (int64_t)kToyThreads * kToyS - 1) / ((int64_t)kToyThreads * kToyS);
  int64_t cap = (int64_t)sm_count() * 2;  // 2 CTAs resident per SM
  int grid = (int)(need < 1 ? 1 : (need < cap ? need : cap));
  kern<<<grid, kToyThreads, smem, s>>>(a);
  FC_LAUNCH_CHECK("coupling_toy_kernel");
  return FLOWCUDA_OK;
}

// Returns FLOWCUDA_EUNSUPPORTED when the spec is outside this fast path (caller falls back
// to the generic dense path).
int coupling_toy_dispatch(const flowcuda_coupling_spec* sp, const float* params, const float* in, float* out_v,
                          float* out_lp, int64_t n, int layer_begin, int layer_end, int with_base, bool inverse,
                          cudaStream_t s) {
  if (sp->event_dim != 2 || sp->num_hidden != 2 || sp->hidden[0] > kToyH || sp->hidden[1] > kToyH)
    return FLOWCUDA_EUNSUPPORTED;
  ToyArgs a;
  a.params = params; a.in = in; a.out_v = out_v; a.out_lp = out_lp; a.n = n;
  a.L = sp->num_layers; a.h1 = sp->hidden[0]; a.h2 = sp->hidden[1];
  a.layer_begin = layer_begin; a.layer_end = layer_end; a.with_base = with_base; a.base = sp->base;
  a.lo = sp->range_min; a.hi = sp->range_max;
  switch (sp->transform) {
    case FLOWCUDA_TRANSFORM_AFFINE_SCALAR:
      a.out_full = 2;
      return inverse ? toy_launch<FLOWCUDA_TRANSFORM_AFFINE_SCALAR, 0, 4, true>(a, s)
                     : toy_launch<FLOWCUDA_TRANSFORM_AFFINE_SCALAR, 0, 4, false>(a, s);
    case FLOWCUDA_TRANSFORM_AFFINE:
      a.out_full = 4;
      return inverse ? toy_launch<FLOWCUDA_TRANSFORM_AFFINE, 0, 4, true>(a, s)
                     : toy_launch<FLOWCUDA_TRANSFORM_AFFINE, 0, 4, false>(a, s);
    case FLOWCUDA_TRANSFORM_RQS:
      a.out_full = 2 * (3 * sp->num_bins + 1);
      if (sp->num_bins == 12)
        return inverse ? toy_launch<FLOWCUDA_TRANSFORM_RQS, 12, 40, true>(a, s)
                       : toy_launch<FLOWCUDA_TRANSFORM_RQS, 12, 40, false>(a, s);
      if (sp->num_bins == 8)
        return inverse ? toy_launch<FLOWCUDA_TRANSFORM_RQS, 8, 28, true>(a, s)
                       : toy_launch<FLOWCUDA_TRANSFORM_RQS, 8, 28, false>(a, s);
      return FLOWCUDA_EUNSUPPORTED;
  }
  return FLOWCUDA_EUNSUPPORTED;
}

}  // namespace flowcuda
