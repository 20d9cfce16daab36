// coupling_toy_grad.cu — rows #7/#9 for the event_dim = 2 MaskedCoupling flows (rows #2/#3):
// value_and_grad of loss = -mean(Transformed.log_prob(x)) (src/real_nvp_toy_densities.py:139-151,170,
// src/maf_toy_densities.py:140-152,171), hand-written reverse mode, FP32.
//
// CTA = 128 threads = one 128-sample tile, thread = sample.  Forward sweep keeps only the
// per-layer 2-float states (shared memory); the backward sweep walks the layers in reverse order
// of application, recomputes each conditioner, runs the inner bijector's reverse mode
// (rqs_grad.cuh) and back-propagates through the MLP.  Parameter gradients are batch
// reductions (dW = H^T·dOut): activations/adjoints of the tile are staged in shared memory and
// every thread owns a fixed slice of the weight gradient, accumulated into this CTA's private
// partial row in the workspace (no atomics); a second kernel sums the rows in a fixed order in
// double precision — the result is deterministic.
#include "common.cuh"
#include "rqs_grad.cuh"

namespace flowcuda {

constexpr int kTgThreads = 128, kTgH = 64, kTgMaxL = 32;
// Row stride of the staged activation / adjoint tiles is T + 4 (see the kernel): the batch reductions read 8-16
// different rows at the same sample index from the lanes of a warp — a 16-way bank conflict at stride 128.

struct TgArgs {
  const float* params;
  const float* x;     // [N,2]
  float* partial;     // [grid][L*PL + 1]
  int64_t n;
  int L, h1, h2, of;  // conditioner 2 -> h1 -> h2 -> of
  int base;
  float lo, hi;
};

template <int OP>
struct TgImg {  // floats: w1v[64] b1[64] W2[64][64] b2[64] W3[64][OP] b3[OP]
  static constexpr int w1 = 0, b1 = 64, W2 = 128, b2 = 128 + 4096, W3 = b2 + 64, b3 = W3 + 64 * OP, total = b3 + OP;
};

template <int OP>
__device__ void tg_stage(float* sw, const TgArgs& a, int l, int col0, int need) {
  using I = TgImg<OP>;
  const int h1 = a.h1, h2 = a.h2, of = a.of, t = l & 1;
  const size_t PL = (size_t)2 * h1 + h1 + (size_t)h1 * h2 + h2 + (size_t)h2 * of + of;
  const float* p = a.params + PL * l;
  const float* pb1 = p + 2 * h1;
  const float* pW2 = pb1 + h1;
  const float* pb2 = pW2 + (size_t)h1 * h2;
  const float* pW3 = pb2 + h2;
  const float* pb3 = pW3 + (size_t)h2 * of;
  for (int i = threadIdx.x; i < I::total; i += blockDim.x) {
    float v = 0.f;
    if (i < I::b1) { if (i < h1) v = p[(1 - t) * h1 + i]; }
    else if (i < I::W2) { int c = i - I::b1; if (c < h1) v = pb1[c]; }
    else if (i < I::b2) { int r = (i - I::W2) >> 6, c = (i - I::W2) & 63; if (r < h1 && c < h2) v = pW2[r * h2 + c]; }
    else if (i < I::W3) { int c = i - I::b2; if (c < h2) v = pb2[c]; }
    else if (i < I::b3) { int r = (i - I::W3) / OP, c = (i - I::W3) % OP; if (r < h2 && c < need) v = pW3[r * of + col0 + c]; }
    else { int c = i - I::b3; if (c < need) v = pb3[col0 + c]; }
    sw[i] = v;
  }
}

// conditioner forward for this thread's sample: h1 -> (regs), h2 -> (regs), o[NEED]
template <int OP, int NEED>
__device__ __forceinline__ void tg_cond(const float* sw, float xm, float* h2, float* o) {
  using I = TgImg<OP>;
  float h1[kTgH];
#pragma unroll
  for (int j = 0; j < kTgH; ++j) h1[j] = fmaxf(fmaf(xm, sw[I::w1 + j], sw[I::b1 + j]), 0.f);
  // packed accumulators: one FFMA2 (fma.rn.f32x2) per two outputs — a weight row arrives as 128-bit loads of two
  // packed pairs, the activation is splatted once per row (scalar FFMA issues at half rate on sm_100)
  f2_t h2p[kTgH / 2];
#pragma unroll
  for (int k = 0; k < kTgH / 2; ++k) h2p[k] = *reinterpret_cast<const f2_t*>(sw + I::b2 + 2 * k);
#pragma unroll 4
  for (int j = 0; j < kTgH; ++j) {
    const f2_t hj = f2_splat(h1[j]);
    const ulonglong2* wr = reinterpret_cast<const ulonglong2*>(sw + I::W2 + j * 64);
#pragma unroll
    for (int k4 = 0; k4 < 16; ++k4) {
      const ulonglong2 w = wr[k4];
      h2p[2 * k4] = f2_fma(hj, w.x, h2p[2 * k4]);
      h2p[2 * k4 + 1] = f2_fma(hj, w.y, h2p[2 * k4 + 1]);
    }
  }
#pragma unroll
  for (int k = 0; k < kTgH / 2; ++k) {
    float lo, hi;
    f2_unpack(h2p[k], lo, hi);
    h2[2 * k] = fmaxf(lo, 0.f);
    h2[2 * k + 1] = fmaxf(hi, 0.f);
  }
  constexpr int NP = NEED / 2;  // packed output pairs (+ one scalar when NEED is odd)
  f2_t op[NP > 0 ? NP : 1];
#pragma unroll
  for (int i = 0; i < NP; ++i) op[i] = *reinterpret_cast<const f2_t*>(sw + I::b3 + 2 * i);
  float olast = (NEED & 1) ? sw[I::b3 + NEED - 1] : 0.f;
#pragma unroll 4
  for (int k = 0; k < kTgH; ++k) {
    const float hk = h2[k];
    const f2_t hk2 = f2_splat(hk);
    const f2_t* wr = reinterpret_cast<const f2_t*>(sw + I::W3 + k * OP);
#pragma unroll
    for (int i = 0; i < NP; ++i) op[i] = f2_fma(hk2, wr[i], op[i]);
    if (NEED & 1) olast = fmaf(hk, sw[I::W3 + k * OP + NEED - 1], olast);
  }
#pragma unroll
  for (int i = 0; i < NP; ++i) f2_unpack(op[i], o[2 * i], o[2 * i + 1]);
  if (NEED & 1) o[NEED - 1] = olast;
}

// TR/KT as in coupling_toy.cu; OP = padded needed-output count; T = threads = samples per tile (128 or 256:
// the larger tile doubles the warps per SM — shared memory allows one CTA — and halves the weight staging
// per sample; it needs (L+1)*2*T floats of states, so deep stacks fall back to 128).
template <int TR, int KT, int OP, int T>
__global__ void __launch_bounds__(T) coupling_toy_grad_kernel(TgArgs a) {
  using I = TgImg<OP>;
  constexpr int kT = T, kLd = T + 4;  // row stride = 4 (mod 32): the 128-bit reads of the batch reductions are conflict-free
  constexpr int R = T / kTgH;  // threads per weight-gradient row
  constexpr int NEED = TR == FLOWCUDA_TRANSFORM_RQS ? 3 * KT + 1 : 2;
  extern __shared__ __align__(16) float smem[];
  float* sw = smem;                                  // layer image
  float* bufH = sw + ((I::total + 3) & ~3);          // [64][128] h2, later dh1
  float* bufG = bufH + kTgH * kLd;                 // [64][129] dh2
  float* bufO = bufG + kTgH * kLd;                 // [OP][129] d(conditioner out)
  float* sxm = bufO + OP * kLd;                    // [128] visible coordinate
  float* sst = sxm + kT;                     // [L+1][2][128] states
  float* sred = sst + (a.L + 1) * 2 * kT;    // [T/32] loss reduction
  const int tid = threadIdx.x, L = a.L;
  const int h1 = a.h1, h2n = a.h2, of = a.of;
  const size_t PL = (size_t)2 * h1 + h1 + (size_t)h1 * h2n + h2n + (size_t)h2n * of + of;
  float* part = a.partial + (size_t)blockIdx.x * (PL * L + 1);
  float loss_acc = 0.f;
  const int64_t ntiles = (a.n + kT - 1) / kT;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t idx = tile * kT + tid;
    const float live = idx < a.n ? 1.f : 0.f;
    float2 xv = idx < a.n ? reinterpret_cast<const float2*>(a.x)[idx] : make_float2(0.f, 0.f);
    float c0 = xv.x, c1 = xv.y, ldsum = 0.f;
    // ---- forward sweep: layer L-1 first ... layer 0 last
    for (int it = 0; it < L; ++it) {
      const int l = L - 1 - it, t = l & 1;
      const int col0 = TR == FLOWCUDA_TRANSFORM_AFFINE_SCALAR ? 0 : t * NEED;
      __syncthreads();
      tg_stage<OP>(sw, a, l, col0, NEED);
      __syncthreads();
      sst[(it * 2 + 0) * kT + tid] = c0;
      sst[(it * 2 + 1) * kT + tid] = c1;
      float h2[kTgH], o[NEED];
      tg_cond<OP, NEED>(sw, t ? c0 : c1, h2, o);
      float nv, l1;
      if constexpr (TR == FLOWCUDA_TRANSFORM_RQS) rqs_apply<false, KT>([&](int i) { return o[i]; }, KT, a.lo, a.hi, t ? c1 : c0, nv, l1);
      else affine_apply<false>(o[0], o[1], t ? c1 : c0, nv, l1);
      if (t) c1 = nv; else c0 = nv;
      ldsum += l1;
    }
    float a0, a1;  // adjoints of -log_prob w.r.t. the current state
    if (a.base == FLOWCUDA_BASE_NORMAL) {
      loss_acc += live * (0.5f * (c0 * c0 + c1 * c1) + 2.f * kHalfLog2Pi - ldsum);
      a0 = c0 * live; a1 = c1 * live;
    } else {
      loss_acc += live * (-ldsum);
      a0 = a1 = 0.f;
    }
    // ---- backward sweep: the layer applied last (layer 0) first
    for (int it = L - 1; it >= 0; --it) {
      const int l = L - 1 - it, t = l & 1;
      const int col0 = TR == FLOWCUDA_TRANSFORM_AFFINE_SCALAR ? 0 : t * NEED;
      __syncthreads();
      tg_stage<OP>(sw, a, l, col0, NEED);
      const float s0 = sst[(it * 2 + 0) * kT + tid], s1 = sst[(it * 2 + 1) * kT + tid];
      const float xm = t ? s0 : s1, xt = t ? s1 : s0;
      sxm[tid] = xm;
      __syncthreads();
      float xbar;
      {  // S1: conditioner forward, inner bijector reverse, dh2
        float h2[kTgH], o[NEED], obar[NEED], yy, ll;
        tg_cond<OP, NEED>(sw, xm, h2, o);
        if constexpr (TR == FLOWCUDA_TRANSFORM_RQS) rqs_fwd_bwd<KT>(o, a.lo, a.hi, xt, t ? a1 : a0, -live, yy, ll, xbar, obar);
        else affine_fwd_bwd(o, xt, t ? a1 : a0, -live, yy, ll, xbar, obar);
#pragma unroll
        for (int i = 0; i < OP; ++i) bufO[i * kLd + tid] = i < NEED ? obar[i] : 0.f;
#pragma unroll 4
        for (int k = 0; k < kTgH; ++k) {
          float g = 0.f;
#pragma unroll
          for (int i = 0; i < NEED; ++i) g = fmaf(sw[I::W3 + k * OP + i], obar[i], g);
          bufH[k * kLd + tid] = h2[k];
          bufG[k * kLd + tid] = h2[k] > 0.f ? g : 0.f;
        }
      }
      __syncthreads();
      float* pl = part + PL * l;
      float* gW1 = pl; float* gb1 = gW1 + 2 * h1; float* gW2 = gb1 + h1; float* gb2 = gW2 + (size_t)h1 * h2n;
      float* gW3 = gb2 + h2n; float* gb3 = gW3 + (size_t)h2n * of;
      {  // C1: dW3 = H2^T dO, db3, dW2 = H1^T dH2, db2 — thread owns row k = tid/2 and every second column
        const int k = tid / R, half = tid % R;  // `half`: which of the row's R interleaved column sets
        constexpr int HC = OP / R, C2N = kTgH / R;
        float acc3[HC];
#pragma unroll
        for (int i = 0; i < HC; ++i) acc3[i] = 0.f;
        float acc2[C2N];
#pragma unroll
        for (int i = 0; i < C2N; ++i) acc2[i] = 0.f;
        const float w1k = sw[I::w1 + k], b1k = sw[I::b1 + k];
        float accb = 0.f;
        for (int s = 0; s < kT; s += 4) {  // four samples per 128-bit shared-memory read
          const float4 hk = *reinterpret_cast<const float4*>(bufH + k * kLd + s);
#pragma unroll
          for (int i = 0; i < HC; ++i) {
            const float4 o4 = *reinterpret_cast<const float4*>(bufO + (R * i + half) * kLd + s);
            acc3[i] = fmaf(hk.w, o4.w, fmaf(hk.z, o4.z, fmaf(hk.y, o4.y, fmaf(hk.x, o4.x, acc3[i]))));
          }
          const float4 x4 = *reinterpret_cast<const float4*>(sxm + s);
          float4 h1k;  // recomputed first-layer activation, row k
          h1k.x = fmaxf(fmaf(x4.x, w1k, b1k), 0.f);
          h1k.y = fmaxf(fmaf(x4.y, w1k, b1k), 0.f);
          h1k.z = fmaxf(fmaf(x4.z, w1k, b1k), 0.f);
          h1k.w = fmaxf(fmaf(x4.w, w1k, b1k), 0.f);
#pragma unroll
          for (int i = 0; i < C2N; ++i) {
            const float4 g4 = *reinterpret_cast<const float4*>(bufG + (R * i + half) * kLd + s);
            acc2[i] = fmaf(h1k.w, g4.w, fmaf(h1k.z, g4.z, fmaf(h1k.y, g4.y, fmaf(h1k.x, g4.x, acc2[i]))));
          }
          if (half < 2) {
            const float4 b4 = half ? *reinterpret_cast<const float4*>(bufG + k * kLd + s)
                                   : (k < OP ? *reinterpret_cast<const float4*>(bufO + k * kLd + s) : make_float4(0.f, 0.f, 0.f, 0.f));
            accb += (b4.x + b4.y) + (b4.z + b4.w);
          }
        }
        if (k < h2n) {
#pragma unroll
          for (int i = 0; i < HC; ++i) {
            const int c = R * i + half;
            if (c < NEED) gW3[(size_t)k * of + col0 + c] += acc3[i];
          }
        }
        if (k < h1) {
#pragma unroll
          for (int i = 0; i < C2N; ++i) {
            const int c = R * i + half;
            if (c < h2n) gW2[(size_t)k * h2n + c] += acc2[i];
          }
        }
        if (half == 1) { if (k < h2n) gb2[k] += accb; }
        else if (half == 0 && k < NEED) gb3[col0 + k] += accb;
      }
      __syncthreads();
      float dxm = 0.f;
      {  // S2: dh1 = relu'(h1) * (W2 dh2), dxm = w1 . dh1   (bufH is free now: reuse for dh1)
        float g2[kTgH];
#pragma unroll
        for (int k = 0; k < kTgH; ++k) g2[k] = bufG[k * kLd + tid];
#pragma unroll 2
        for (int j = 0; j < kTgH; ++j) {
          // four interleaved partial sums: one 64-long dependent FMA chain per output left the four warps of
          // the CTA waiting on the FMA latency
          float ga = 0.f, gb = 0.f, gc = 0.f, gd = 0.f;
#pragma unroll
          for (int k = 0; k < kTgH; k += 4) {
            const float4 w4 = *reinterpret_cast<const float4*>(sw + I::W2 + j * 64 + k);
            ga = fmaf(w4.x, g2[k], ga);
            gb = fmaf(w4.y, g2[k + 1], gb);
            gc = fmaf(w4.z, g2[k + 2], gc);
            gd = fmaf(w4.w, g2[k + 3], gd);
          }
          float g = (ga + gb) + (gc + gd);
          const float pre = fmaf(xm, sw[I::w1 + j], sw[I::b1 + j]);
          g = pre > 0.f ? g : 0.f;
          bufH[j * kLd + tid] = g;
          dxm = fmaf(sw[I::w1 + j], g, dxm);
        }
      }
      __syncthreads();
      if (tid < 2 * kTgH) {  // C2: dW1 (visible row) and db1
        const int j = tid & 63, isb = tid >> 6;
        float acc = 0.f;
        for (int s = 0; s < kT; ++s) acc = fmaf(isb ? 1.f : sxm[s], bufH[j * kLd + s], acc);
        if (j < h1) {
          if (isb) gb1[j] += acc; else gW1[(1 - t) * h1 + j] += acc;
        }
      }
      if (t) { a1 = xbar; a0 += dxm; } else { a0 = xbar; a1 += dxm; }
    }
  }
  // loss: fixed-order block reduction
  float ls = warp_sum(loss_acc);
  __syncthreads();
  if ((tid & 31) == 0) sred[tid >> 5] = ls;
  __syncthreads();
  if (tid == 0) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < T / 32; ++w) t += sred[w];
    part[PL * L] += t;
  }
}

// out[c] = scale * sum_rows partial[r][c] in double, fixed order; column `cols-1` is the loss.
__global__ void __launch_bounds__(128) tg_reduce_kernel(const float* __restrict__ partial, int rows, int64_t cols, double scale,
                                                       float* __restrict__ loss, float* __restrict__ grad) {
  const int64_t col = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (col >= cols) return;
  double acc = 0.0;
  for (int r = lane; r < rows; r += 32) acc += (double)partial[(size_t)r * cols + col];
  acc = warp_sum_d(acc);
  if (lane == 0) {
    if (col == cols - 1) { if (loss) loss[0] = (float)(acc * scale); }
    else if (grad) grad[col] = (float)(acc * scale);
  }
}

static bool tg_supported(const flowcuda_coupling_spec* sp) {
  return sp && sp->event_dim == 2 && sp->num_hidden == 2 && sp->hidden[0] >= 1 && sp->hidden[0] <= kTgH &&
         sp->hidden[1] >= 1 && sp->hidden[1] <= kTgH && sp->num_layers >= 1 && sp->num_layers <= kTgMaxL &&
         (sp->transform != FLOWCUDA_TRANSFORM_RQS || sp->num_bins == 12 || sp->num_bins == 8);
}
static int tg_of(const flowcuda_coupling_spec* sp) {
  return sp->transform == FLOWCUDA_TRANSFORM_AFFINE_SCALAR ? 2 : 2 * (sp->transform == FLOWCUDA_TRANSFORM_RQS ? 3 * sp->num_bins + 1 : 2);
}
static size_t tg_pl(const flowcuda_coupling_spec* sp) {
  const size_t h1 = sp->hidden[0], h2 = sp->hidden[1], of = tg_of(sp);
  return 2 * h1 + h1 + h1 * h2 + h2 + h2 * of + of;
}
static int tg_grid(int64_t n) {
  int64_t tiles = (n + kTgThreads - 1) / kTgThreads;
  int64_t cap = sm_count();
  return (int)(tiles < 1 ? 1 : (tiles < cap ? tiles : cap));
}

template <int TR, int KT, int OP, int T>
static int tg_launch_t(TgArgs a, size_t fl, cudaStream_t s) {
  auto kern = coupling_toy_grad_kernel<TR, KT, OP, T>;
  FC_ENSURE_SMEM(kern, fl * 4);
  const int64_t tiles = (a.n + T - 1) / T;
  const int grid = (int)(tiles < sm_count() ? tiles : sm_count());  // never more rows than tg_grid(n) zeroed
  kern<<<grid, T, fl * 4, s>>>(a);
  FC_LAUNCH_CHECK("coupling_toy_grad_kernel");
  return FLOWCUDA_OK;
}

template <int TR, int KT, int OP>
static int tg_launch(const flowcuda_coupling_spec* sp, TgArgs a, int grid, cudaStream_t s) {
  using I = TgImg<OP>;
  (void)sp; (void)grid;
  auto floats = [&](int T) {
    return (size_t)((I::total + 3) & ~3) + 2 * (size_t)kTgH * (T + 4) + (size_t)OP * (T + 4) + T + (size_t)(a.L + 1) * 2 * T + 8;
  };
  // 256-sample tiles when every SM still gets a tile and the states fit the 227 KB of shared memory
  if (a.n >= (int64_t)256 * sm_count() && floats(256) * 4 <= 232448) return tg_launch_t<TR, KT, OP, 256>(a, floats(256), s);
  return tg_launch_t<TR, KT, OP, 128>(a, floats(128), s);
}

}  // namespace flowcuda

using namespace flowcuda;

namespace flowcuda {
bool coupling_dense_grad_supported(const flowcuda_coupling_spec* sp);
size_t coupling_dense_grad_workspace_bytes(const flowcuda_coupling_spec* sp, int64_t n);
int coupling_dense_grad_run(const flowcuda_coupling_spec* sp, const float* params, const float* x, int64_t n, float* loss,
                            float* grad, void* workspace, size_t wsb, cudaStream_t s, flowcuda_layer_grad_fn on_layer, void* user);
}

// The stored-activation engine serves every shape the fused toy kernel does not, and the toy shapes too when
// the caller asks for the tensor-core GEMMs (precision BF16X3).
static bool tg_use_dense(const flowcuda_coupling_spec* sp) {
  return !tg_supported(sp) || (sp->precision == FLOWCUDA_PRECISION_BF16X3 && coupling_dense_grad_supported(sp));
}

extern "C" {

size_t flowcuda_coupling_nll_grad_workspace_bytes(const flowcuda_coupling_spec* spec, int64_t n) {
  if (spec && n >= 1 && tg_use_dense(spec)) return coupling_dense_grad_workspace_bytes(spec, n);
  if (!tg_supported(spec) || n < 1) return 0;
  return (size_t)tg_grid(n) * (tg_pl(spec) * spec->num_layers + 1) * sizeof(float) + 256;
}

int flowcuda_coupling_nll_grad(const flowcuda_coupling_spec* spec, const float* params, const float* x, int64_t n,
                               float* loss, float* grad, void* workspace, size_t workspace_bytes, void* stream) {
  return flowcuda_coupling_nll_grad_layers(spec, params, x, n, loss, grad, workspace, workspace_bytes, stream, nullptr, nullptr);
}

int flowcuda_coupling_nll_grad_layers(const flowcuda_coupling_spec* spec, const float* params, const float* x, int64_t n,
                                      float* loss, float* grad, void* workspace, size_t workspace_bytes, void* stream,
                                      flowcuda_layer_grad_fn on_layer, void* user) {
  FC_REQUIRE(spec && params && x && workspace, "coupling_nll_grad: NULL pointer");
  if (tg_use_dense(spec)) {  // generic event size: stored-activation reverse mode (coupling_dense_grad.cu)
    FC_REQUIRE(n >= 1 && loss && grad, "coupling_nll_grad: bad arguments");
    return coupling_dense_grad_run(spec, params, x, n, loss, grad, workspace, workspace_bytes, (cudaStream_t)stream, on_layer, user);
  }
  FC_REQUIRE(n >= 1, "coupling_nll_grad: n must be >= 1");
  FC_REQUIRE(((uintptr_t)x & 7) == 0, "coupling_nll_grad: x must be 8-byte aligned");
  const size_t need = flowcuda_coupling_nll_grad_workspace_bytes(spec, n);
  if (workspace_bytes < need) {
    set_error("coupling_nll_grad: workspace %zu < %zu bytes", workspace_bytes, need);
    return FLOWCUDA_EWORKSPACE;
  }
  cudaStream_t s = (cudaStream_t)stream;
  float* part = (float*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  const int grid = tg_grid(n);
  const int64_t cols = (int64_t)tg_pl(spec) * spec->num_layers + 1;
  FC_CUDA(cudaMemsetAsync(part, 0, (size_t)grid * cols * sizeof(float), s));
  TgArgs a;
  a.params = params; a.x = x; a.partial = part; a.n = n; a.L = spec->num_layers;
  a.h1 = spec->hidden[0]; a.h2 = spec->hidden[1]; a.of = tg_of(spec); a.base = spec->base;
  a.lo = spec->range_min; a.hi = spec->range_max;
  int rc = FLOWCUDA_EUNSUPPORTED;
  switch (spec->transform) {
    case FLOWCUDA_TRANSFORM_AFFINE_SCALAR: rc = tg_launch<FLOWCUDA_TRANSFORM_AFFINE_SCALAR, 0, 4>(spec, a, grid, s); break;
    case FLOWCUDA_TRANSFORM_AFFINE: rc = tg_launch<FLOWCUDA_TRANSFORM_AFFINE, 0, 4>(spec, a, grid, s); break;
    case FLOWCUDA_TRANSFORM_RQS:
      rc = spec->num_bins == 12 ? tg_launch<FLOWCUDA_TRANSFORM_RQS, 12, 40>(spec, a, grid, s)
                                : tg_launch<FLOWCUDA_TRANSFORM_RQS, 8, 28>(spec, a, grid, s);
      break;
  }
  if (rc) return rc;
  tg_reduce_kernel<<<(int)((cols + 3) / 4), 128, 0, s>>>(part, grid, cols, 1.0 / (double)n, loss, grad);
  FC_LAUNCH_CHECK("tg_reduce_kernel");
  if (on_layer)  // one fused kernel pair: every layer's slice is complete at the same point of the caller's stream
    for (int l = 0; l < spec->num_layers; ++l) on_layer(user, l, (int64_t)tg_pl(spec) * l, (int64_t)tg_pl(spec), stream);
  return FLOWCUDA_OK;
}

}  // extern "C"
