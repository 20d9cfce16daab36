// realnvp_grad.cu — rows #1/#7/#9: value_and_grad of -mean(log_prob_nvp_chains) for the
// hand-rolled Real NVP chain (src/real_nvp.py:292-298), hand-written reverse mode.
//
// No activations are stored: after the data->base pass the backward sweep walks the layers
// in reverse order of application and rebuilds each layer's input by applying that layer's
// forward_nvp (couplings are exactly invertible).  Parameter gradients are reduced
// deterministically: thread registers -> warp shuffle tree -> per-warp shared rows ->
// per-CTA partial rows in the workspace -> fixed-order double-precision column sums (stage 2,
// which also applies the optional fused SGD step).
#include "common.cuh"

namespace flowcuda {

constexpr int kGradThreads = 128;
constexpr int kGradS = 4;       // samples per thread per sweep
constexpr int kGradMaxH = 32;
constexpr int kGradMaxBlocks = 148 * 8;

template <int HT>
__global__ void __launch_bounds__(kGradThreads)
realnvp_grad_stage1(const float* __restrict__ params, int L, int Hrt, uint64_t flip_mask,
                    const float* __restrict__ y, int64_t n, float* __restrict__ partial /* [grid][L*P+1] */) {
  extern __shared__ float smem[];
  constexpr int HA = HT > 0 ? HT : kGradMaxH;  // array extent
  const int H = HT > 0 ? HT : Hrt;
  const int P = 4 * H + 2;
  const int LP = L * P;
  constexpr int warps = kGradThreads / 32;
  float* sw = smem;        // [LP] parameters, reference order W1[H], b1[H], W2[H][2], b2[2]
  float* acc = smem + LP;  // [warps][LP+1] per-warp accumulators (last column = loss)
  for (int i = threadIdx.x; i < LP; i += blockDim.x) sw[i] = params[i];
  for (int i = threadIdx.x; i < warps * (LP + 1); i += blockDim.x) acc[i] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* myacc = acc + warp * (LP + 1);

  const int64_t span = (int64_t)gridDim.x * kGradThreads * kGradS;
  for (int64_t base = (int64_t)blockIdx.x * kGradThreads * kGradS; base < n; base += span) {
    float v0[kGradS], v1[kGradS], a0[kGradS], a1[kGradS], lv[kGradS];
    float loss_t = 0.f;
#pragma unroll
    for (int s = 0; s < kGradS; ++s) {
      int64_t idx = base + (int64_t)s * kGradThreads + threadIdx.x;  // coalesced per s
      bool live = idx < n;
      lv[s] = live ? 1.f : 0.f;
      float2 yy = live ? ld_stream_f2(reinterpret_cast<const float2*>(y) + idx) : make_float2(0.f, 0.f);
      v0[s] = yy.x;
      v1[s] = yy.y;
    }
    // ---- data -> base: layers L-1 .. 0 (real_nvp.py:268-283) ----
#pragma unroll
    for (int s = 0; s < kGradS; ++s) {
      float lsum = 0.f;
      for (int l = L - 1; l >= 0; --l) {
        const float* w = sw + l * P;
        bool flip = (flip_mask >> l) & 1ull;
        float pass = flip ? v1[s] : v0[s], tr = flip ? v0[s] : v1[s];
        float sh = w[4 * H], ls = w[4 * H + 1];
#pragma unroll
        for (int j = 0; j < HA; ++j) {
          if (j < H) {
            float h = leaky_relu_f(fmaf(pass, w[j], w[H + j]));
            sh = fmaf(h, w[2 * H + 2 * j], sh);
            ls = fmaf(h, w[2 * H + 2 * j + 1], ls);
          }
        }
        tr = (tr - sh) * expf(-ls);
        lsum += ls;
        if (flip) v0[s] = tr; else v1[s] = tr;
      }
      // -log_prob = 0.5|z|^2 + log(2 pi) + sum(log_scale)
      loss_t += lv[s] * (0.5f * (v0[s] * v0[s] + v1[s] * v1[s]) + 2.f * kHalfLog2Pi + lsum);
      a0[s] = v0[s] * lv[s];  // d(-log_prob)/dz = z   (dead lanes contribute exact zeros)
      a1[s] = v1[s] * lv[s];
    }
    // ---- backward: layer 0 (applied last) first ----
    for (int l = 0; l < L; ++l) {
      const float* w = sw + l * P;
      bool flip = (flip_mask >> l) & 1ull;
      float gW1[HA], gb1[HA], gW2s[HA], gW2l[HA], gbs = 0.f, gbl = 0.f;
#pragma unroll
      for (int j = 0; j < HA; ++j) gW1[j] = gb1[j] = gW2s[j] = gW2l[j] = 0.f;
#pragma unroll
      for (int s = 0; s < kGradS; ++s) {
        float x1 = flip ? v1[s] : v0[s], x2 = flip ? v0[s] : v1[s];
        float ap = flip ? a1[s] : a0[s], at = flip ? a0[s] : a1[s];
        // conditioner on the pass-through value (identical for the layer's input and output)
        float sh = w[4 * H], ls = w[4 * H + 1];
        float pre[HA], hh[HA];
#pragma unroll
        for (int j = 0; j < HA; ++j) {
          if (j < H) {
            pre[j] = fmaf(x1, w[j], w[H + j]);
            hh[j] = leaky_relu_f(pre[j]);
            sh = fmaf(hh[j], w[2 * H + 2 * j], sh);
            ls = fmaf(hh[j], w[2 * H + 2 * j + 1], ls);
          }
        }
        float e = expf(-ls);
        float y2 = fmaf(x2, expf(ls), sh);  // forward_nvp (real_nvp.py:96): this layer's input
        float gs = -at * e;                 // adjoint of shift
        float gl = fmaf(-at, x2, lv[s]);    // adjoint of log_scale (+1 from the -log-det term)
        float gy2 = at * e;
        float gy1 = ap;
        gbs += gs;
        gbl += gl;
#pragma unroll
        for (int j = 0; j < HA; ++j) {
          if (j < H) {
            float gh = fmaf(w[2 * H + 2 * j], gs, w[2 * H + 2 * j + 1] * gl);
            float gp = pre[j] >= 0.f ? gh : 0.01f * gh;
            gW2s[j] = fmaf(hh[j], gs, gW2s[j]);
            gW2l[j] = fmaf(hh[j], gl, gW2l[j]);
            gW1[j] = fmaf(gp, x1, gW1[j]);
            gb1[j] += gp;
            gy1 = fmaf(gp, w[j], gy1);
          }
        }
        if (flip) { v1[s] = x1; v0[s] = y2; a1[s] = gy1; a0[s] = gy2; }
        else      { v0[s] = x1; v1[s] = y2; a0[s] = gy1; a1[s] = gy2; }
      }
      float* row = myacc + l * P;
#pragma unroll
      for (int j = 0; j < HA; ++j) {
        if (j < H) {
          float r0 = warp_sum(gW1[j]), r1 = warp_sum(gb1[j]), r2 = warp_sum(gW2s[j]), r3 = warp_sum(gW2l[j]);
          if (lane == 0) {
            row[j] += r0;
            row[H + j] += r1;
            row[2 * H + 2 * j] += r2;
            row[2 * H + 2 * j + 1] += r3;
          }
        }
      }
      float r4 = warp_sum(gbs), r5 = warp_sum(gbl);
      if (lane == 0) {
        row[4 * H] += r4;
        row[4 * H + 1] += r5;
      }
    }
    float lt = warp_sum(loss_t);
    if (lane == 0) myacc[LP] += lt;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < LP + 1; i += blockDim.x) {
    float t = 0.f;
#pragma unroll
    for (int w2 = 0; w2 < warps; ++w2) t += acc[w2 * (LP + 1) + i];
    partial[(size_t)blockIdx.x * (LP + 1) + i] = t;
  }
}

// Stage 2: one warp per column; fixed-order double sums over the CTA partial rows.
__global__ void __launch_bounds__(128)
realnvp_grad_stage2(const float* __restrict__ partial, int nblocks, int LP, double inv_n, float* __restrict__ loss,
                    float* __restrict__ grad, float lr, const float* __restrict__ params, float* __restrict__ params_out) {
  int col = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
  int lane = threadIdx.x & 31;
  if (col > LP) return;
  double a = 0.0;
  for (int b = lane; b < nblocks; b += 32) a += (double)partial[(size_t)b * (LP + 1) + col];
  a = warp_sum_d(a);
  if (lane == 0) {
    float g = (float)(a * inv_n);
    if (col == LP) {
      if (loss) loss[0] = g;
    } else {
      if (grad) grad[col] = g;
      if (params_out) params_out[col] = params[col] - lr * g;  // real_nvp.py:298
    }
  }
}

static int grad_grid(int64_t n) {
  int64_t need = (n + (int64_t)kGradThreads * kGradS - 1) / ((int64_t)kGradThreads * kGradS);
  if (need < 1) need = 1;
  return (int)(need < kGradMaxBlocks ? need : kGradMaxBlocks);
}

}  // namespace flowcuda

using namespace flowcuda;

extern "C" {

size_t flowcuda_realnvp_nll_grad_workspace_bytes(int n_layers, int hidden, int64_t n) {
  if (n_layers < 1 || hidden < 1 || n < 0) return 0;
  return (size_t)grad_grid(n) * ((size_t)n_layers * (4 * hidden + 2) + 1) * sizeof(float);
}

int flowcuda_realnvp_nll_grad(const float* params, int n_layers, int hidden, uint64_t flip_mask, const float* y,
                              int64_t n, float* loss, float* grad, float lr, float* params_out, void* workspace,
                              size_t workspace_bytes, void* stream) {
  FC_REQUIRE(params && y && workspace, "realnvp_nll_grad: NULL pointer");
  FC_REQUIRE(n_layers >= 1 && n_layers <= 64, "realnvp_nll_grad: n_layers=%d outside [1,64]", n_layers);
  FC_REQUIRE(hidden >= 1 && hidden <= kGradMaxH, "realnvp_nll_grad: hidden=%d outside [1,%d]", hidden, kGradMaxH);
  FC_REQUIRE(n >= 1, "realnvp_nll_grad: n must be >= 1");
  FC_REQUIRE(((uintptr_t)y & 7) == 0, "realnvp_nll_grad: y must be 8-byte aligned");
  if (workspace_bytes < flowcuda_realnvp_nll_grad_workspace_bytes(n_layers, hidden, n)) {
    set_error("realnvp_nll_grad: workspace %zu < %zu bytes", workspace_bytes,
              flowcuda_realnvp_nll_grad_workspace_bytes(n_layers, hidden, n));
    return FLOWCUDA_EWORKSPACE;
  }
  cudaStream_t s = (cudaStream_t)stream;
  const int P = 4 * hidden + 2, LP = n_layers * P;
  const int grid = grad_grid(n);
  size_t smem = ((size_t)LP + (kGradThreads / 32) * ((size_t)LP + 1)) * sizeof(float);
  if (hidden == 4) {
    realnvp_grad_stage1<4><<<grid, kGradThreads, smem, s>>>(params, n_layers, hidden, flip_mask, y, n, (float*)workspace);
  } else {
    if (smem > 48 * 1024)
      FC_ENSURE_SMEM(realnvp_grad_stage1<0>, smem);
    realnvp_grad_stage1<0><<<grid, kGradThreads, smem, s>>>(params, n_layers, hidden, flip_mask, y, n, (float*)workspace);
  }
  FC_LAUNCH_CHECK("realnvp_grad_stage1");
  int cols = LP + 1;
  realnvp_grad_stage2<<<(cols + 3) / 4, 128, 0, s>>>((const float*)workspace, grid, LP, 1.0 / (double)n, loss, grad,
                                                    lr, params, lr != 0.f ? params_out : nullptr);
  FC_LAUNCH_CHECK("realnvp_grad_stage2");
  return FLOWCUDA_OK;
}

}  // extern "C"
