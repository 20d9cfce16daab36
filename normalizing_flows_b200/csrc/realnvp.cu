// realnvp.cu — row #1: the hand-rolled Real NVP chain of src/real_nvp.py, D = 2.
//
// One kernel runs the whole L-layer chain per sample and fuses the N(0,I) log-density:
// HBM traffic is y in (8 B) and log_prob out (4 B) per sample.  The 18-float conditioner
// [1 -> H -> 2] of every layer is staged once per CTA in shared memory, pre-splatted as
// {w,w} pairs so that the Blackwell packed-fp32 pipe (fma.rn.f32x2 -> FFMA2) processes two
// samples per issue slot.
#include "common.cuh"

namespace flowcuda {

constexpr int kNvpThreads = 256;
constexpr int kNvpMaxLayers = 64;
constexpr int kNvpMaxHidden = 32;

// smem image of one layer (in float2 units, each weight duplicated):
//   [0,H) W1   [H,2H) b1   [2H,3H) W2[:,0] (shift)   [3H,4H) W2[:,1] (log_scale)   4H: b2s  4H+1: b2l
__host__ __device__ constexpr int nvp_layer_f2(int H) { return 4 * H + 2; }

template <int H>
__device__ __forceinline__ void stage_params(float2* sp, const float* __restrict__ params, int L, int Hrt) {
  const int Hh = H > 0 ? H : Hrt;
  const int P = 4 * Hh + 2;
  for (int i = threadIdx.x; i < L * P; i += blockDim.x) {
    int l = i / P, r = i - l * P;
    const float* p = params + (size_t)l * P;
    float v;
    if (r < Hh) v = p[r];                          // W1[0, j]
    else if (r < 2 * Hh) v = p[r];                 // b1[j]
    else if (r < 3 * Hh) v = p[2 * Hh + 2 * (r - 2 * Hh)];      // W2[j, 0]
    else if (r < 4 * Hh) v = p[2 * Hh + 2 * (r - 3 * Hh) + 1];  // W2[j, 1]
    else v = p[4 * Hh + (r - 4 * Hh)];             // b2
    sp[i] = make_float2(v, v);
  }
}

// Conditioner on the pass-through value of two samples at once (src/real_nvp.py:161-171).
template <int H>
__device__ __forceinline__ void cond_pair(const float2* __restrict__ w, int Hrt, float2 y1, float2& shift,
                                          float2& ls) {
  const int Hh = H > 0 ? H : Hrt;
  shift = w[4 * Hh];
  ls = w[4 * Hh + 1];
#pragma unroll
  for (int j = 0; j < Hh; ++j) {
    float2 pre = ffma2(y1, w[j], w[Hh + j]);
    float2 sm = fmul2(pre, splat2(0.01f));
    float2 h = make_float2(fmaxf(pre.x, sm.x), fmaxf(pre.y, sm.y));  // jax.nn.leaky_relu
    shift = ffma2(h, w[2 * Hh + j], shift);
    ls = ffma2(h, w[3 * Hh + j], ls);
  }
}

// MODE 0: data -> base (inverse_nvp chain, layers L-1..0) ; MODE 1: base -> data (forward_nvp, 0..L-1)
template <int H, int MODE>
__global__ void __launch_bounds__(kNvpThreads)
realnvp_chain_kernel(const float* __restrict__ params, int L, int Hrt, uint64_t flip_mask,
                     const float* __restrict__ in, float* __restrict__ out_v, float* __restrict__ out_lp,
                     float* __restrict__ out_ls, int64_t n) {
  extern __shared__ float2 sp[];
  stage_params<H>(sp, params, L, Hrt);
  __syncthreads();
  const int Hh = H > 0 ? H : Hrt;
  const int P2 = nvp_layer_f2(Hh);

  // one "pair" = samples (2q, 2q+1) = one float4 of the [N,2] input
  const int64_t npairs = n >> 1;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const float4* in4 = reinterpret_cast<const float4*>(in);
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < npairs; q += stride) {
    float4 v = ld_stream_f4(in4 + q);
    float2 c0 = make_float2(v.x, v.z);  // component 0 of samples A,B
    float2 c1 = make_float2(v.y, v.w);  // component 1
    float2 lsum = make_float2(0.f, 0.f);
#pragma unroll 1
    for (int i = 0; i < L; ++i) {
      const int l = MODE == 0 ? L - 1 - i : i;
      const float2* w = sp + l * P2;
      const bool flip = (flip_mask >> l) & 1ull;
      float2 pass = flip ? c1 : c0;
      float2 tr = flip ? c0 : c1;
      float2 shift, ls;
      cond_pair<H>(w, Hrt, pass, shift, ls);
      if (MODE == 0) {
        // x2 = (y2 - shift) * exp(-log_scale)   (real_nvp.py:129)
        float2 d = fadd2(tr, make_float2(-shift.x, -shift.y));
        float2 e = make_float2(exp2f(-kLog2e * ls.x), exp2f(-kLog2e * ls.y));
        tr = fmul2(d, e);
      } else {
        // y2 = x2 * exp(log_scale) + shift      (real_nvp.py:96)
        float2 e = make_float2(exp2f(kLog2e * ls.x), exp2f(kLog2e * ls.y));
        tr = ffma2(tr, e, shift);
      }
      lsum = fadd2(lsum, ls);
      if (flip) c0 = tr; else c1 = tr;
    }
    if (out_v) st_stream_f4(reinterpret_cast<float4*>(out_v) + q, make_float4(c0.x, c1.x, c0.y, c1.y));
    if (out_lp) {
      // log N(z;0,I) - sum(log_scale)   (real_nvp.py:57-68,157)
      float2 r2 = ffma2(c0, c0, fmul2(c1, c1));
      float2 lp = ffma2(r2, splat2(-0.5f), make_float2(-2.f * kHalfLog2Pi - lsum.x, -2.f * kHalfLog2Pi - lsum.y));
      st_stream_f2(reinterpret_cast<float2*>(out_lp) + q, lp);
    }
    if (out_ls) st_stream_f2(reinterpret_cast<float2*>(out_ls) + q, lsum);
  }
  // odd tail sample
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
    int64_t s = n - 1;
    float2 c0 = splat2(in[2 * s]), c1 = splat2(in[2 * s + 1]), lsum = splat2(0.f);
    for (int i = 0; i < L; ++i) {
      const int l = MODE == 0 ? L - 1 - i : i;
      const bool flip = (flip_mask >> l) & 1ull;
      float2 pass = flip ? c1 : c0, tr = flip ? c0 : c1, shift, ls;
      cond_pair<H>(sp + l * P2, Hrt, pass, shift, ls);
      if (MODE == 0) tr = splat2((tr.x - shift.x) * exp2f(-kLog2e * ls.x));
      else tr = splat2(fmaf(tr.x, exp2f(kLog2e * ls.x), shift.x));
      lsum.x += ls.x;
      if (flip) c0 = tr; else c1 = tr;
    }
    if (out_v) { out_v[2 * s] = c0.x; out_v[2 * s + 1] = c1.x; }
    if (out_lp) out_lp[s] = -0.5f * (c0.x * c0.x + c1.x * c1.x) - 2.f * kHalfLog2Pi - lsum.x;
    if (out_ls) out_ls[s] = lsum.x;
  }
}

template <int MODE>
static int launch_chain(const float* params, int L, int H, uint64_t flip_mask, const float* in, float* out_v,
                        float* out_lp, float* out_ls, int64_t n, cudaStream_t s) {
  FC_REQUIRE(n >= 0, "realnvp: negative n");
  if (n == 0) return FLOWCUDA_OK;
  FC_REQUIRE(params && in, "realnvp: NULL params or input");
  FC_REQUIRE(L >= 1 && L <= kNvpMaxLayers, "realnvp: n_layers=%d outside [1,%d]", L, kNvpMaxLayers);
  FC_REQUIRE(H >= 1 && H <= kNvpMaxHidden, "realnvp: hidden=%d outside [1,%d]", H, kNvpMaxHidden);
  FC_REQUIRE(n >= 0, "realnvp: negative n");
  FC_REQUIRE(((uintptr_t)in & 15) == 0 && ((uintptr_t)out_v & 15) == 0 && ((uintptr_t)out_lp & 7) == 0 &&
                 ((uintptr_t)out_ls & 7) == 0,
             "realnvp: buffers must be 16-byte ([N,2]) / 8-byte ([N]) aligned");
  if (n == 0) return FLOWCUDA_OK;
  size_t smem = (size_t)L * nvp_layer_f2(H) * sizeof(float2);
  // grid: a multiple of the SM count; 8 CTAs of 256 threads are resident per SM
  int64_t need = ((n >> 1) + kNvpThreads - 1) / kNvpThreads;
  int64_t cap = (int64_t)sm_count() * 8;
  int grid = (int)(need < 1 ? 1 : (need < cap ? need : cap));
  if (H == 4) {
    realnvp_chain_kernel<4, MODE><<<grid, kNvpThreads, smem, s>>>(params, L, H, flip_mask, in, out_v, out_lp, out_ls, n);
  } else {
    if (smem > 48 * 1024)
      FC_CUDA(cudaFuncSetAttribute(realnvp_chain_kernel<0, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    realnvp_chain_kernel<0, MODE><<<grid, kNvpThreads, smem, s>>>(params, L, H, flip_mask, in, out_v, out_lp, out_ls, n);
  }
  FC_LAUNCH_CHECK("realnvp_chain_kernel");
  return FLOWCUDA_OK;
}

}  // namespace flowcuda

using namespace flowcuda;

extern "C" {

int flowcuda_realnvp_log_prob(const float* params, int n_layers, int hidden, uint64_t flip_mask, const float* y,
                              float* log_prob, float* z, int64_t n, void* stream) {
  if (n == 0) return FLOWCUDA_OK;
  FC_REQUIRE(log_prob || z, "realnvp_log_prob: no output requested");
  return launch_chain<0>(params, n_layers, hidden, flip_mask, y, z, log_prob, nullptr, n, (cudaStream_t)stream);
}

int flowcuda_realnvp_sample(const float* params, int n_layers, int hidden, uint64_t flip_mask, const float* z,
                            float* x, int64_t n, void* stream) {
  if (n == 0) return FLOWCUDA_OK;
  FC_REQUIRE(x, "realnvp_sample: NULL output");
  return launch_chain<1>(params, n_layers, hidden, flip_mask, z, x, nullptr, nullptr, n, (cudaStream_t)stream);
}

int flowcuda_realnvp_forward(const float* layer_params, int hidden, int flip, const float* x, float* y, int64_t n,
                             void* stream) {
  if (n == 0) return FLOWCUDA_OK;
  FC_REQUIRE(y, "realnvp_forward: NULL output");
  return launch_chain<1>(layer_params, 1, hidden, flip ? 1ull : 0ull, x, y, nullptr, nullptr, n, (cudaStream_t)stream);
}

int flowcuda_realnvp_inverse(const float* layer_params, int hidden, int flip, const float* y, float* x,
                             float* log_scale, int64_t n, void* stream) {
  if (n == 0) return FLOWCUDA_OK;
  FC_REQUIRE(x || log_scale, "realnvp_inverse: no output requested");
  return launch_chain<0>(layer_params, 1, hidden, flip ? 1ull : 0ull, y, x, nullptr, log_scale, n, (cudaStream_t)stream);
}

}  // extern "C"
