// realnvp.cu — row #1: the hand-rolled Real NVP chain of src/real_nvp.py, D = 2.
//
// One kernel runs the whole L-layer chain per sample and fuses the N(0,I) log-density:
// HBM traffic is y in (8 B) and log_prob out (4 B) per sample.  The 18-float conditioner
// [1 -> H -> 2] of every layer is staged once per CTA in shared memory; a thread carries eight
// samples as four packed pairs, so a weight is loaded and splatted once per eight samples and the
// Blackwell packed-fp32 pipe (fma.rn.f32x2 -> FFMA2) processes two samples per issue slot.
#include "common.cuh"

namespace flowcuda {

constexpr int kNvpThreads = 256;
constexpr int kNvpMaxLayers = 64;
constexpr int kNvpMaxHidden = 32;

// smem image of one layer, 4H+4 plain floats (16-byte aligned rows):
//   [0,H) W1   [H,2H) b1   [2H,3H) 0.99·W2[:,0]   [3H,4H) 0.99·W2[:,1]   then As, Bs, Al, Bl
// with leaky_relu(p) = 0.01·p + 0.99·relu(p) (exact) the conditioner output is
//   shift(y) = Bs + As·y + Σ_j 0.99·W2[j,0]·relu(W1_j·y + b1_j),
//   As = 0.01·Σ_j W2[j,0]·W1_j,  Bs = b2_0 + 0.01·Σ_j W2[j,0]·b1_j        (likewise log_scale),
// i.e. one FMNMX instead of FMUL+FMNMX per hidden unit and sample.
__host__ __device__ constexpr int nvp_layer_floats(int H) { return (4 * H + 4 + 3) & ~3; }

template <int H>
__device__ __forceinline__ void stage_params(float* sp, const float* __restrict__ params, int L, int Hrt) {
  const int Hh = H > 0 ? H : Hrt;
  const int P = 4 * Hh + 2, PS = nvp_layer_floats(Hh);
  for (int i = threadIdx.x; i < L * PS; i += blockDim.x) {
    const int l = i / PS, r = i - l * PS;
    const float* p = params + (size_t)l * P;
    float v = 0.f;
    if (r < 2 * Hh) v = p[r];                                             // W1[0, j], b1[j]
    else if (r < 3 * Hh) v = 0.99f * p[2 * Hh + 2 * (r - 2 * Hh)];        // W2[j, 0]
    else if (r < 4 * Hh) v = 0.99f * p[2 * Hh + 2 * (r - 3 * Hh) + 1];    // W2[j, 1]
    else if (r < 4 * Hh + 4) {
      const int which = (r - 4 * Hh) >> 1, isB = (r - 4 * Hh) & 1;        // As, Bs, Al, Bl
      float acc = isB ? p[4 * Hh + which] : 0.f;
      for (int j = 0; j < Hh; ++j) acc = fmaf(0.01f * p[2 * Hh + 2 * j + which], isB ? p[Hh + j] : p[j], acc);
      v = acc;
    }
    sp[i] = v;
  }
}

// One coupling layer on NP packed pairs (2·NP samples): `pass` conditions, `tr` is transformed.
template <int H, int MODE, int NP>
__device__ __forceinline__ void nvp_layer(const float* __restrict__ w, int Hrt, const f2_t (&pass)[NP], f2_t (&tr)[NP],
                                          f2_t (&lsum)[NP]) {
  const int Hh = H > 0 ? H : Hrt;
  f2_t shift[NP], ls[NP];
  {
    const float4 k = *reinterpret_cast<const float4*>(w + 4 * Hh);  // As, Bs, Al, Bl
    const f2_t as = f2_splat(k.x), bs = f2_splat(k.y), al = f2_splat(k.z), bl = f2_splat(k.w);
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      shift[p] = f2_fma(pass[p], as, bs);
      ls[p] = f2_fma(pass[p], al, bl);
    }
  }
#pragma unroll
  for (int j = 0; j < Hh; ++j) {
    const f2_t w1 = f2_splat(w[j]), b1 = f2_splat(w[Hh + j]), ws = f2_splat(w[2 * Hh + j]), wl = f2_splat(w[3 * Hh + j]);
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      float a, b;
      f2_unpack(f2_fma(pass[p], w1, b1), a, b);
      const f2_t h = f2_pack(fmaxf(a, 0.f), fmaxf(b, 0.f));
      shift[p] = f2_fma(h, ws, shift[p]);
      ls[p] = f2_fma(h, wl, ls[p]);
    }
  }
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    float la, lb;
    f2_unpack(ls[p], la, lb);
    if (MODE == 0) {
      // x2 = (y2 - shift) * exp(-log_scale)   (real_nvp.py:129)
      const f2_t e = f2_pack(ex2_ftz(-kLog2e * la), ex2_ftz(-kLog2e * lb));
      tr[p] = f2_mul(f2_fma(shift[p], f2_splat(-1.f), tr[p]), e);
    } else {
      // y2 = x2 * exp(log_scale) + shift      (real_nvp.py:96)
      const f2_t e = f2_pack(ex2_ftz(kLog2e * la), ex2_ftz(kLog2e * lb));
      tr[p] = f2_fma(tr[p], e, shift[p]);
    }
    lsum[p] = f2_add(lsum[p], ls[p]);
  }
}

constexpr int kNvpPairs = 4;  // packed pairs per thread and iteration (8 samples): amortises the weight loads

// MODE 0: data -> base (inverse_nvp chain, layers L-1..0) ; MODE 1: base -> data (forward_nvp, 0..L-1)
template <int H, int MODE>
__global__ void __launch_bounds__(kNvpThreads, 4)
realnvp_chain_kernel(const float* __restrict__ params, int L, int Hrt, uint64_t flip_mask,
                     const float* __restrict__ in, float* __restrict__ out_v, float* __restrict__ out_lp,
                     float* __restrict__ out_ls, int64_t n) {
  extern __shared__ __align__(16) float sp[];
  stage_params<H>(sp, params, L, Hrt);
  __syncthreads();
  const int Hh = H > 0 ? H : Hrt;
  const int PS = nvp_layer_floats(Hh);
  constexpr int NP = kNvpPairs;

  // one "pair" = samples (2q, 2q+1) = one float4 of the [N,2] input; a thread takes NP pairs that are
  // blockDim apart so that every 128-bit access of a warp is contiguous
  const int64_t npairs = n >> 1;
  const float4* in4 = reinterpret_cast<const float4*>(in);
  const int64_t tile = (int64_t)blockDim.x * NP;
  for (int64_t base = (int64_t)blockIdx.x * tile; base < npairs; base += (int64_t)gridDim.x * tile) {
    f2_t c0[NP], c1[NP], lsum[NP];
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      const int64_t q = base + (int64_t)p * blockDim.x + threadIdx.x;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (q < npairs) v = ld_stream_f4(in4 + q);
      c0[p] = f2_pack(v.x, v.z);  // component 0 of samples A,B
      c1[p] = f2_pack(v.y, v.w);  // component 1
      lsum[p] = f2_splat(0.f);
    }
#pragma unroll 1
    for (int i = 0; i < L; ++i) {
      const int l = MODE == 0 ? L - 1 - i : i;
      const float* w = sp + l * PS;
      if ((flip_mask >> l) & 1ull) nvp_layer<H, MODE, NP>(w, Hrt, c1, c0, lsum);  // flip: component 1 conditions
      else nvp_layer<H, MODE, NP>(w, Hrt, c0, c1, lsum);
    }
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      const int64_t q = base + (int64_t)p * blockDim.x + threadIdx.x;
      if (q >= npairs) continue;
      float a0, b0, a1, b1, la, lb;
      f2_unpack(c0[p], a0, b0);
      f2_unpack(c1[p], a1, b1);
      f2_unpack(lsum[p], la, lb);
      if (out_v) st_stream_f4(reinterpret_cast<float4*>(out_v) + q, make_float4(a0, a1, b0, b1));
      if (out_lp) {
        // log N(z;0,I) - sum(log_scale)   (real_nvp.py:57-68,157)
        const float pa = fmaf(-0.5f, fmaf(a0, a0, a1 * a1), -2.f * kHalfLog2Pi - la);
        const float pb = fmaf(-0.5f, fmaf(b0, b0, b1 * b1), -2.f * kHalfLog2Pi - lb);
        st_stream_f2(reinterpret_cast<float2*>(out_lp) + q, make_float2(pa, pb));
      }
      if (out_ls) st_stream_f2(reinterpret_cast<float2*>(out_ls) + q, make_float2(la, lb));
    }
  }
  // odd tail sample
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
    const int64_t s = n - 1;
    f2_t c0[1] = {f2_splat(in[2 * s])}, c1[1] = {f2_splat(in[2 * s + 1])}, lsum[1] = {f2_splat(0.f)};
    for (int i = 0; i < L; ++i) {
      const int l = MODE == 0 ? L - 1 - i : i;
      if ((flip_mask >> l) & 1ull) nvp_layer<H, MODE, 1>(sp + l * PS, Hrt, c1, c0, lsum);
      else nvp_layer<H, MODE, 1>(sp + l * PS, Hrt, c0, c1, lsum);
    }
    float a0, a1, la, t;
    f2_unpack(c0[0], a0, t);
    f2_unpack(c1[0], a1, t);
    f2_unpack(lsum[0], la, t);
    if (out_v) { out_v[2 * s] = a0; out_v[2 * s + 1] = a1; }
    if (out_lp) out_lp[s] = fmaf(-0.5f, fmaf(a0, a0, a1 * a1), -2.f * kHalfLog2Pi - la);
    if (out_ls) out_ls[s] = la;
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
  size_t smem = (size_t)L * nvp_layer_floats(H) * sizeof(float);
  // persistent grid: exactly the CTAs resident at once (no partial last wave), grid-stride over tiles
  const int64_t work = ((n >> 1) + kNvpPairs - 1) / kNvpPairs;
  if (H == 4) {
    int grid = grid_resident(realnvp_chain_kernel<4, MODE>, kNvpThreads, smem, work, kNvpThreads);
    realnvp_chain_kernel<4, MODE><<<grid, kNvpThreads, smem, s>>>(params, L, H, flip_mask, in, out_v, out_lp, out_ls, n);
  } else {
    if (smem > 48 * 1024)
      FC_ENSURE_SMEM((realnvp_chain_kernel<0, MODE>), smem);
    int grid = grid_resident(realnvp_chain_kernel<0, MODE>, kNvpThreads, smem, work, kNvpThreads);
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
