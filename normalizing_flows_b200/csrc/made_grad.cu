// made_grad.cu — rows #7/#8: value_and_grad of loss = -mean(log_prob(x)) for the MADE-conditioned MAF
// (the "train samples/s (MAF-MNIST)" step of BASELINE.json; jax.value_and_grad(loss_fn) of
// src/maf_mnist.py:134-166 applied to the north_star's MADE flow; oracle: nll_and_grads over
// oracle/flows_oracle.py:576-590).  Hand-written reverse mode, FP32 throughout:
//   forward  (per layer, stored): xp (layer input in the layer's ordering), h1, h2, o = [mu, alpha]
//   backward (per layer): dmu = -du·e^-alpha, dalpha = -du·u + 1/N, dx = du·e^-alpha + dh1·W1ᵀ,
//            dW_i = mask_i ⊙ (inᵀ·dout), db_i = colsum(dout), relu masks from the stored activations.
// Every product runs through ONE generic tiled SGEMM (strided operands, fixed-order K loop, no split-K
// atomics), so gradients are bit-reproducible run to run.  The tensor-core density path (made.cu) is
// not used here: training needs the stored FP32 activations.
#include "common.cuh"

namespace flowcuda {

__host__ __device__ __forceinline__ int mg_hdeg(int k, int H, int D) {
  const int d = (int)(((unsigned)(k + 1) * (unsigned)(D - 1) + (unsigned)H) / (unsigned)(H + 1));  // sizes <= 32768
  return d < 1 ? 1 : d;
}

// MADE mask of a Linear at (in i, out j) from degree tables (NULL table: inputs have degree i+1, outputs
// (j mod D)+1 with a STRICT comparison — the [shift | log_scale] output Linear)
struct MgMask {
  const int* din;   // [k_in] or NULL (inputs: degree i+1)
  const int* dout;  // [n_out]; for the output Linear dout[j] = j mod D (strict "<" folded in as "<= dout")
  int D, on;        // on = 0: no mask
  int wrap;         // 1: the output Linear, whose dout table restarts at column D
};
__device__ __forceinline__ int mg_din(const MgMask& m, int i) { return m.din ? m.din[i] : i + 1; }
__device__ __forceinline__ int mg_dout(const MgMask& m, int j) { return m.dout[j]; }
__device__ __forceinline__ bool mg_mask(const MgMask& m, int i, int j) { return mg_din(m, i) <= mg_dout(m, j); }
// largest dout over [j0, j1] (degree tables are non-decreasing; the output Linear wraps at D)
__device__ __forceinline__ int mg_dout_max(const MgMask& m, int j0, int j1) {
  if (m.wrap && j0 / m.D != j1 / m.D) return m.D - 1;
  return m.dout[j1];
}

struct MgGemm {
  const float* A; int64_t ars, acs;  // A(i,k) = A[i*ars + k*acs]
  const float* B; int64_t brs, bcs;  // B(k,j) = B[k*brs + j*bcs]
  float* C; int64_t ldc;             // C[i*ldc + j]
  int M, N; int64_t K;
  const float* bias;                 // + bias[j] (or NULL)
  const float* gate; int64_t ldg;    // multiply by (gate[i*ldg + j] > 0) (relu backward), or NULL
  int relu;                          // relu on the result
  int accumulate;                    // C += result
  MgMask mask;                       // mask.on: weight product, apply the MADE mask with the i/j roles below
  int splits; float* part;           // splits > 1: split-K, raw partials to part[z][M][N]; mg_splitk_reduce finishes (fixed order)
  int mask_on_b;                     // 1: mask B(k,j) as W[k (in), j (out)];  2: mask B(k,j) as Wᵀ: in = j, out = k; 0: mask C(i,j) as W[i,j]
};

constexpr int kMgT = 128, kMgK = 8, kMgSplits = 8;

// C = op(A)·op(B): 128 x 128 tile, 8 k per step, 256 threads x (8 x 8) outputs held as four 4 x 4 blocks so
// that every shared-memory read is a 128-bit load; the next step's operands are fetched into registers while
// the current step is computed (double-buffered shared memory, one barrier per step).
__global__ void __launch_bounds__(256) mg_sgemm_kernel(MgGemm g) {
  __shared__ __align__(16) float As[2][kMgK][kMgT + 4];
  __shared__ __align__(16) float Bs[2][kMgK][kMgT + 4];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int i0 = blockIdx.y * kMgT, j0 = blockIdx.x * kMgT;
  const int jl = min(j0 + kMgT, g.N) - 1;
  // a masked Linear is block-triangular in degree order: skip what is entirely masked out
  if (g.mask.on && g.mask_on_b == 0 && mg_din(g.mask, i0) > mg_dout_max(g.mask, j0, jl)) {
    if (!g.accumulate && blockIdx.z == 0) {
      for (int e = tid; e < kMgT * kMgT; e += 256) {
        const int gi = i0 + (e >> 7), gj = j0 + (e & 127);
        if (gi < g.M && gj < g.N) g.C[gi * g.ldc + gj] = 0.f;
      }
    }
    return;
  }
  // the k steps this tile needs (masks make whole k ranges vanish; split-K takes a slice)
  int64_t kbeg = 0, kend = g.K;
  if (g.splits > 1) {
    const int64_t chunk = ((g.K + g.splits - 1) / g.splits + kMgK - 1) & ~(int64_t)(kMgK - 1);
    kbeg = chunk * blockIdx.z;
    kend = min(g.K, kbeg + chunk);
  }
  if (g.mask.on && g.mask_on_b == 1) {  // B = W[k (in), j (out)]: inputs with degree above the tile's largest output degree
    const int dmax = mg_dout_max(g.mask, j0, jl);
    int64_t lo = 0, hi = g.K;  // first k with din(k) > dmax (din is non-decreasing)
    while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (mg_din(g.mask, (int)mid) > dmax) hi = mid; else lo = mid + 1; }
    kend = lo;
  } else if (g.mask.on && g.mask_on_b == 2) {  // B = Wᵀ: in = j, out = k: outputs with degree below the tile's smallest input degree
    const int dmin = mg_din(g.mask, j0);
    if (!g.mask.wrap) {
      int64_t lo = 0, hi = g.K;  // first k with dout(k) >= dmin
      while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (g.mask.dout[mid] >= dmin) hi = mid; else lo = mid + 1; }
      kbeg = lo & ~(int64_t)(kMgK - 1);
    }
  }
  // accumulators as packed pairs along j: one FFMA2 (fma.rn.f32x2) per two outputs — the 3-register scalar
  // FFMA issues at half rate on sm_100, so the packed form is what reaches the FP32 roof
  f2_t acc2[8][4];
#pragma unroll
  for (int a = 0; a < 8; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc2[a][b] = f2_splat(0.f);
  const bool a_kfast = g.acs == 1, b_jfast = g.bcs == 1;
  float ra[4], rb[4];
  // per-thread element coordinates and base pointers, hoisted out of the k loop
  const float* ap[4]; const float* bp[4];
  int ak[4], bk[4], bj[4];
  bool aok[4], bok[4];
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const int e = tid + t * 256;
    int i, k;
    if (a_kfast) { i = e >> 3; k = e & 7; } else { i = e & 127; k = e >> 7; }
    ak[t] = k; aok[t] = i0 + i < g.M;
    ap[t] = g.A + (int64_t)(i0 + i) * g.ars + (kbeg + k) * g.acs;
    int j, kb;
    if (b_jfast) { j = e & 127; kb = e >> 7; } else { j = e >> 3; kb = e & 7; }
    bk[t] = kb; bj[t] = j0 + j; bok[t] = j0 + j < g.N;
    bp[t] = g.B + (kbeg + kb) * g.brs + (int64_t)(j0 + j) * g.bcs;
  }
  const int64_t astep = (int64_t)kMgK * g.acs, bstep = (int64_t)kMgK * g.brs;
  // smallest output degree / largest input degree of the tile's j range (the k range varies per step)
  auto dout_min = [&](int lo, int hi) { return (g.mask.wrap && lo / g.mask.D != hi / g.mask.D) ? 0 : g.mask.dout[lo]; };
  auto fetch = [&](int64_t k0) {
    const int64_t off = (k0 - kbeg) / kMgK;
    bool need_mask = false;
    if (g.mask.on && g.mask_on_b) {  // uniform: is any element of this k-block x j-range masked at all?
      const int kl = (int)min(k0 + kMgK, kend) - 1;
      if (g.mask_on_b == 1) need_mask = mg_din(g.mask, kl) > dout_min(j0, jl);
      else need_mask = mg_din(g.mask, jl) > dout_min((int)k0, kl);
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      ra[t] = (aok[t] && k0 + ak[t] < kend) ? ap[t][off * astep] : 0.f;
      float bv = 0.f;
      const int64_t gkb = k0 + bk[t];
      if (bok[t] && gkb < kend) {
        bv = bp[t][off * bstep];
        if (need_mask) {
          if (g.mask_on_b == 1 && !mg_mask(g.mask, (int)gkb, bj[t])) bv = 0.f;
          if (g.mask_on_b == 2 && !mg_mask(g.mask, bj[t], (int)gkb)) bv = 0.f;
        }
      }
      rb[t] = bv;
    }
  };
  auto stash = [&](int buf) {
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int e = tid + t * 256;
      int i, k;
      if (a_kfast) { i = e >> 3; k = e & 7; } else { i = e & 127; k = e >> 7; }
      As[buf][k][i] = ra[t];
      int j, kb;
      if (b_jfast) { j = e & 127; kb = e >> 7; } else { j = e >> 3; kb = e & 7; }
      Bs[buf][kb][j] = rb[t];
    }
  };
  int buf = 0;
  if (kbeg < kend) {
    fetch(kbeg);
    stash(0);
  }
  __syncthreads();
  for (int64_t k0 = kbeg; k0 < kend; k0 += kMgK) {
    const bool more = k0 + kMgK < kend;
    if (more) fetch(k0 + kMgK);
#pragma unroll
    for (int k = 0; k < kMgK; ++k) {
      const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][k][64 + ty * 4]);
      const ulonglong2 b0 = *reinterpret_cast<const ulonglong2*>(&Bs[buf][k][tx * 4]);
      const ulonglong2 b1 = *reinterpret_cast<const ulonglong2*>(&Bs[buf][k][64 + tx * 4]);
      const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const f2_t bv[4] = {b0.x, b0.y, b1.x, b1.y};
#pragma unroll
      for (int a = 0; a < 8; ++a) {
        const f2_t as = f2_splat(av[a]);
#pragma unroll
        for (int b = 0; b < 4; ++b) acc2[a][b] = f2_fma(as, bv[b], acc2[a][b]);
      }
    }
    if (more) stash(buf ^ 1);
    __syncthreads();
    buf ^= 1;
  }
  float acc[8][8];
#pragma unroll
  for (int a = 0; a < 8; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) f2_unpack(acc2[a][b], acc[a][2 * b], acc[a][2 * b + 1]);
#pragma unroll
  for (int a = 0; a < 8; ++a) {
    const int gi = i0 + (a < 4 ? ty * 4 + a : 64 + ty * 4 + a - 4);
    if (gi >= g.M) continue;
#pragma unroll
    for (int b = 0; b < 8; ++b) {
      const int gj = j0 + (b < 4 ? tx * 4 + b : 64 + tx * 4 + b - 4);
      if (gj >= g.N) continue;
      float r = acc[a][b];
      if (g.splits > 1) {
        g.part[((int64_t)blockIdx.z * g.M + gi) * g.N + gj] = r;
        continue;
      }
      if (g.bias) r += g.bias[gj];
      if (g.relu) r = fmaxf(r, 0.f);
      if (g.gate && !(g.gate[gi * g.ldg + gj] > 0.f)) r = 0.f;
      if (g.mask.on && g.mask_on_b == 0 && !mg_mask(g.mask, gi, gj)) r = 0.f;
      float* c = g.C + gi * g.ldc + gj;
      *c = g.accumulate ? *c + r : r;
    }
  }
}

// C[i,j] = mask ⊙ sum_z part[z][i][j]   (z in ascending order: deterministic)
__global__ void __launch_bounds__(256) mg_splitk_reduce_kernel(MgGemm g) {
  const int64_t tot = (int64_t)g.M * g.N;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < tot; e += (int64_t)gridDim.x * blockDim.x) {
    const int gi = (int)(e / g.N), gj = (int)(e - (int64_t)gi * g.N);
    float r = 0.f;
    if (!g.mask.on || mg_mask(g.mask, gi, gj))
      for (int z = 0; z < g.splits; ++z) r += g.part[(int64_t)z * tot + e];
    g.C[gi * g.ldc + gj] = r;
  }
}

static int mg_launch(const MgGemm& g, cudaStream_t s) {
  dim3 grid((g.N + kMgT - 1) / kMgT, (g.M + kMgT - 1) / kMgT, g.splits > 1 ? g.splits : 1);
  mg_sgemm_kernel<<<grid, 256, 0, s>>>(g);
  FC_LAUNCH_CHECK("mg_sgemm_kernel");
  if (g.splits > 1) {
    mg_splitk_reduce_kernel<<<grid_for((int64_t)g.M * g.N, 256, 8), 256, 0, s>>>(g);
    FC_LAUNCH_CHECK("mg_splitk_reduce_kernel");
  }
  return FLOWCUDA_OK;
}

__global__ void mg_degree_kernel(int* __restrict__ deg, int H, int D) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < H) deg[k] = mg_hdeg(k, H, D);
}
__global__ void mg_out_degree_kernel(int* __restrict__ deg, int D) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < 2 * D) deg[j] = j % D;
}

// out[r, c] = in[r, rev ? D-1-c : c]
__global__ void __launch_bounds__(256) mg_copy_perm_kernel(const float* __restrict__ in, float* __restrict__ out, int64_t N,
                                                          int D, int rev, int accumulate) {
  const int64_t tot = N * D;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / D; const int c = (int)(i - r * D);
    const float v = in[r * D + (rev ? D - 1 - c : c)];
    out[i] = accumulate ? out[i] + v : v;
  }
}

// forward affine: u = (xp - mu) e^-alpha ; ld[r] (+)= -sum_q alpha   (one warp per row)
__global__ void __launch_bounds__(256) mg_affine_fwd_kernel(const float* __restrict__ xp, const float* __restrict__ o,
                                                           float* __restrict__ u, float* __restrict__ ld, int64_t N, int D,
                                                           int first) {
  const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (r >= N) return;
  float s = 0.f;
  for (int q = lane; q < D; q += 32) {
    const float mu = o[r * 2 * D + q], al = o[r * 2 * D + D + q];
    u[r * D + q] = (xp[r * D + q] - mu) * expf(-al);
    s -= al;
  }
  s = warp_sum(s);
  if (lane == 0) ld[r] = first ? s : ld[r] + s;
}

// lp[r] = ld[r] + sum_q (-u²/2 - log(2π)/2) ; du[r,q] = u[r,q] / N   (gradient of -mean(lp) wrt the base point)
__global__ void __launch_bounds__(256) mg_finish_kernel(const float* __restrict__ u, const float* __restrict__ ld,
                                                       float* __restrict__ lp, float* __restrict__ du, int64_t N, int D) {
  const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (r >= N) return;
  const float invn = 1.f / (float)N;
  float s = 0.f;
  for (int q = lane; q < D; q += 32) {
    const float z = u[r * D + q];
    s += fmaf(-0.5f * z, z, -kHalfLog2Pi);
    du[r * D + q] = z * invn;
  }
  s = warp_sum(s);
  if (lane == 0) lp[r] = ld[r] + s;
}

// backward affine: given du (in place -> dx_direct), writes dO = [dmu, dalpha]
__global__ void __launch_bounds__(256) mg_affine_bwd_kernel(const float* __restrict__ xp, const float* __restrict__ o,
                                                           float* __restrict__ du, float* __restrict__ dO, int64_t N, int D) {
  const int64_t tot = N * D;
  const float invn = 1.f / (float)N;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / D; const int q = (int)(i - r * D);
    const float mu = o[r * 2 * D + q], al = o[r * 2 * D + D + q];
    const float e = expf(-al), g = du[i];
    const float uu = (xp[i] - mu) * e;
    dO[r * 2 * D + q] = -g * e;
    dO[r * 2 * D + D + q] = fmaf(-g, uu, invn);  // d(-mean ld)/dalpha = +1/N
    du[i] = g * e;
  }
}

// db[j] = sum_r dY[r, j]  (fixed order: deterministic).  Block = 32 columns x 8 row groups: a warp reads 32
// consecutive columns of one row (coalesced); thread (g, c) sums rows r = g (mod 8) in four interleaved
// accumulators, then the 8 groups are added in a fixed order.
__global__ void __launch_bounds__(256) mg_colsum_kernel(const float* __restrict__ dY, int64_t N, int n, float* __restrict__ db) {
  __shared__ double part[8][33];
  const int c = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + c;
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
  if (j < n) {
    int64_t r = g;
    for (; r + 24 < N; r += 32) {
      a0 += (double)dY[r * n + j];
      a1 += (double)dY[(r + 8) * n + j];
      a2 += (double)dY[(r + 16) * n + j];
      a3 += (double)dY[(r + 24) * n + j];
    }
    for (; r < N; r += 8) a0 += (double)dY[r * n + j];
  }
  part[g][c] = (a0 + a1) + (a2 + a3);
  __syncthreads();
  if (g == 0 && j < n) {
    double t = 0.0;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += part[k][c];
    db[j] = (float)t;
  }
}


// ---- host side ------------------------------------------------------------------------------------
struct MgLayout {
  int D, H1, H2, L;
  int64_t w1, b1, w2, b2, w3, b3, per_layer;
  size_t per_layer_act;  // floats of stored activations per layer
  size_t off_state, off_tmp, off_ld, off_lp, off_dO, off_dH2, off_dH1, off_sum, off_deg, off_part, off_act, total;  // bytes
};

static size_t mg_al(size_t x) { return (x + 255) & ~(size_t)255; }

static bool mg_layout(const flowcuda_made_spec* sp, int64_t n, MgLayout& L) {
  if (!sp || sp->num_hidden != 2) return false;
  L.D = sp->event_dim; L.H1 = sp->hidden[0]; L.H2 = sp->hidden[1]; L.L = sp->num_layers;
  int64_t o = 0;
  L.w1 = o; o += (int64_t)L.D * L.H1; L.b1 = o; o += L.H1;
  L.w2 = o; o += (int64_t)L.H1 * L.H2; L.b2 = o; o += L.H2;
  L.w3 = o; o += (int64_t)L.H2 * 2 * L.D; L.b3 = o; o += 2 * L.D;
  L.per_layer = o;
  L.per_layer_act = (size_t)n * (3 * (size_t)L.D + L.H1 + L.H2);
  size_t off = 0;
  L.off_state = off; off = mg_al(off + (size_t)n * L.D * 4);
  L.off_tmp = off; off = mg_al(off + (size_t)n * L.D * 4);
  L.off_ld = off; off = mg_al(off + (size_t)n * 4);
  L.off_lp = off; off = mg_al(off + (size_t)n * 4);
  L.off_dO = off; off = mg_al(off + (size_t)n * 2 * L.D * 4);
  L.off_dH2 = off; off = mg_al(off + (size_t)n * L.H2 * 4);
  L.off_dH1 = off; off = mg_al(off + (size_t)n * L.H1 * 4);
  L.off_sum = off; off = mg_al(off + 1024 * 8);
  L.off_deg = off; off = mg_al(off + (size_t)(L.H1 + L.H2 + 2 * L.D) * 4);
  {
    size_t mx = (size_t)L.D * L.H1;
    if ((size_t)L.H1 * L.H2 > mx) mx = (size_t)L.H1 * L.H2;
    if ((size_t)L.H2 * 2 * L.D > mx) mx = (size_t)L.H2 * 2 * L.D;
    L.off_part = off; off = mg_al(off + mx * 4 * kMgSplits);
  }
  L.off_act = off; off = mg_al(off + L.per_layer_act * 4 * (size_t)L.L);
  L.total = off + 256;
  return true;
}

static MgGemm mg_gemm(const float* A, int64_t ars, int64_t acs, const float* B, int64_t brs, int64_t bcs, float* C,
                      int64_t ldc, int M, int N, int64_t K) {
  MgGemm g = {};
  g.A = A; g.ars = ars; g.acs = acs; g.B = B; g.brs = brs; g.bcs = bcs; g.C = C; g.ldc = ldc; g.M = M; g.N = N; g.K = K;
  return g;
}

static int mg_run(const flowcuda_made_spec* sp, const float* params, const float* x, int64_t n, float* loss, float* grad,
                  void* workspace, size_t wsb, cudaStream_t s) {
  MgLayout L;
  if (!mg_layout(sp, n, L)) {
    set_error("made_nll_grad: needs exactly two hidden layers");
    return FLOWCUDA_EUNSUPPORTED;
  }
  if (!workspace || wsb < L.total) {
    set_error("made_nll_grad: workspace %zu < %zu bytes", wsb, L.total);
    return FLOWCUDA_EWORKSPACE;
  }
  uint8_t* ws = (uint8_t*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  const int D = L.D, H1 = L.H1, H2 = L.H2, N = (int)n;
  float* S = (float*)(ws + L.off_state);
  float* T = (float*)(ws + L.off_tmp);
  float* LD = (float*)(ws + L.off_ld);
  float* LP = (float*)(ws + L.off_lp);
  float* dO = (float*)(ws + L.off_dO);
  float* dH2 = (float*)(ws + L.off_dH2);
  float* dH1 = (float*)(ws + L.off_dH1);
  float* act = (float*)(ws + L.off_act);
  const int eb = grid_for((int64_t)n * D, 256, 16), rb = (int)((n * 32 + 255) / 256);
  auto XP = [&](int l) { return act + (size_t)l * L.per_layer_act; };
  auto HA = [&](int l) { return XP(l) + (size_t)n * D; };
  auto HB = [&](int l) { return HA(l) + (size_t)n * H1; };
  auto OO = [&](int l) { return HB(l) + (size_t)n * H2; };
  int* deg1 = (int*)(ws + L.off_deg);
  int* deg2 = deg1 + H1;
  mg_degree_kernel<<<(H1 + 255) / 256, 256, 0, s>>>(deg1, H1, D);
  mg_degree_kernel<<<(H2 + 255) / 256, 256, 0, s>>>(deg2, H2, D);
  float* part = (float*)(ws + L.off_part);
  const int splits = n >= 1024 ? kMgSplits : 1;
  int* deg3 = deg2 + H2;
  mg_out_degree_kernel<<<(2 * D + 255) / 256, 256, 0, s>>>(deg3, D);
  const MgMask m0 = {nullptr, deg1, D, 1, 0}, m1 = {deg1, deg2, D, 1, 0}, m2 = {deg2, deg3, D, 1, 1};
  int rc;
  FC_CUDA(cudaMemcpyAsync(S, x, (size_t)n * D * 4, cudaMemcpyDeviceToDevice, s));
  // ---- forward, data -> base: layer L-1 first
  for (int it = 0; it < L.L; ++it) {
    const int l = L.L - 1 - it, rev = l & 1;
    const float* P = params + L.per_layer * l;
    mg_copy_perm_kernel<<<eb, 256, 0, s>>>(S, XP(l), n, D, rev, 0);
    MgGemm g = mg_gemm(XP(l), D, 1, P + L.w1, H1, 1, HA(l), H1, N, H1, D);
    g.bias = P + L.b1; g.relu = 1; g.mask = m0; g.mask_on_b = 1;
    if ((rc = mg_launch(g, s))) return rc;
    g = mg_gemm(HA(l), H1, 1, P + L.w2, H2, 1, HB(l), H2, N, H2, H1);
    g.bias = P + L.b2; g.relu = 1; g.mask = m1; g.mask_on_b = 1;
    if ((rc = mg_launch(g, s))) return rc;
    g = mg_gemm(HB(l), H2, 1, P + L.w3, 2 * D, 1, OO(l), 2 * D, N, 2 * D, H2);
    g.bias = P + L.b3; g.mask = m2; g.mask_on_b = 1;
    if ((rc = mg_launch(g, s))) return rc;
    mg_affine_fwd_kernel<<<rb, 256, 0, s>>>(XP(l), OO(l), T, LD, n, D, it == 0);
    mg_copy_perm_kernel<<<eb, 256, 0, s>>>(T, S, n, D, rev, 0);
    FC_LAUNCH_CHECK("made_nll_grad forward");
  }
  mg_finish_kernel<<<rb, 256, 0, s>>>(S, LD, LP, T, n, D);  // T = d loss / d base point (natural order)
  FC_LAUNCH_CHECK("mg_finish_kernel");
  if ((rc = flowcuda_sum_scaled(LP, n, -1.0 / (double)n, loss, ws + L.off_sum, s))) return rc;
  // ---- backward: layer 0 first
  for (int l = 0; l < L.L; ++l) {
    const int rev = l & 1;
    const float* P = params + L.per_layer * l;
    float* G = grad + L.per_layer * l;
    mg_copy_perm_kernel<<<eb, 256, 0, s>>>(T, S, n, D, rev, 0);  // S = gradient in the layer's ordering
    mg_affine_bwd_kernel<<<eb, 256, 0, s>>>(XP(l), OO(l), S, dO, n, D);
    FC_LAUNCH_CHECK("mg_affine_bwd_kernel");
    // output Linear
    MgGemm g = mg_gemm(HB(l), 1, H2, dO, 2 * D, 1, G + L.w3, 2 * D, H2, 2 * D, n);
    g.mask = m2; g.splits = splits; g.part = part;
    if ((rc = mg_launch(g, s))) return rc;
    mg_colsum_kernel<<<(2 * D + 31) / 32, 256, 0, s>>>(dO, n, 2 * D, G + L.b3);
    g = mg_gemm(dO, 2 * D, 1, P + L.w3, 1, 2 * D, dH2, H2, N, H2, 2 * D);
    g.gate = HB(l); g.ldg = H2; g.mask = m2; g.mask_on_b = 2;
    if ((rc = mg_launch(g, s))) return rc;
    // hidden 2
    g = mg_gemm(HA(l), 1, H1, dH2, H2, 1, G + L.w2, H2, H1, H2, n);
    g.mask = m1; g.splits = splits; g.part = part;
    if ((rc = mg_launch(g, s))) return rc;
    mg_colsum_kernel<<<(H2 + 31) / 32, 256, 0, s>>>(dH2, n, H2, G + L.b2);
    g = mg_gemm(dH2, H2, 1, P + L.w2, 1, H2, dH1, H1, N, H1, H2);
    g.gate = HA(l); g.ldg = H1; g.mask = m1; g.mask_on_b = 2;
    if ((rc = mg_launch(g, s))) return rc;
    // hidden 1
    g = mg_gemm(XP(l), 1, D, dH1, H1, 1, G + L.w1, H1, D, H1, n);
    g.mask = m0; g.splits = splits; g.part = part;
    if ((rc = mg_launch(g, s))) return rc;
    mg_colsum_kernel<<<(H1 + 31) / 32, 256, 0, s>>>(dH1, n, H1, G + L.b1);
    g = mg_gemm(dH1, H1, 1, P + L.w1, 1, H1, S, D, N, D, H1);
    g.accumulate = 1; g.mask = m0; g.mask_on_b = 2;
    if ((rc = mg_launch(g, s))) return rc;
    mg_copy_perm_kernel<<<eb, 256, 0, s>>>(S, T, n, D, rev, 0);  // back to the natural order
    FC_LAUNCH_CHECK("made_nll_grad backward");
  }
  return FLOWCUDA_OK;
}

}  // namespace flowcuda

using namespace flowcuda;

extern "C" {

size_t flowcuda_made_nll_grad_workspace_bytes(const flowcuda_made_spec* spec, int64_t n) {
  MgLayout L;
  if (n < 0 || !mg_layout(spec, n, L)) return 0;
  return L.total;
}

int flowcuda_made_nll_grad(const flowcuda_made_spec* spec, const float* params, const float* x, int64_t n, float* loss,
                           float* grad, void* workspace, size_t workspace_bytes, void* stream) {
  FC_REQUIRE(spec && params && x && loss && grad, "made_nll_grad: NULL pointer");
  FC_REQUIRE(n >= 1 && n < (1ll << 31), "made_nll_grad: n=%lld outside [1, 2^31)", (long long)n);
  FC_REQUIRE(spec->event_dim >= 1 && spec->num_layers >= 1, "made_nll_grad: bad spec");
  FC_REQUIRE(spec->event_dim <= 32768 && spec->hidden[0] <= 32768 && spec->hidden[1] <= 32768, "made_nll_grad: sizes above 32768");
  return mg_run(spec, params, x, n, loss, grad, workspace, workspace_bytes, (cudaStream_t)stream);
}

}  // extern "C"
