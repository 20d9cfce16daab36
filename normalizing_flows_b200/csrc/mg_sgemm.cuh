// mg_sgemm.cuh — the generic FP32 SGEMM of the hand-written reverse-mode paths (made_grad.cu,
// coupling_dense_grad.cu): C = op(A)·op(B) with strided operands (so dY·Wᵀ and Xᵀ·dY need no transposes),
// optional bias / relu / relu-gate epilogue, optional MADE degree masks with block-triangular skipping, and
// deterministic split-K.  128 x 128 tile, 8 k per step, 8 x 8 outputs per thread as packed f32x2
// accumulators, double-buffered shared memory.
#pragma once
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

// Block-strided column subset: compact column j stands for full column off + (j / blk)*stride + j % blk.
// (The last Linear of a coupling layer only needs the columns of the dims it transforms: blocks of P every 2P.)
// Understood by the tensor-core engine only (mgtc.cuh); mode 0 = none.
struct MgColMap {
  int mode;  // 1: B's columns j and the bias; 2: B's K index; 3: C's columns
  int blk, stride, off;
};
__host__ __device__ __forceinline__ int mg_full_col(const MgColMap& m, int j) {
  const int b = j / m.blk;
  return m.off + b * m.stride + (j - b * m.blk);
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
  MgColMap map;                      // tensor-core engine only
};

constexpr int kMgT = 128, kMgK = 8, kMgSplits = 8;

// C = op(A)·op(B): 128 x 128 tile, 8 k per step, 256 threads x (8 x 8) outputs held as four 4 x 4 blocks so
// that every shared-memory read is a 128-bit load; the next step's operands are fetched into registers while
// the current step is computed (double-buffered shared memory, one barrier per step).
static __global__ void __launch_bounds__(256) mg_sgemm_kernel(MgGemm g) {
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
static __global__ void __launch_bounds__(256) mg_splitk_reduce_kernel(MgGemm g) {
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
  if (g.map.mode) {
    set_error("mg_launch: column maps need the tensor-core engine");
    return FLOWCUDA_EUNSUPPORTED;
  }
  dim3 grid((g.N + kMgT - 1) / kMgT, (g.M + kMgT - 1) / kMgT, g.splits > 1 ? g.splits : 1);
  mg_sgemm_kernel<<<grid, 256, 0, s>>>(g);
  FC_LAUNCH_CHECK("mg_sgemm_kernel");
  if (g.splits > 1) {
    mg_splitk_reduce_kernel<<<grid_for((int64_t)g.M * g.N, 256, 8), 256, 0, s>>>(g);
    FC_LAUNCH_CHECK("mg_splitk_reduce_kernel");
  }
  return FLOWCUDA_OK;
}

// db[j] = sum_r dY[r, j]  (fixed order: deterministic), two stages so that the grid fills the machine:
// stage 1, block (column block x, row slice z) = 32 columns x 8 row groups: a warp reads 32 consecutive columns
// of one row (coalesced); thread (g, c) sums rows r = g (mod 8) of its slice in four interleaved double
// accumulators, the 8 groups are added in a fixed order -> part[z][j]; stage 2 adds the slices in order.
constexpr int kMgColSlices = 32;
static __global__ void __launch_bounds__(256) mg_colsum_stage1(const float* __restrict__ dY, int64_t N, int n,
                                                              double* __restrict__ part) {
  __shared__ double sm[8][33];
  const int c = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + c;
  const int64_t chunk = ((N + gridDim.y - 1) / gridDim.y + 7) & ~(int64_t)7;
  const int64_t r0 = chunk * blockIdx.y, r1 = min(N, r0 + chunk);
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
  if (j < n) {
    int64_t r = r0 + g;
    for (; r + 24 < r1; r += 32) {
      a0 += (double)dY[r * n + j];
      a1 += (double)dY[(r + 8) * n + j];
      a2 += (double)dY[(r + 16) * n + j];
      a3 += (double)dY[(r + 24) * n + j];
    }
    for (; r < r1; r += 8) a0 += (double)dY[r * n + j];
  }
  sm[g][c] = (a0 + a1) + (a2 + a3);
  __syncthreads();
  if (g == 0 && j < n) {
    double t = 0.0;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += sm[k][c];
    part[(int64_t)blockIdx.y * n + j] = t;
  }
}
static __global__ void __launch_bounds__(256) mg_colsum_stage2(const double* __restrict__ part, int slices, int n,
                                                              float* __restrict__ db, MgColMap map) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  double t = 0.0;
  for (int z = 0; z < slices; ++z) t += part[(int64_t)z * n + j];
  db[map.mode ? mg_full_col(map, j) : j] = (float)t;
}
// scratch: kMgColSlices * n doubles
static int mg_colsum(const float* dY, int64_t N, int n, float* db, void* scratch, cudaStream_t s, MgColMap map = MgColMap{}) {
  int slices = (int)((N + 255) / 256);
  if (slices > kMgColSlices) slices = kMgColSlices;
  if (slices < 1) slices = 1;
  mg_colsum_stage1<<<dim3((n + 31) / 32, slices), 256, 0, s>>>(dY, N, n, (double*)scratch);
  FC_LAUNCH_CHECK("mg_colsum_stage1");
  mg_colsum_stage2<<<(n + 255) / 256, 256, 0, s>>>((const double*)scratch, slices, n, db, map);
  FC_LAUNCH_CHECK("mg_colsum_stage2");
  return FLOWCUDA_OK;
}

static MgGemm mg_gemm(const float* A, int64_t ars, int64_t acs, const float* B, int64_t brs, int64_t bcs, float* C,
                      int64_t ldc, int M, int N, int64_t K) {
  MgGemm g = {};
  g.A = A; g.ars = ars; g.acs = acs; g.B = B; g.brs = brs; g.bcs = bcs; g.C = C; g.ldc = ldc; g.M = M; g.N = N; g.K = K;
  return g;
}

}  // namespace flowcuda
