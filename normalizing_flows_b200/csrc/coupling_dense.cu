// coupling_dense.cu — rows #2-#4/#6, generic event_dim and conditioner shape, FP32 CUDA-core
// path (FLOWCUDA_PRECISION_FP32): distrax.MaskedCoupling stacks of src/maf_mnist.py:52-89 and
// the C5 deep stacks.  Per flow layer: masked Linear+relu layers as tiled SGEMMs with the
// mask folded into the A-operand load, then ONE kernel that computes the last Linear only for
// the columns of the unmasked dimensions and applies the inner bijector in its epilogue — the
// [N, D·(3K+1)] conditioner output never touches HBM.  Log-det partials are reduced in a fixed
// order (deterministic).  The tcgen05 variant lives in coupling_dense_tc.cu.
#include "common.cuh"
#include "transforms.cuh"

namespace flowcuda {

constexpr int kBM = 64, kBK = 16, kBNmax = 128, kDenseThreads = 256;

// C[N,n] = act(A'[N,K] · W[K,n] + b),  A' = A with columns of parity != vis_parity zeroed
// (vis_parity < 0: no masking).
__global__ void __launch_bounds__(kDenseThreads)
dense_linear_kernel(const float* __restrict__ A, int lda, const float* __restrict__ W, const float* __restrict__ b,
                    float* __restrict__ C, int ldc, int64_t N, int K, int n, int vis_parity, int relu) {
  __shared__ float As[kBK][kBM + 4];
  __shared__ float Bs[kBK][kBNmax];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t row0 = (int64_t)blockIdx.x * kBM;
  const int col0 = blockIdx.y * kBNmax;
  float acc[4][8];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < K; k0 += kBK) {
    {  // A tile: 64 x 16, thread -> (row = tid/4, 4 consecutive k)
      int r = tid >> 2, kq = (tid & 3) * 4;
      int64_t gr = row0 + r;
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        int k = k0 + kq + e;
        float v = 0.f;
        if (gr < N && k < K && (vis_parity < 0 || (k & 1) == vis_parity)) v = A[gr * lda + k];
        As[kq + e][r] = v;
      }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {  // B tile: 16 x 128
      int idx = tid + e * kDenseThreads;
      int kr = idx >> 7, c = idx & 127;
      int k = k0 + kr, gc = col0 + c;
      Bs[kr][c] = (k < K && gc < n) ? W[(size_t)k * n + gc] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < kBK; ++kk) {
      float4 a4 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      float av[4] = {a4.x, a4.y, a4.z, a4.w};
      float bv[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) bv[j] = Bs[kk][tx + 16 * j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int64_t gr = row0 + ty * 4 + i;
    if (gr >= N) continue;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      int gc = col0 + tx + 16 * j;
      if (gc < n) {
        float v = acc[i][j] + b[gc];
        C[gr * ldc + gc] = relu ? fmaxf(v, 0.f) : v;
      }
    }
  }
}

struct DenseTrArgs {
  const float* H;      // [N, K] last hidden activations
  const float* W;      // [K, out_full]
  const float* b;      // [out_full]
  float* v;            // [N, D] state, transformed dims updated in place
  float* ldp;          // [N, tiles] log-det partials
  int64_t N;
  int K, D, ppd, out_full;
  int parity;          // transformed dims are d = parity + 2*i
  int ndims;           // number of transformed dims
  int dpt;             // dims per column tile
  int tiles;
  int nbins;
  float lo, hi;
};

// Last Linear restricted to the transformed dims' columns + inner bijector epilogue.
template <int TR, bool INV>
__global__ void __launch_bounds__(kDenseThreads) dense_transform_kernel(DenseTrArgs a) {
  __shared__ float As[kBK][kBM + 4];
  __shared__ float Bs[kBK][kBNmax];
  __shared__ float Cs[kBM][kBNmax + 1];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t row0 = (int64_t)blockIdx.x * kBM;
  const int tile = blockIdx.y;
  const int dim0 = tile * a.dpt;                       // index among transformed dims
  const int nd = min(a.dpt, a.ndims - dim0);
  const int ncols = nd * a.ppd;
  float acc[4][8];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < a.K; k0 += kBK) {
    {
      int r = tid >> 2, kq = (tid & 3) * 4;
      int64_t gr = row0 + r;
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        int k = k0 + kq + e;
        As[kq + e][r] = (gr < a.N && k < a.K) ? a.H[gr * a.K + k] : 0.f;
      }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      int idx = tid + e * kDenseThreads;
      int kr = idx >> 7, c = idx & 127;
      int k = k0 + kr;
      float w = 0.f;
      if (k < a.K && c < ncols) {
        int di = c / a.ppd, j = c - di * a.ppd;
        int d = a.parity + 2 * (dim0 + di);
        w = a.W[(size_t)k * a.out_full + (size_t)d * a.ppd + j];
      }
      Bs[kr][c] = w;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < kBK; ++kk) {
      float4 a4 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      float av[4] = {a4.x, a4.y, a4.z, a4.w};
      float bv[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) bv[j] = Bs[kk][tx + 16 * j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      int c = tx + 16 * j;
      float bias = 0.f;
      if (c < ncols) {
        int di = c / a.ppd, jj = c - di * a.ppd;
        bias = a.b[(size_t)(a.parity + 2 * (dim0 + di)) * a.ppd + jj];
      }
      Cs[ty * 4 + i][c] = acc[i][j] + bias;
    }
  __syncthreads();
  // epilogue: 4 threads per row, thread `sub` handles dims sub, sub+4, ...
  const int r = tid >> 2, sub = tid & 3;
  const int64_t gr = row0 + r;
  float ldsum = 0.f;
  if (gr < a.N) {
    for (int di = sub; di < nd; di += 4) {
      const int d = a.parity + 2 * (dim0 + di);
      const float* pr = &Cs[r][di * a.ppd];
      float xin = a.v[gr * a.D + d], nv, l1;
      if (TR == FLOWCUDA_TRANSFORM_RQS) {
        rqs_apply<INV, 0>([&](int i) { return pr[i]; }, a.nbins, a.lo, a.hi, xin, nv, l1);
      } else {
        affine_apply<INV>(pr[0], pr[1], xin, nv, l1);
      }
      a.v[gr * a.D + d] = nv;
      ldsum += l1;
    }
  }
  // fixed-order reduction over the 4 sub-threads of the row
  float s1 = __shfl_down_sync(0xffffffffu, ldsum, 1);
  float s2 = __shfl_down_sync(0xffffffffu, ldsum, 2);
  float s3 = __shfl_down_sync(0xffffffffu, ldsum, 3);
  if (sub == 0 && gr < a.N) a.ldp[gr * a.tiles + tile] = ((ldsum + s1) + s2) + s3;
}

// ScalarAffine with ONE (shift, log_scale) per sample broadcast over all dims
// (real_nvp_toy_densities.py:84-86 semantics at any D): cond[N,2] -> update unmasked dims.
template <bool INV>
__global__ void __launch_bounds__(256)
dense_scalar_affine_kernel(const float* __restrict__ cond, float* __restrict__ v, float* __restrict__ ldp, int64_t N,
                           int D, int parity, int ndims) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= N) return;
  float shift = cond[2 * r], ls = cond[2 * r + 1];
  float e = expf(INV ? -ls : ls);
  for (int d = parity; d < D; d += 2) {
    float x = v[r * D + d];
    v[r * D + d] = INV ? (x - shift) * e : fmaf(x, e, shift);
  }
  ldp[r] = (INV ? -ls : ls) * (float)ndims;
}

// ld[r] (+)= sum_t ldp[r][t] in fixed order; optionally finish with the base log-density.
__global__ void __launch_bounds__(256)
dense_ld_accum_kernel(const float* __restrict__ ldp, int tiles, float* __restrict__ ld, int64_t N, int first) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= N) return;
  float s = first ? 0.f : ld[r];
  for (int t = 0; t < tiles; ++t) s += ldp[r * tiles + t];
  ld[r] = s;
}

__global__ void __launch_bounds__(256)
dense_finish_kernel(const float* __restrict__ v, const float* __restrict__ ld, int64_t N, int D, int with_base, int base,
                    float* __restrict__ out_lp, float* __restrict__ out_ld) {
  // one warp per row: sum of base terms over D in a fixed lane order
  int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  if (r >= N) return;
  float s = 0.f;
  if (with_base)
    for (int d = lane; d < D; d += 32) s += base_term(base, v[r * D + d]);
  s = warp_sum(s);
  if (lane == 0) {
    float l = ld ? ld[r] : 0.f;
    if (out_ld) out_ld[r] = l;
    if (out_lp) out_lp[r] = l + s;
  }
}

struct DenseLayout {
  int nlin;            // number of Linear layers (num_hidden + 1)
  int sizes[FLOWCUDA_MAX_HIDDEN_LAYERS + 2];
  int64_t per_layer;
  int ppd, out_full, hmax;
};

static DenseLayout dense_layout(const flowcuda_coupling_spec* sp) {
  DenseLayout L;
  L.ppd = sp->transform == FLOWCUDA_TRANSFORM_RQS ? 3 * sp->num_bins + 1 : 2;
  L.out_full = sp->transform == FLOWCUDA_TRANSFORM_AFFINE_SCALAR ? 2 : sp->event_dim * L.ppd;
  L.nlin = sp->num_hidden + 1;
  L.sizes[0] = sp->event_dim;
  L.hmax = 2;
  for (int i = 0; i < sp->num_hidden; ++i) {
    L.sizes[i + 1] = sp->hidden[i];
    if (sp->hidden[i] > L.hmax) L.hmax = sp->hidden[i];
  }
  L.sizes[L.nlin] = L.out_full;
  L.per_layer = 0;
  for (int i = 0; i < L.nlin; ++i) L.per_layer += (int64_t)L.sizes[i] * L.sizes[i + 1] + L.sizes[i + 1];
  return L;
}

int64_t coupling_param_count(const flowcuda_coupling_spec* sp) { return dense_layout(sp).per_layer * sp->num_layers; }

static int dense_tiles(const flowcuda_coupling_spec* sp, const DenseLayout& L, int* dpt_out) {
  int dpt = kBNmax / L.ppd;
  if (dpt < 1) dpt = 0;
  if (dpt_out) *dpt_out = dpt;
  int ndmax = (sp->event_dim + 1) / 2;
  return dpt ? (ndmax + dpt - 1) / dpt : 0;
}

size_t coupling_dense_workspace_bytes(const flowcuda_coupling_spec* sp, int64_t n) {
  DenseLayout L = dense_layout(sp);
  int tiles = dense_tiles(sp, L, nullptr);
  if (tiles < 1) tiles = 1;
  size_t f = (size_t)n * sp->event_dim + 2 * (size_t)n * L.hmax + (size_t)n * tiles + (size_t)n;
  return f * sizeof(float) + 256;
}

int coupling_dense_run(const flowcuda_coupling_spec* sp, const float* params, const float* in, float* out_v,
                       float* out_lp, float* out_ld, int64_t n, int layer_begin, int layer_end, int with_base,
                       bool inverse, void* workspace, size_t workspace_bytes, cudaStream_t s) {
  DenseLayout L = dense_layout(sp);
  int dpt = 0;
  int tiles = dense_tiles(sp, L, &dpt);
  const bool scalar = sp->transform == FLOWCUDA_TRANSFORM_AFFINE_SCALAR;
  if (!scalar && dpt < 1) {
    set_error("coupling: %d params per dim exceed the %d-column tile", L.ppd, kBNmax);
    return FLOWCUDA_EUNSUPPORTED;
  }
  if (scalar) tiles = 1;
  if (workspace_bytes < coupling_dense_workspace_bytes(sp, n) || !workspace) {
    set_error("coupling: workspace %zu < %zu bytes", workspace_bytes, coupling_dense_workspace_bytes(sp, n));
    return FLOWCUDA_EWORKSPACE;
  }
  const int D = sp->event_dim;
  float* v = (float*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  float* Ha = v + (size_t)n * D;
  float* Hb = Ha + (size_t)n * L.hmax;
  float* ldp = Hb + (size_t)n * L.hmax;
  float* ld = ldp + (size_t)n * tiles;
  FC_CUDA(cudaMemcpyAsync(v, in, (size_t)n * D * sizeof(float), cudaMemcpyDeviceToDevice, s));
  const int rows_grid = (int)((n + kBM - 1) / kBM);
  const int nl = layer_end - layer_begin;
  for (int it = 0; it < nl; ++it) {
    const int l = inverse ? layer_begin + it : layer_end - 1 - it;
    // mask True (visible) at odd indices on even layers, at even indices on odd layers
    const int vis = (l & 1) ? 0 : 1;
    const int tp = 1 - vis;  // parity of the transformed dims
    const int ndims = (D - tp + 1) / 2;
    const float* p = params + L.per_layer * l;
    const float* A = v;
    int lda = D, K = D;
    float* bufs[2] = {Ha, Hb};
    for (int i = 0; i < sp->num_hidden; ++i) {
      int nout = L.sizes[i + 1];
      const float* W = p;
      const float* b = p + (size_t)K * nout;
      dim3 grid(rows_grid, (nout + kBNmax - 1) / kBNmax);
      dense_linear_kernel<<<grid, kDenseThreads, 0, s>>>(A, lda, W, b, bufs[i & 1], nout, n, K, nout, i == 0 ? vis : -1, 1);
      FC_LAUNCH_CHECK("dense_linear_kernel");
      p = b + nout;
      A = bufs[i & 1];
      lda = K = nout;
    }
    const float* W = p;
    const float* b = p + (size_t)K * L.out_full;
    if (scalar) {
      float* cond = bufs[sp->num_hidden & 1];
      dim3 grid(rows_grid, 1);
      dense_linear_kernel<<<grid, kDenseThreads, 0, s>>>(A, lda, W, b, cond, 2, n, K, 2, sp->num_hidden == 0 ? vis : -1, 0);
      FC_LAUNCH_CHECK("dense_linear_kernel(final)");
      int g = (int)((n + 255) / 256);
      if (inverse) dense_scalar_affine_kernel<true><<<g, 256, 0, s>>>(cond, v, ldp, n, D, tp, ndims);
      else dense_scalar_affine_kernel<false><<<g, 256, 0, s>>>(cond, v, ldp, n, D, tp, ndims);
      FC_LAUNCH_CHECK("dense_scalar_affine_kernel");
    } else {
      if (sp->num_hidden == 0) {
        set_error("coupling: at least one hidden layer is required");
        return FLOWCUDA_EUNSUPPORTED;
      }
      DenseTrArgs a;
      a.H = A; a.W = W; a.b = b; a.v = v; a.ldp = ldp; a.N = n; a.K = K; a.D = D; a.ppd = L.ppd;
      a.out_full = L.out_full; a.parity = tp; a.ndims = ndims; a.dpt = dpt; a.tiles = tiles;
      a.nbins = sp->num_bins; a.lo = sp->range_min; a.hi = sp->range_max;
      int used_tiles = (ndims + dpt - 1) / dpt;
      // tiles beyond used_tiles (odd D) must read as zero
      if (used_tiles < tiles) FC_CUDA(cudaMemsetAsync(ldp, 0, (size_t)n * tiles * sizeof(float), s));
      dim3 grid(rows_grid, used_tiles);
      if (sp->transform == FLOWCUDA_TRANSFORM_RQS) {
        if (inverse) dense_transform_kernel<FLOWCUDA_TRANSFORM_RQS, true><<<grid, kDenseThreads, 0, s>>>(a);
        else dense_transform_kernel<FLOWCUDA_TRANSFORM_RQS, false><<<grid, kDenseThreads, 0, s>>>(a);
      } else {
        if (inverse) dense_transform_kernel<FLOWCUDA_TRANSFORM_AFFINE, true><<<grid, kDenseThreads, 0, s>>>(a);
        else dense_transform_kernel<FLOWCUDA_TRANSFORM_AFFINE, false><<<grid, kDenseThreads, 0, s>>>(a);
      }
      FC_LAUNCH_CHECK("dense_transform_kernel");
    }
    dense_ld_accum_kernel<<<(int)((n + 255) / 256), 256, 0, s>>>(ldp, tiles, ld, n, it == 0);
    FC_LAUNCH_CHECK("dense_ld_accum_kernel");
  }
  if (nl == 0) FC_CUDA(cudaMemsetAsync(ld, 0, (size_t)n * sizeof(float), s));
  if (out_v) FC_CUDA(cudaMemcpyAsync(out_v, v, (size_t)n * D * sizeof(float), cudaMemcpyDeviceToDevice, s));
  if (out_lp || out_ld) {
    int64_t threads = n * 32;
    dense_finish_kernel<<<(int)((threads + 255) / 256), 256, 0, s>>>(v, ld, n, D, with_base, sp->base, out_lp, out_ld);
    FC_LAUNCH_CHECK("dense_finish_kernel");
  }
  return FLOWCUDA_OK;
}

}  // namespace flowcuda
