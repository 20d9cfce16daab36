// dense_tc.cuh — the tcgen05 bf16x3 GEMM shared by the wide-conditioner flows
// (coupling_dense_tc.cu: distrax.MaskedCoupling stacks, src/maf_mnist.py:34-89;
//  made.cu: MADE-affine MAF).
//
// Every Linear is C[128 x 256 tile] = A · W with bf16 operands in error-compensated 3-pass form
// (hi·hi + hi·lo + lo·hi, fp32 accumulate in TMEM; single-pass bf16 misses the 1e-3 log_prob
// budget by 7x — profiles/r01/precision_study.txt).  Operands live in HBM as ready-made
// shared-memory tile images (tc.cuh), so a pipeline stage is two cp.async.bulk (TMA) copies:
//   A image  [m_tile][k_stage][hi|lo][4 chunks][128 rows][16 B]   (activations, 32 K per stage)
//   B image  [n_tile][k_stage][hi|lo][4 chunks][256 rows][16 B]   (weights, masks folded in)
// CTA = loader warp + MMA warp + 2 x 8 epilogue warps, persistent (one per SM) with two TMEM accumulators so
// that a tile's epilogue overlaps the next tile's MMAs.  Epilogues (thread = sample row = TMEM lane):
//   hidden: +bias, relu, bf16 hi/lo split, written straight into the next Linear's A image;
//   last:   the inner bijector (RQS / affine) on the tile's dims, state updated in place,
//           log-det partial per (row, tile) — the [N, D*params] conditioner output never exists.
// A masked (MADE) Linear is block lower-triangular once its units are sorted by degree, so an
// n tile only walks the first ks_tile[nt] K stages: the all-zero weight blocks are never loaded.
#pragma once
#include "tc.cuh"
#include "transforms.cuh"

namespace flowcuda {

constexpr int kDtBN = 256, kDtBK = 32;
// threads per CTA: loader warp, MMA warp, two epilogue groups of DtCfg<EPI>::kEpiWarps warps
constexpr int kDtABytes = 128 * kDtBK * 2;   // one of hi/lo of an A stage (8 KB)
constexpr int kDtBBytes = kDtBN * kDtBK * 2; // one of hi/lo of a B stage (16 KB)
constexpr int kDtStageBytes = 2 * kDtABytes + 2 * kDtBBytes;  // 48 KB
constexpr int kDtMaxKsTiles = 48;

// ---- activation packing -------------------------------------------------------------------
// v[N, D] fp32 (flow state) -> A image whose K index j reads dim d = off + stride*j (j < nk),
// K padded to KS*32.  Coupling layers: off = visible parity, stride 2.  MADE: (0, 1) or, for the
// layers that see the reversed ordering, (D-1, -1).
static __global__ void __launch_bounds__(256) dense_tc_pack_x_kernel(const float* __restrict__ v, int64_t N, int D,
                                                                    int off, int stride, int nk, int KS,
                                                                    uint8_t* __restrict__ img) {
  // v is written by the previous kernel of the chain.  No early pdl_launch_dependents() here: the dependent is a GEMM
  // whose resident CTAs (200 KB of shared memory each) shrink L1 to its minimum while this kernel's strided reads of
  // v want it — measured 1.88 -> 2.41 ms on the 16-layer affine stack at batch 16384 with an early trigger.
  pdl_wait();
  const int64_t rows = (N + 255) / 256 * 256;  // whole CTA-pair tiles (dense_tc2.cuh): MT is always even
  const int64_t total = rows * (KS * 4);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = i % rows;
    const int ck = (int)(i / rows);  // global chunk index along K (8 elements each)
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int j = ck * 8 + e;
      f[e] = (row < N && j < nk) ? v[row * D + off + stride * j] : 0.f;
    }
    uint32_t h[4], l[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) tc::split_bf16x2(f[2 * e], f[2 * e + 1], h[e], l[e]);
    const int64_t mt = row >> 7;
    const int ks = ck >> 2, c = ck & 3;
    uint8_t* st = img + ((mt * KS + ks) * 2) * (int64_t)kDtABytes + tc::img_off(128, (int)(row & 127), c);
    *reinterpret_cast<uint4*>(st) = make_uint4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<uint4*>(st + kDtABytes) = make_uint4(l[0], l[1], l[2], l[3]);
  }
}

// ---- the GEMM ---------------------------------------------------------------------------------
struct DtGemm {
  const uint8_t* a_img;  // [m_tiles][KS][hi|lo][8 KB]
  const uint8_t* b_img;  // [n_tiles][KS][hi|lo][16 KB]
  const float* bias;     // this Linear's bias (inside the flow parameters)
  int KS;                // K stages of the images (the image strides)
  int n_out;             // valid output columns (hidden epilogue) / conditioner width (transform epilogue)
  uint8_t* out_img;      // hidden epilogue: next Linear's A image, out_KS stages
  int out_KS;
  float* v;              // transform epilogue: state [N, D], updated in place
  float* ldp;            // [N, tiles] log-det partials
  int64_t N;
  int D, tp, dstride;    // tile dim index `dim` -> state dim d = tp + dstride*dim
  int ndims, dpt, ppd, tiles;
  int next_rev;          // affine epilogue with out_img: the next layer reads the dims reversed
  int made, rev;         // made: bias columns are [shift_0..shift_{D-1}, log_scale_0..]; rev: this layer sees dims reversed
  float lo, hi;
  float koff[9];         // RQS epilogue: lo + i * min_bin_size (constant-bank operands of rqs_apply_mufu)
  unsigned char ks_tile[kDtMaxKsTiles];  // K stages n tile nt really needs (0 = all KS)
  // Training forward (coupling_dense_grad.cu): what the reverse mode reads back, stored by the same epilogues.
  float* store_h;        // hidden epilogue: relu(A·W + b) as fp32 [N, store_ld] (NULL: not stored)
  int64_t store_ld;
  float* store_o;        // RQS epilogue: the RAW accumulators (A·W, no bias), fp32 [N, store_ldo]: 32 floats per transformed
  int64_t store_ldo;     //   dim (its 25 spline parameters + 7 zeros), dim k of the layer at column 32*k
};

constexpr int kEpiHidden = 0, kEpiRqs8 = 1, kEpiAffine = 2;
// the RQS epilogue that also stores the conditioner accumulators (training forward).  Its own instantiation: compiled
// into the density kernel, the store path cost that kernel 50 registers and 5 % of its cycles with the pointer NULL
constexpr int kEpiRqs8Store = 3;
__host__ __device__ constexpr bool dt_is_rqs(int epi) { return epi == kEpiRqs8 || epi == kEpiRqs8Store; }

// Diagnostics (tools/dt_trace.py; compiled only into a -DFLOWCUDA_TRACE build, never into the shipped library): where the
// single-thread roles of dense_tc_gemm_kernel spend their cycles.  Per CTA 8 counters: kernel cycles, issuer waiting for
// a free accumulator / for operands, producer waiting for a free stage, epilogue warp 0 / warp EW waiting for an accumulator,
// and their item counts.  n-half (hidden) launches write a second block of 148 x 8 counters behind the first.
#ifdef FLOWCUDA_TRACE
__device__ unsigned long long* g_dt_trace = nullptr;
#define DT_TRACE_T0(var) const long long var = clock64()
#define DT_TRACE_ADD(acc, var) acc += clock64() - var
#else
#define DT_TRACE_T0(var)
#define DT_TRACE_ADD(acc, var)
#endif

// Column layout of a transform tile (last Linear): the 256 columns are two halves of 128; half g holds the
// dpt/2 dims [g*dpt/2, (g+1)*dpt/2) of the tile, ppd consecutive parameters each (affine: 64 dims x 2 = 128 columns;
// RQS with 8 bins: 5 dims x 25 = 125 columns + 3 pad).  Halves are self-contained so that a CTA can compute ONE half
// of a tile (128-column MMA) when the tail of the tile list is split (dense_tc_gemm_kernel, TS).
__host__ __device__ __forceinline__ bool dt_col_to_dim(int n, int dpt, int ppd, int& di, int& pp) {
  const int dph = dpt >> 1, r = n & 127;
  const int q = r / ppd;
  di = (n >> 7) * dph + q;
  pp = r - q * ppd;
  return q < dph;
}
__host__ __device__ __forceinline__ int dt_dim_to_col(int di, int dpt, int ppd) {
  const int dph = dpt >> 1;
  return ((di / dph) << 7) + (di % dph) * ppd;
}

// accumulator columns n tile `nt` really needs (a multiple of 16, the MMA N granularity): all 256 except for the last,
// ragged tile of a transform Linear, whose unused columns hold zero weights — a narrower MMA skips them
__device__ __forceinline__ int dt_tile_cols(const DtGemm& a, int nt) {
  if (a.dpt <= 0) return kDtBN;
  const int vd = a.ndims - nt * a.dpt;  // valid dims of this tile
  if (vd >= a.dpt) return kDtBN;
  const int dph = a.dpt >> 1;
  const int cols = vd <= dph ? vd * a.ppd : 128 + (vd - dph) * a.ppd;
  const int n16 = (cols + 15) & ~15;
  return n16 < 16 ? 16 : n16;
}

// bias index of tile column n (transform epilogues)
__device__ __forceinline__ float dt_tile_bias(const DtGemm& a, int nt, int n) {
  int di, pp;
  if (!dt_col_to_dim(n, a.dpt, a.ppd, di, pp)) return 0.f;
  const int dim = nt * a.dpt + di;
  if (dim >= a.ndims) return 0.f;
  if (a.made) {
    const int q = a.rev ? a.D - 1 - dim : dim;
    return a.bias[pp * a.D + q];
  }
  return a.bias[(a.tp + a.dstride * dim) * a.ppd + pp];
}

// Epilogue of tile columns [c_lo, c_hi) of an accumulator (thread = sample row = TMEM lane).  `tb` is the TMEM address
// of this warp's lane quarter of the accumulator, whose column 0 holds tile column `cshift` (0, or 128 for the second
// half of a tile computed as a half item); `T` is this warp's 32 x 36-float transposition scratch (affine).  Column
// ranges are multiples of 16 (hidden) / 32 (affine); a warp's log-det partial goes to slot 4*nt + c_lo/64 of its row.
template <int EPI, bool INV>
__device__ __forceinline__ void dt_epilogue(const DtGemm& a, int mt, int nt, uint32_t tb, const float* __restrict__ sbias,
                                            float* T, int warp, int lane, int c_lo, int c_hi, int cshift) {
  const int row = (warp & 3) * 32 + lane;
  const int64_t grow = (int64_t)mt * 128 + row;
  tb -= (uint32_t)cshift;
  if (EPI == kEpiHidden) {
    uint8_t* og = a.out_img + (int64_t)mt * a.out_KS * 2 * kDtABytes;
#pragma unroll 1
    for (int c = c_lo; c < c_hi; c += 16) {
      uint32_t vv[16];
      tc::tmem_ld_x16(tb + (uint32_t)c, vv);
      tc::tmem_ld_wait();
      uint32_t h[8], l[8];
      float fv[16];
#pragma unroll
      for (int e = 0; e < 16; ++e) fv[e] = fmaxf(__uint_as_float(vv[e]) + sbias[c + e], 0.f);
#pragma unroll
      for (int e = 0; e < 8; ++e) tc::split_bf16x2(fv[2 * e], fv[2 * e + 1], h[e], l[e]);
      if (a.store_h != nullptr && grow < a.N) {  // training: the reverse mode's relu gate and weight-gradient operand
        const int col = nt * kDtBN + c;
        float* dst = a.store_h + grow * a.store_ld + col;
        if (col + 16 <= a.n_out && ((uintptr_t)dst & 15) == 0) {
#pragma unroll
          for (int q = 0; q < 4; ++q)
            reinterpret_cast<float4*>(dst)[q] = make_float4(fv[4 * q], fv[4 * q + 1], fv[4 * q + 2], fv[4 * q + 3]);
        } else {
#pragma unroll
          for (int e = 0; e < 16; ++e)
            if (col + e < a.n_out) dst[e] = fv[e];
        }
      }
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int ck = (nt * kDtBN + c) / 8 + q;  // chunk index along the next Linear's K
        if (ck < a.out_KS * 4) {
          uint8_t* st = og + (int64_t)(ck >> 2) * 2 * kDtABytes + tc::img_off(128, row, ck & 3);
          *reinterpret_cast<uint4*>(st) = make_uint4(h[4 * q], h[4 * q + 1], h[4 * q + 2], h[4 * q + 3]);
          *reinterpret_cast<uint4*>(st + kDtABytes) = make_uint4(l[4 * q], l[4 * q + 1], l[4 * q + 2], l[4 * q + 3]);
        }
      }
    }
  } else {
    // Affine epilogue.  The accumulator comes out of TMEM as thread = sample row, but the state v[N, D]
    // is row-major: a thread-per-row access touches 32 cache lines per request.  Each warp therefore
    // transposes its 32 rows x 32 columns through shared memory (the pipeline stages are idle by now)
    // and applies the bijector as lane = dimension, so the state is read and written in 64-byte runs;
    // a third phase (thread = row again) sums the log-det and, for MADE stacks, writes the NEXT
    // layer's activation image directly (bf16 hi/lo), so no separate packing pass re-reads v.
    constexpr int TS = 36;  // row stride of the transposition tile in floats (16-byte aligned rows)
    const int r0 = (warp & 3) * 32;
    const int k = lane & 15, hp = lane >> 4;
    uint8_t* og = a.out_img ? a.out_img + (int64_t)mt * a.out_KS * 2 * kDtABytes : nullptr;
    float ldsum = 0.f;
#pragma unroll 1
    for (int c = c_lo; c < c_hi; c += 32) {
      const int dim0 = nt * 128 + c / 2;
      if (dim0 >= a.ndims) break;  // uniform across the CTA
      uint32_t vv[32];
      tc::tmem_ld_x32(tb + (uint32_t)c, vv);
      tc::tmem_ld_wait();
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        float4 t4;
        t4.x = __uint_as_float(vv[4 * e]) + sbias[c + 4 * e];
        t4.y = __uint_as_float(vv[4 * e + 1]) + sbias[c + 4 * e + 1];
        t4.z = __uint_as_float(vv[4 * e + 2]) + sbias[c + 4 * e + 2];
        t4.w = __uint_as_float(vv[4 * e + 3]) + sbias[c + 4 * e + 3];
        *reinterpret_cast<float4*>(T + lane * TS + 4 * e) = t4;
      }
      __syncwarp();
      {  // lane = (dimension k, row parity hp)
        const int dim = dim0 + k;
        const bool dvalid = dim < a.ndims;
        const int d = a.tp + a.dstride * dim;
        // this lane's column of the state: rows r0 + hp, r0 + hp + 2, ...  — one pointer walked by a constant
        // stride (per-element 64-bit index arithmetic was a quarter of the kernel's instructions)
        const int64_t row0 = (int64_t)mt * 128 + r0;
        const int64_t left = a.N - row0;
        const int nrows = dvalid ? (left < 32 ? (left < 0 ? 0 : (int)left) : 32) : 0;  // rows of this block the lane may touch
        float* vp = a.v + (row0 + hp) * a.D + d;
        const int64_t rstride = 2 * (int64_t)a.D;
        float xv[16];
        {
          const float* q = vp;
#pragma unroll
          for (int i = 0; i < 16; ++i, q += rstride) xv[i] = (2 * i + hp < nrows) ? *q : 0.f;
        }
#pragma unroll
        for (int i = 0; i < 16; ++i, vp += rstride) {
          const int r = 2 * i + hp;
          const float2 p = *reinterpret_cast<const float2*>(T + r * TS + 2 * k);
          float nv, l1;
          affine_apply<INV>(p.x, p.y, xv[i], nv, l1);
          if (r < nrows) {
            *vp = nv;
          } else {
            nv = 0.f;
            l1 = 0.f;
          }
          *reinterpret_cast<float2*>(T + r * TS + 2 * k) = make_float2(nv, l1);
        }
      }
      __syncwarp();
      {  // thread = row again: log-det partial, next layer's activation image
        float u[16];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float4 t4 = *reinterpret_cast<const float4*>(T + lane * TS + 4 * e);
          u[2 * e] = t4.x;
          u[2 * e + 1] = t4.z;
          ldsum += t4.y + t4.w;
        }
        if (og != nullptr) {
          uint32_t h[8], l[8];
          int ck;
          if (a.next_rev) {
#pragma unroll
            for (int e = 0; e < 8; ++e) tc::split_bf16x2(u[15 - 2 * e], u[14 - 2 * e], h[e], l[e]);
            ck = (a.D - 16 - dim0) >> 3;
          } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) tc::split_bf16x2(u[2 * e], u[2 * e + 1], h[e], l[e]);
            ck = dim0 >> 3;
          }
#pragma unroll
          for (int q = 0; q < 2; ++q) {
            if (ck + q < 0 || ck + q >= a.out_KS * 4) continue;  // beyond the next Linear's K range (its padding stays zero)
            uint8_t* st = og + (int64_t)((ck + q) >> 2) * 2 * kDtABytes + tc::img_off(128, row, (ck + q) & 3);
            *reinterpret_cast<uint4*>(st) = make_uint4(h[4 * q], h[4 * q + 1], h[4 * q + 2], h[4 * q + 3]);
            *reinterpret_cast<uint4*>(st + kDtABytes) = make_uint4(l[4 * q], l[4 * q + 1], l[4 * q + 2], l[4 * q + 3]);
          }
        }
      }
      __syncwarp();
    }
    if (grow < a.N) a.ldp[grow * a.tiles + 4 * nt + (c_lo >> 6)] = ldsum;
  }
}

// RQS epilogue by one warp per TMEM lane quarter (thread = sample row; 25 spline parameters per dim, two 128-column
// halves of 5 dims per tile).  The warp takes the tile's dims [d_lo, d_hi) (at most 6), paired (d, d + 1): a pair is
// evaluated as two independent splines in lockstep (rqs_apply_mufu_x2), their 2 x 25 accumulator columns fetched
// together.  `cshift` = tile column held at accumulator column 0 (128 for a half item of half 1).  The row's state
// values are loaded BEFORE the accumulator is waited for (dt_rqs_load_x): they do not depend on it, and the
// thread-per-row access (32 cache lines per warp request) used to sit inside every spline.  The log-det partial of the
// warp's dims goes to slot `ldslot` of the row's four slots of tile nt.
struct DtRqsX {
  float xa[3], xb[3];  // first / second member of pair j = 0..2
};
__device__ __forceinline__ void dt_rqs_load_x(const DtGemm& a, int mt, int nt, int d_lo, int d_hi, int warp, int lane, DtRqsX& x) {
  const int64_t grow = (int64_t)mt * 128 + (warp & 3) * 32 + lane;
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const int d0 = d_lo + 2 * j, d1 = d0 + 1;
    const int dim0 = nt * 10 + d0, dim1 = nt * 10 + d1;
    x.xa[j] = (d0 < d_hi && grow < a.N && dim0 < a.ndims) ? a.v[grow * a.D + a.tp + a.dstride * dim0] : 0.f;
    x.xb[j] = (d1 < d_hi && grow < a.N && dim1 < a.ndims) ? a.v[grow * a.D + a.tp + a.dstride * dim1] : 0.f;
  }
}

template <bool INV, bool STORE = false>
__device__ __forceinline__ void dt_epilogue_rqs_tile(const DtGemm& a, int mt, int nt, int d_lo, int d_hi, int ldslot, int cshift,
                                                     uint32_t tb, const float* __restrict__ sbias, int warp, int lane, DtRqsX& x) {
  const int row = (warp & 3) * 32 + lane;
  const int64_t grow = (int64_t)mt * 128 + row;
  float ldsum[2] = {0.f, 0.f};
  // the loop is not unrolled (one copy of the spline code): the state values are consumed from the head of two
  // register queues that shift by one per iteration — a runtime index would put the arrays in local memory
#pragma unroll 1
  for (int d0 = d_lo; d0 < d_hi; d0 += 2) {
    const int d1 = d0 + 1;
    const int dim0 = nt * 10 + d0, dim1 = nt * 10 + d1;
    if (dim0 >= a.ndims) break;  // uniform across the CTA
    const bool two = d1 < d_hi && dim1 < a.ndims;
    const int c0 = dt_dim_to_col(d0, 10, 25), c1 = dt_dim_to_col(two ? d1 : d0, 10, 25);
    uint32_t vv[2][25];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const uint32_t c = (uint32_t)((u ? c1 : c0) - cshift);
      tc::tmem_ld_x16(tb + c, vv[u]);
      tc::tmem_ld_x8(tb + c + 16, vv[u] + 16);
      tc::tmem_ld_x1(tb + c + 24, vv[u] + 24);
    }
    tc::tmem_ld_wait();
    if (grow < a.N) {
      if (STORE && a.store_o != nullptr) {
        // training: the spline adjoint reads the conditioner output back.  The RAW accumulators of a dim (the reader
        // adds the bias) go out from the registers the spline is about to consume, 32 floats per (row, dim) — 25
        // values + 7 zeros — so that every store is a whole, aligned 32-byte sector (STG.256); no second pass over TMEM
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          if (u == 1 && !two) break;
          float* dst = a.store_o + grow * a.store_ldo + (int64_t)(u ? dim1 : dim0) * 32;
          const uint32_t tail[8] = {vv[u][24], 0u, 0u, 0u, 0u, 0u, 0u, 0u};
          st_global_v8(dst, vv[u]);
          st_global_v8(dst + 8, vv[u] + 8);
          st_global_v8(dst + 16, vv[u] + 16);
          st_global_v8(dst + 24, tail);
        }
      }
      float o[2][25];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const float* sb = sbias + (u ? c1 : c0);
#pragma unroll
        for (int i = 0; i < 25; ++i)
          o[u][i] = i < 16 ? fmaf(__uint_as_float(vv[u][i]), kLog2e, sb[i]) : __uint_as_float(vv[u][i]) + sb[i];
      }
      const float v2[2] = {x.xa[0], two ? x.xb[0] : 0.5f * (a.lo + a.hi)};
      float nv[2], l2[2];
      rqs_apply_mufu_x2<INV, 8>(o[0], o[1], a.koff, a.lo, a.hi, v2, nv, l2);
      a.v[grow * a.D + a.tp + a.dstride * dim0] = nv[0];
      ldsum[0] += l2[0];
      if (two) {
        a.v[grow * a.D + a.tp + a.dstride * dim1] = nv[1];
        ldsum[1] += l2[1];
      }
      if (a.out_img != nullptr) {
        // the dims this layer transforms are exactly the dims the NEXT layer's conditioner sees (the masks alternate):
        // write the new values straight into its activation image (K index = dim index, bf16 hi/lo), so that no
        // packing pass re-reads v between the layers
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          if (u == 1 && !two) break;
          const int jj = u ? dim1 : dim0;
          const __nv_bfloat16 h = __float2bfloat16_rn(nv[u]);
          const __nv_bfloat16 l = __float2bfloat16_rn(nv[u] - __bfloat162float(h));
          uint8_t* st = a.out_img + ((int64_t)mt * a.out_KS + (jj >> 5)) * 2 * kDtABytes + tc::img_off(128, row, (jj >> 3) & 3) + (jj & 7) * 2;
          *reinterpret_cast<__nv_bfloat16*>(st) = h;
          *reinterpret_cast<__nv_bfloat16*>(st + kDtABytes) = l;
        }
      }
    }
    x.xa[0] = x.xa[1];
    x.xa[1] = x.xa[2];
    x.xb[0] = x.xb[1];
    x.xb[1] = x.xb[2];
  }
  if (grow < a.N) a.ldp[grow * a.tiles + 4 * nt + ldslot] = ldsum[0] + ldsum[1];
}

// dims of a whole RQS tile that epilogue group `g` takes on the CTA's w-th item: 10 dims = 5 pairs for two groups — three
// pairs and two, alternating between the groups from item to item so that each averages 2.5 (the accumulators' double
// buffering absorbs the one-pair difference)
__device__ __forceinline__ void dt_rqs_group_dims(int w, int g, int& d_lo, int& d_hi) {
  const int cut = (w & 1) ? 4 : 6;
  d_lo = g == 0 ? 0 : cut;
  d_hi = g == 0 ? cut : 10;
}

constexpr int kDtTileScratchFloats = 8 * 32 * 36;  // per-warp 32 x 36 transposition tiles of one epilogue group (affine)
// The affine epilogue needs the transposition scratch (two epilogue groups) and runs a 3-stage ring; the
// other epilogues have the room for 4 stages.
template <int EPI, bool WIDE> struct DtCfg {
  // Warps per epilogue group.  Hidden and affine epilogues are bound by single-warp issue: eight warps share a
  // tile (two per TMEM lane quarter, 128 columns each; maf_made log_prob 1.28 -> 1.10 ms).  The RQS epilogue
  // keeps four (measured slower with eight: 1.23 -> 1.32 ms at the maf_mnist shape; its 25-parameter spline
  // needs the registers the larger CTA takes away).
  // WIDE is chosen per launch: RQS never, the others when a CTA gets more than one tile (a 64-tile hidden
  // Linear at batch 4096 ran 2 us slower with the larger CTA).
  static constexpr int kEpiWarps = WIDE ? 8 : 4;
  static constexpr int kThreads = 64 + 2 * 32 * kEpiWarps;
  static constexpr int kStages = EPI == 2 ? 3 : 4;
  static constexpr int kScratch = EPI == 2 ? 2 * kDtTileScratchFloats : 0;
  static constexpr size_t kSmem = (size_t)kStages * kDtStageBytes + kScratch * 4 + 2 * kDtBN * 4 + (2 * kStages + 4) * 8 + 16;
};

// Work items of a persistent CTA.  A full item is one 128 x 256 tile; a half item (`part` = 0 / 1) is the 128 x 128 half
// of a tile whose B operand is the eight 2 KB runs [rows 128*part, +128) of the tile's weight image (MMA with N = 128).
//   * NH (hidden epilogue): EVERY tile is split — a launch with fewer tiles than SMs (the 512-wide hidden Linears at
//     batch 4096: 64 tiles, each a latency-bound chain of 16 K stages) becomes twice as many CTAs of half the MMA time;
//   * TS (RQS epilogue): only the tiles of the last, PARTIAL wave are split — 1248 tiles on 148 CTAs are 8 full waves
//     plus 64 tiles; unsplit, 84 CTAs idle through the ninth tile and the kernel ends with a whole tile's epilogue;
//     split, every CTA ends with at most half a tile (MMAs 9 -> 8.5 tile times, final epilogue halved).
struct DtItem {
  int mt, nt, part;  // part < 0: the whole tile
};
// Tile order.  Coupling Linears: m-tile major (concurrent CTAs spread over the n tiles; n-tile major order put 128 CTAs
// on the same weight stages at once and the issuer's operand waits grew from 14 K to 21 K cycles on the deep16 stack).
// MADE Linears (`desc`): n-tile major, DESCENDING — the masks leave n tile 0 the fewest K stages, so the expensive tiles
// go first and the tail of the walk is cheap (slowest issuer of the affine Linear 85 K -> 80 K cycles, tools/dt_trace.py).
template <bool NH, bool TS>
__device__ __forceinline__ bool dt_item(int w, int MT, int NT, bool desc, DtItem& it) {
  const int grid = (int)gridDim.x, cta = (int)blockIdx.x;
  if (NH) {
    const int t = cta + w * grid;
    if (t >= MT * NT * 2) return false;
    int c;
    if (desc) {
      c = t / MT;
      it.mt = t - c * MT;
      c = 2 * NT - 1 - c;
    } else {
      it.mt = t / (2 * NT);
      c = t - it.mt * 2 * NT;
    }
    it.nt = c >> 1;
    it.part = c & 1;
    return true;
  }
  const int ntiles = MT * NT;
  int t = cta + w * grid;
  it.part = -1;
  if (TS) {
    const int nfull = ntiles / grid, rem = ntiles - nfull * grid;
    if (w > nfull) return false;
    if (w == nfull) {
      if (rem == 0) return false;
      if (2 * rem <= grid) {  // split the tail: half `cta & 1` of tail tile `cta >> 1`
        if (cta >= 2 * rem) return false;
        t = nfull * grid + (cta >> 1);
        it.part = cta & 1;
      } else if (cta >= rem) {
        return false;
      }
    }
  } else if (t >= ntiles) {
    return false;
  }
  if (desc) {
    it.nt = t / MT;
    it.mt = t - it.nt * MT;
    it.nt = NT - 1 - it.nt;
  } else {
    it.mt = t / NT;
    it.nt = t - it.mt * NT;
  }
  return true;
}

// Persistent kernel: one CTA per SM walks its work items; a TMA ring feeds the MMA warp across item boundaries, and the
// two 256-column TMEM accumulators alternate between two epilogue warpgroups, so an epilogue may take twice the MMA time
// of a tile before it stalls the tensor pipe.
template <int EPI, bool INV, bool WIDE, bool NH = false>
__global__ void __launch_bounds__(DtCfg<EPI, WIDE>::kThreads, 1) dense_tc_gemm_kernel(DtGemm a, int MT, int NT) {
  using Cfg = DtCfg<EPI, WIDE>;
  static_assert(!NH || (EPI == kEpiHidden && !WIDE), "n-half tiles exist for the hidden epilogue with 4-warp groups");
  constexpr bool TS = EPI != kEpiHidden;  // transform Linears: split the tiles of the partial last wave
  constexpr int kDtStages = Cfg::kStages;
  constexpr int EW = Cfg::kEpiWarps;
  // Warp roles: the epilogue warps come FIRST (warp w reads TMEM lane quarter w % 4), the TMA producer and the MMA issuer
  // LAST.  The schedulers pick the highest eligible warp id first: with the issuer at warp 1 every one of its ~30
  // instructions per MMA queued behind four busy epilogue warps of its scheduler, and the tensor pipe idled between
  // MMAs (36-41 % active on the affine Linears with the issuer never waiting on a barrier).
  constexpr int kLoadWarp = 2 * EW, kMmaWarp = 2 * EW + 1;
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* stages = smem;                                                          // [stages][48 KB]
  float* Tbase = reinterpret_cast<float*>(smem + kDtStages * kDtStageBytes);       // epilogue scratch
  float* sbias2 = Tbase + Cfg::kScratch;                                    // [2][256]
  uint64_t* full = reinterpret_cast<uint64_t*>(sbias2 + 2 * kDtBN);                // [stages]
  uint64_t* empty = full + kDtStages;                                              // [stages]
  uint64_t* acc_full = empty + kDtStages;                                          // [2]
  uint64_t* acc_empty = acc_full + 2;                                              // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
#ifdef FLOWCUDA_TRACE
  const long long tr_entry = clock64();
#endif

  if (warp == kMmaWarp) tc::tmem_alloc(tmem_slot, 2 * kDtBN);
  if (tid == kLoadWarp * 32) {
    for (int s = 0; s < kDtStages; ++s) {
      tc::mbar_init(&full[s], 1);
      tc::mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(&acc_full[s], 1);
      tc::mbar_init(&acc_empty[s], 2 * EW);  // one arrival per epilogue warp: both groups drain every accumulator
    }
    tc::fence_mbar_init();
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  // everything above touched only this CTA's shared memory and TMEM: from here on the kernel reads what the previous
  // kernel of the chain wrote (activation images, the state v) and writes what it may still be reading
  pdl_wait();
  pdl_launch_dependents();

  if (warp == kLoadWarp) {
    if (lane == 0) {  // ===== TMA loader
      uint32_t it = 0;
      DtItem wi;
#ifdef FLOWCUDA_TRACE
      long long tr_wait = 0;
#endif
      for (int w = 0; dt_item<NH, TS>(w, MT, NT, a.made != 0, wi); ++w) {
        const int mt = wi.mt, nt = wi.nt;
        int KS = a.KS;
        if (nt < kDtMaxKsTiles && a.ks_tile[nt] != 0) KS = a.ks_tile[nt];
        const uint8_t* ag = a.a_img + (int64_t)mt * a.KS * 2 * kDtABytes;
        const uint8_t* bg = a.b_img + (int64_t)nt * a.KS * 2 * kDtBBytes;
        for (int ks = 0; ks < KS; ++ks, ++it) {
          const uint32_t st = it % kDtStages, ph = (it / kDtStages) & 1u;
          DT_TRACE_T0(t0);
          tc::mbar_wait(&empty[st], ph ^ 1u);  // passes at once on a fresh barrier
          DT_TRACE_ADD(tr_wait, t0);
          uint8_t* dst = stages + st * kDtStageBytes;
          if ((NH || TS) && wi.part >= 0) {
            tc::mbar_arrive_expect_tx(&full[st], 4 * kDtABytes);
            tc::tma_bulk_g2s(dst, ag + (int64_t)ks * 2 * kDtABytes, 2 * kDtABytes, &full[st]);
            // rows [128 part, +128) of each of the 2 x 4 K chunks: 2 KB runs, 4 KB apart in the image
            const uint8_t* bs = bg + (int64_t)ks * 2 * kDtBBytes + wi.part * 2048;
#pragma unroll
            for (int c = 0; c < 8; ++c) tc::tma_bulk_g2s(dst + 2 * kDtABytes + c * 2048, bs + c * 4096, 2048, &full[st]);
          } else {
            tc::mbar_arrive_expect_tx(&full[st], kDtStageBytes);
            tc::tma_bulk_g2s(dst, ag + (int64_t)ks * 2 * kDtABytes, 2 * kDtABytes, &full[st]);
            tc::tma_bulk_g2s(dst + 2 * kDtABytes, bg + (int64_t)ks * 2 * kDtBBytes, 2 * kDtBBytes, &full[st]);
          }
        }
      }
#ifdef FLOWCUDA_TRACE
      (void)tr_wait;
#endif
    }
  } else if (warp == kMmaWarp) {
    {  // ===== MMA issuer: 3 passes x 2 K steps per stage.  The WHOLE warp runs this loop, converged; the MMAs and commits
       // are issued by its elected lane (tc::mma_ss_warp).  Inside `if (lane == 0)` the compiler wraps every tcgen05.mma —
       // an instruction with uniform-register operands — in an ELECT / BRA.U.ANY loop and cannot keep the descriptors in
       // uniform registers across it: the single thread then needs ~95 cycles per MMA (tools/mma_timing.py), more than
       // the 64 tensor cycles of a 128-column MMA.
      constexpr uint32_t lbo_a = 16u * 128u;
      uint32_t it = 0;
      DtItem wi;
#ifdef FLOWCUDA_TRACE
      long long tr_acc = 0, tr_full = 0, tr_items = 0;
      const long long tr_start = clock64();
#endif
      for (int w = 0; dt_item<NH, TS>(w, MT, NT, a.made != 0, wi); ++w) {
        const int nt = wi.nt;
        const bool halfn = (NH || TS) && wi.part >= 0;
        // MMA width: the half / the whole tile, narrowed to the columns a ragged last tile of a transform Linear needs
        int ncols = kDtBN;
        if (EPI != kEpiHidden) {
          ncols = dt_tile_cols(a, nt);
          if (halfn) ncols = wi.part == 0 ? min(ncols, kDtBN / 2) : max(ncols - kDtBN / 2, 16);
        } else if (halfn) {
          ncols = kDtBN / 2;
        }
        const uint32_t idesc = tc::make_idesc(tc::kFmtBF16, 128, ncols);
        const uint32_t lbo_b = halfn ? 16u * 128u : 32u * 128u;       // rows / 8 * 128 of the staged B image
        const uint32_t b_lo = halfn ? (uint32_t)kDtABytes : (uint32_t)kDtBBytes;  // offset of its lo part
        int KS = a.KS;
        if (nt < kDtMaxKsTiles && a.ks_tile[nt] != 0) KS = a.ks_tile[nt];
        const uint32_t slot = (uint32_t)w & 1u, aph = ((uint32_t)w >> 1) & 1u;
        DT_TRACE_T0(t0);
        tc::mbar_wait(&acc_empty[slot], aph ^ 1u);  // the epilogue has drained this accumulator
        DT_TRACE_ADD(tr_acc, t0);
#ifdef FLOWCUDA_TRACE
        ++tr_items;
#endif
        tc::tc_fence_after();
        const uint32_t dtm = tmem + slot * (uint32_t)kDtBN;
        uint32_t acc = 0;
        for (int ks = 0; ks < KS; ++ks, ++it) {
          const uint32_t st = it % kDtStages, ph = (it / kDtStages) & 1u;
          DT_TRACE_T0(t1);
          tc::mbar_wait(&full[st], ph);
          DT_TRACE_ADD(tr_full, t1);
          tc::tc_fence_after();
          const uint32_t sa = tc::smem_u32(stages + st * kDtStageBytes), sb = sa + 2 * kDtABytes;
          // ONE descriptor pair per stage; the six MMAs add their operand offsets (>> 4) to the start-address field —
          // rebuilding both descriptors per MMA was ~15 dependent uniform-datapath instructions, more than the 64 tensor
          // cycles of a 128-column MMA leave (the addresses stay below 2^18, the field cannot carry)
          const uint64_t da = tc::make_smem_desc(sa, lbo_a, 128u), db = tc::make_smem_desc(sb, lbo_b, 128u);
#pragma unroll
          for (int pass = 0; pass < 3; ++pass) {
            const uint32_t ao = pass == 2 ? (uint32_t)kDtABytes : 0u, bo = pass == 1 ? b_lo : 0u;
#pragma unroll
            for (int s = 0; s < kDtBK / 16; ++s) {
              tc::mma_ss_warp<tc::kFmtBF16>(dtm, da + ((ao + s * 2u * lbo_a) >> 4), db + ((bo + s * 2u * lbo_b) >> 4), idesc, acc);
              acc = 1;
            }
          }
          tc::mma_commit_warp(&empty[st]);
        }
        tc::mma_commit_warp(&acc_full[slot]);
      }
#ifdef FLOWCUDA_TRACE
      if (g_dt_trace && lane == 0) {
        g_dt_trace[(NH ? 148 * 8 : 0) + blockIdx.x * 8 + 0] = (unsigned long long)(clock64() - tr_start);
        g_dt_trace[(NH ? 148 * 8 : 0) + blockIdx.x * 8 + 3] = (unsigned long long)(tr_start - tr_entry);  // kernel entry -> issuer loop
        g_dt_trace[(NH ? 148 * 8 : 0) + blockIdx.x * 8 + 1] = (unsigned long long)tr_acc;
        g_dt_trace[(NH ? 148 * 8 : 0) + blockIdx.x * 8 + 2] = (unsigned long long)tr_full;
        g_dt_trace[(NH ? 148 * 8 : 0) + blockIdx.x * 8 + 6] = (unsigned long long)tr_items;
        g_dt_trace[(NH ? 148 * 8 : 0) + blockIdx.x * 8 + 7] = (unsigned long long)it;
      }
#endif
    }
  } else {
    // ===== epilogue warps.  BOTH groups work on EVERY item, each on its half of the accumulator's columns: with the
    // two accumulators alternating between items, the MMAs of item i+1 then run under the epilogue of item i and a
    // group never waits for "its" accumulator to be refilled — per-item time max(MMA, epilogue/2).  (Round 1 gave
    // whole items to alternating groups; a group's accumulator could only be refilled after the group had drained
    // it, so the per-item time was (MMA + epilogue)/2: tensor pipe 76 % at best on the RQS Linear.)
    // Inside a group of eight warps, two warps share a TMEM lane quarter and split the group's columns again.
    const uint32_t group = (uint32_t)warp / EW;
    const int sub = (warp >> 2) & 1;  // EW == 8 only: which half of the group's column range
    const int et = tid % (EW * 32);
    auto group_sync = [&]() {  // this epilogue group only
      if (group == 0) asm volatile("bar.sync 1, %0;" ::"n"(EW * 32) : "memory");
      else asm volatile("bar.sync 2, %0;" ::"n"(EW * 32) : "memory");
    };
    float* sbias = sbias2 + group * kDtBN;
    float* Tw = Tbase + ((int)group * EW + warp % EW) * 32 * 36;  // this warp's transposition scratch (affine)
    DtItem wi;
#ifdef FLOWCUDA_TRACE
    long long tr_wait = 0;
#endif
    for (int w = 0; dt_item<NH, TS>(w, MT, NT, a.made != 0, wi); ++w) {
      const int mt = wi.mt, nt = wi.nt;
      const uint32_t slot = (uint32_t)w & 1u, aph = ((uint32_t)w >> 1) & 1u;
      const bool halfitem = (NH || TS) && wi.part >= 0;
      const int cshift = halfitem ? wi.part * (kDtBN / 2) : 0;
      // this warp's columns of the tile: the group's half of what the accumulator holds, halved again by `sub` in
      // eight-warp groups (64-column quarters: the granularity of the log-det partial slots); of a half item's 128
      // columns each group takes 64, which only the `sub` = 0 warps of an eight-warp group work on
      const bool split8 = EW == 8 && !halfitem;
      const bool colwork = EW == 4 || !halfitem || sub == 0;
      const int span = halfitem ? 64 : (split8 ? 64 : 128);
      const int c_lo = cshift + ((int)group * (split8 ? 2 : 1) + (split8 ? sub : 0)) * span, c_hi = c_lo + span;
      // RQS: the groups share a whole tile's 10 dims (dt_rqs_group_dims); a half item's 5 dims go 3 + 2
      int rd_lo = 0, rd_hi = 0;
      if (dt_is_rqs(EPI)) {
        if (halfitem) {
          rd_lo = 5 * wi.part + (group == 0 ? 0 : 3);
          rd_hi = 5 * wi.part + (group == 0 ? 3 : 5);
        } else {
          dt_rqs_group_dims(w, (int)group, rd_lo, rd_hi);
        }
      }
      for (int n = et; n < kDtBN; n += EW * 32) {  // bias of this tile's columns
        float b = 0.f;
        if (EPI == kEpiHidden) {
          const int col = nt * kDtBN + n;
          if (col < a.n_out) b = a.bias[col];
        } else {
          b = dt_tile_bias(a, nt, n);
          // RQS: the 16 bin logits of a dimension reach the spline in log2 units (rqs_apply_mufu)
          if (dt_is_rqs(EPI) && (n & 127) % 25 < 16) b *= kLog2e;
        }
        sbias[n] = b;
      }
      group_sync();
      DtRqsX xin;
      if (dt_is_rqs(EPI)) dt_rqs_load_x(a, mt, nt, rd_lo, rd_hi, warp, lane, xin);
      if (EPI == kEpiAffine) {
        // the bijector reads this warp's slice of the state v (it left L2 while the Linears streamed their
        // images): pull the thread's row of the slice into L2 now, under the tile's MMAs
        const int64_t prow = (int64_t)mt * 128 + (warp & 3) * 32 + lane;
        const int dlo = nt * a.dpt + c_lo / a.ppd, nd = min(span / a.ppd, a.ndims - dlo);
        if (prow < a.N && nd > 0) {
          const float* p0 = a.v + prow * a.D + a.tp + a.dstride * dlo;
          const uintptr_t q1 = (uintptr_t)(p0 + a.dstride * (nd - 1));
          for (uintptr_t q = (uintptr_t)p0 & ~(uintptr_t)127; q <= q1; q += 128)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(q));
        }
      }
      DT_TRACE_T0(t0);
      tc::mbar_wait(&acc_full[slot], aph);
      DT_TRACE_ADD(tr_wait, t0);
      tc::tc_fence_after();
      const uint32_t tb = tmem + slot * (uint32_t)kDtBN + ((uint32_t)((warp & 3) * 32) << 16);
      if (dt_is_rqs(EPI)) {
        // log-det slots: a whole tile's groups use slots 0 / 2, a half item's groups the two slots of its half
        dt_epilogue_rqs_tile<INV, EPI == kEpiRqs8Store>(a, mt, nt, rd_lo, rd_hi, halfitem ? 2 * wi.part + (int)group : 2 * (int)group, cshift, tb, sbias,
                                  warp, lane, xin);
      } else if (colwork) {
        dt_epilogue<EPI, INV>(a, mt, nt, tb, sbias, Tw, warp, lane, c_lo, c_hi, cshift);
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&acc_empty[slot]);
      // every warp of the group must be done with sbias before the group's next item overwrites it
      group_sync();
    }
#ifdef FLOWCUDA_TRACE
    if (g_dt_trace && lane == 0 && warp == 0) g_dt_trace[(NH ? 148 * 8 : 0) + blockIdx.x * 8 + 4] = (unsigned long long)tr_wait;
#endif
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) {
    __syncwarp();
    tc::tmem_dealloc(tmem, 2 * kDtBN);
#ifdef FLOWCUDA_TRACE
    if (g_dt_trace && lane == 0) g_dt_trace[(NH ? 148 * 8 : 0) + blockIdx.x * 8 + 5] = (unsigned long long)(clock64() - tr_entry);  // whole CTA
#endif
  }
}

template <int EPI, bool INV, bool WIDE>
static int dt_launch_gemm_w(const DtGemm& g, int64_t MT, int NT, cudaStream_t s) {
  auto kern = dense_tc_gemm_kernel<EPI, INV, WIDE>;
  constexpr size_t kSmem = DtCfg<EPI, WIDE>::kSmem;
  FC_ENSURE_SMEM(kern, kSmem);
  const int64_t tiles = MT * NT;
  const int grid = (int)(tiles < sm_count() ? tiles : sm_count());
  FC_CUDA(launch_chain(kern, dim3(grid), dim3(DtCfg<EPI, WIDE>::kThreads), kSmem, s, g, (int)MT, NT));
  return FLOWCUDA_OK;
}

template <int EPI, bool INV>
static int dt_launch_gemm_nh(const DtGemm& g, int64_t MT, int NT, cudaStream_t s) {
  auto kern = dense_tc_gemm_kernel<EPI, INV, false, true>;
  constexpr size_t kSmem = DtCfg<EPI, false>::kSmem;
  FC_ENSURE_SMEM(kern, kSmem);
  const int64_t tiles = MT * NT * 2;
  const int grid = (int)(tiles < sm_count() ? tiles : sm_count());
  FC_CUDA(launch_chain(kern, dim3(grid), dim3(DtCfg<EPI, false>::kThreads), kSmem, s, g, (int)MT, NT));
  return FLOWCUDA_OK;
}

template <int EPI, bool INV>
static int dt_launch_gemm(const DtGemm& g, int64_t MT, int NT, cudaStream_t s) {
  if (EPI == kEpiHidden) {
    // fewer tiles than SMs: split every n tile in two (128 x 128 CTA tiles) when that fills more of the device
    if (MT * NT < sm_count() * 3 / 4) return dt_launch_gemm_nh<kEpiHidden, false>(g, MT, NT, s);
  }
  if (!dt_is_rqs(EPI) && MT * NT > sm_count()) return dt_launch_gemm_w<EPI, INV, true>(g, MT, NT, s);
  return dt_launch_gemm_w<EPI, INV, false>(g, MT, NT, s);
}

// ---- log-det bookkeeping (fixed summation order: deterministic) ----------------------------------
// Every transform GEMM leaves two log-det partials per (row, tile) in ITS layer's slab of ldp
// ([layers processed][N][tiles]); one kernel at the end of the chain sums a row's `nslabs * used` partials in a
// fixed order (lane-strided, then the shuffle tree), adds the base log-density and writes the outputs.
static __global__ void __launch_bounds__(256) dense_tc_finish_kernel(const float* __restrict__ v, const float* __restrict__ ldp,
                                                                    int nslabs, int tiles, int used, int64_t N, int D,
                                                                    int with_base, int base, float* __restrict__ out_lp,
                                                                    float* __restrict__ out_ld) {
  pdl_wait();
  int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;  // one warp per row
  int lane = threadIdx.x & 31;
  if (r >= N) return;
  float ld = 0.f;
  for (int sl = 0; sl < nslabs; ++sl) {
    const float* p = ldp + ((int64_t)sl * N + r) * tiles;
    float s = 0.f;
    for (int t = lane; t < used; t += 32) s += p[t];
    ld += warp_sum(s);
  }
  float s = 0.f;
  if (with_base)
    for (int d = lane; d < D; d += 32) s += base_term(base, v[r * D + d]);
  s = warp_sum(s);
  if (lane == 0) {
    if (out_ld) out_ld[r] = ld;
    if (out_lp) out_lp[r] = ld + s;
  }
}

static inline size_t dt_al256(size_t x) { return (x + 255) & ~(size_t)255; }

}  // namespace flowcuda
