// mgtc.cuh — tensor-core engine for the reverse-mode GEMMs (made_grad.cu, coupling_dense_grad.cu).
//
// Runs the SAME problem descriptor as the FP32 engine (MgGemm, mg_sgemm.cuh) — C = op(A)·op(B) with strided
// operands, bias / relu / relu-gate / accumulate epilogue, MADE masks, deterministic split-K — on tcgen05
// with bf16 operands in error-compensated 3-pass form (hi·hi + hi·lo + lo·hi, fp32 accumulate in TMEM), the
// arithmetic of the density path (dense_tc.cuh).  Per product:
//   1. mgtc_pack_kernel (x2): op(A) -> [m_tile][k_stage][hi|lo] 128-row tile images, op(B) -> 256-row tile
//      images with the MADE mask folded in.  Strides are generic, so Xᵀ·dY and dY·Wᵀ need no transposes.
//   2. mgtc_gemm_kernel: persistent (one CTA per SM), TMA-bulk ring of 48 KB stages, two 256-column TMEM
//      accumulators alternating between two epilogue warpgroups; a work item is (split, m tile, n tile) and
//      walks only the K stages its split and its mask need (a masked Linear is block-triangular).
//   3. split-K (weight gradients: K = batch, few output tiles): raw partials, then mg_splitk_reduce_kernel
//      adds them in ascending order and applies the mask — bit-reproducible.
#pragma once
#include "dense_tc.cuh"
#include "mg_sgemm.cuh"

namespace flowcuda {

constexpr int kMgTcSplits = 16;
constexpr int kMgTcStages = 4;
constexpr int kMgTcThreads = 320;  // loader warp, MMA warp, two epilogue groups of four warps
constexpr size_t kMgTcSmem = (size_t)kMgTcStages * kDtStageBytes + 2 * kDtBN * 4 + (2 * kMgTcStages + 4) * 8 + 16;

// image row r (tiles of TR rows), K index k  <-  src[r*rs + k*cs]  (r < R, k < K; zero padding)
// mask_mode 1: the element is W[in = k, out = r]; 2: W[in = r, out = k]; 0: none.
// KC (cs == 1, K contiguous in memory): four lanes take the four 8-element chunks of one row's K stage, so a
// warp reads eight full 128-byte lines and writes four 128-byte runs of the image.  Otherwise (rows contiguous:
// a transposed source) consecutive lanes take consecutive rows: every load and the store are coalesced.
template <int TR, bool KC>
static __global__ void __launch_bounds__(256) mgtc_pack_kernel(const float* __restrict__ src, int64_t rs, int64_t cs,
                                                              int R, int64_t K, int KS, int tiles, MgMask mask,
                                                              int mask_mode, MgColMap map, uint8_t* __restrict__ img) {
  constexpr int kHalf = TR * kDtBK * 2;  // bytes of one of hi/lo of a stage
  const int64_t rows = (int64_t)tiles * TR;
  const int64_t total = rows * (KS * 4);
  const bool vec = KC && map.mode != 2 && (rs & 3) == 0 && ((uintptr_t)src & 15) == 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t row;
    int ck;  // chunk of 8 along K
    if (KC) {
      const int64_t t = i >> 2;
      row = t % rows;
      ck = (int)(t / rows) * 4 + (int)(i & 3);
    } else {
      row = i % rows;
      ck = (int)(i / rows);
    }
    float f[8];
    const float* p = src + (map.mode == 1 && row < R ? (int64_t)mg_full_col(map, (int)row) : row) * rs;
    const int64_t k0 = (int64_t)ck * 8;
    if (vec && row < R && k0 + 8 <= K) {
      const float4 lo4 = *reinterpret_cast<const float4*>(p + k0), hi4 = *reinterpret_cast<const float4*>(p + k0 + 4);
      f[0] = lo4.x; f[1] = lo4.y; f[2] = lo4.z; f[3] = lo4.w; f[4] = hi4.x; f[5] = hi4.y; f[6] = hi4.z; f[7] = hi4.w;
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int64_t k = k0 + e;
        f[e] = (row < R && k < K) ? p[(map.mode == 2 ? (int64_t)mg_full_col(map, (int)k) : k) * cs] : 0.f;
      }
    }
    if (mask_mode != 0 && row < R) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int64_t k = k0 + e;
        if (k < K) {
          if (mask_mode == 1 && !mg_mask(mask, (int)k, (int)row)) f[e] = 0.f;
          if (mask_mode == 2 && !mg_mask(mask, (int)row, (int)k)) f[e] = 0.f;
        }
      }
    }
    uint32_t h[4], l[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) tc::split_bf16x2(f[2 * e], f[2 * e + 1], h[e], l[e]);
    const int64_t t = row / TR;
    const int ks = ck >> 2, c = ck & 3;
    uint8_t* st = img + ((t * KS + ks) * 2) * (int64_t)kHalf + tc::img_off(TR, (int)(row % TR), c);
    *reinterpret_cast<uint4*>(st) = make_uint4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<uint4*>(st + kHalf) = make_uint4(l[0], l[1], l[2], l[3]);
  }
}

struct MgTcArgs {
  const uint8_t* a_img;
  const uint8_t* b_img;
  int KS, MT, NT, S, ks_chunk;
  MgGemm g;
};

// K stages [ks0, ks1) of work item (z, nt): the split's slice intersected with what the mask leaves.
__device__ __forceinline__ void mgtc_krange(const MgTcArgs& a, int z, int mt, int nt, int& ks0, int& ks1) {
  ks0 = z * a.ks_chunk;
  ks1 = min(a.KS, ks0 + a.ks_chunk);
  const MgGemm& g = a.g;
  if (!g.mask.on) return;
  const int j0 = nt * kDtBN, jl = min(j0 + kDtBN, g.N) - 1;
  if (g.mask_on_b == 1) {  // B = W[k (in), j (out)]: inputs with degree above the tile's largest output degree never count
    const int dmax = mg_dout_max(g.mask, j0, jl);
    int64_t lo = 0, hi = g.K;  // first k with din(k) > dmax
    while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (mg_din(g.mask, (int)mid) > dmax) hi = mid; else lo = mid + 1; }
    ks1 = min(ks1, (int)((lo + kDtBK - 1) / kDtBK));
    if (ks1 <= ks0) ks1 = ks0 + 1;  // keep one (all-zero) stage: the accumulator must be written
  } else if (g.mask_on_b == 2 && !g.mask.wrap) {  // B = Wᵀ (in = j, out = k): outputs below the tile's smallest input degree
    const int dmin = mg_din(g.mask, j0);
    int64_t lo = 0, hi = g.K;  // first k with dout(k) >= dmin
    while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (g.mask.dout[mid] >= dmin) hi = mid; else lo = mid + 1; }
    ks0 = max(ks0, min((int)(lo / kDtBK), ks1 - 1));
  } else if (g.mask_on_b == 0 && a.S > 1) {  // weight gradient: a tile the mask empties is never reduced
    if (mg_din(g.mask, mt * 128) > mg_dout_max(g.mask, j0, jl)) ks1 = ks0;
  }
}

static __global__ void __launch_bounds__(kMgTcThreads, 1) mgtc_gemm_kernel(MgTcArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* stages = smem;
  float* sbias = reinterpret_cast<float*>(smem + kMgTcStages * kDtStageBytes);  // [2][256]
  uint64_t* full = reinterpret_cast<uint64_t*>(sbias + 2 * kDtBN);
  uint64_t* empty = full + kMgTcStages;
  uint64_t* acc_full = empty + kMgTcStages;
  uint64_t* acc_empty = acc_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int per_split = a.MT * a.NT, nitems = per_split * a.S;

  if (warp == 1) tc::tmem_alloc(tmem_slot, 2 * kDtBN);
  if (tid == 0) {
    for (int s = 0; s < kMgTcStages; ++s) {
      tc::mbar_init(&full[s], 1);
      tc::mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(&acc_full[s], 1);
      tc::mbar_init(&acc_empty[s], 4);
    }
    tc::fence_mbar_init();
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {  // ===== TMA loader
      uint32_t it = 0;
      for (int t = blockIdx.x; t < nitems; t += gridDim.x) {
        const int z = t / per_split, r = t - z * per_split, mt = r / a.NT, nt = r - mt * a.NT;
        int ks0, ks1;
        mgtc_krange(a, z, mt, nt, ks0, ks1);
        const uint8_t* ag = a.a_img + (int64_t)mt * a.KS * 2 * kDtABytes;
        const uint8_t* bg = a.b_img + (int64_t)nt * a.KS * 2 * kDtBBytes;
        for (int ks = ks0; ks < ks1; ++ks, ++it) {
          const uint32_t st = it % kMgTcStages, ph = (it / kMgTcStages) & 1u;
          tc::mbar_wait(&empty[st], ph ^ 1u);
          uint8_t* dst = stages + st * kDtStageBytes;
          tc::mbar_arrive_expect_tx(&full[st], kDtStageBytes);
          tc::tma_bulk_g2s(dst, ag + (int64_t)ks * 2 * kDtABytes, 2 * kDtABytes, &full[st]);
          tc::tma_bulk_g2s(dst + 2 * kDtABytes, bg + (int64_t)ks * 2 * kDtBBytes, 2 * kDtBBytes, &full[st]);
        }
      }
    }
  } else if (warp == 1) {
    {  // ===== MMA issuer: the whole warp runs the loop converged, its elected lane issues (dense_tc.cuh)
      constexpr uint32_t idesc = tc::make_idesc(tc::kFmtBF16, 128, kDtBN);
      constexpr uint32_t lbo_a = 16u * 128u, lbo_b = (uint32_t)(kDtBN >> 3) * 128u;
      uint32_t it = 0, tl = 0;
      for (int t = blockIdx.x; t < nitems; t += gridDim.x) {
        const int z = t / per_split, r = t - z * per_split, mt = r / a.NT, nt = r - mt * a.NT;
        int ks0, ks1;
        mgtc_krange(a, z, mt, nt, ks0, ks1);
        if (ks1 <= ks0) continue;
        const uint32_t slot = tl & 1u, aph = (tl >> 1) & 1u;
        ++tl;
        tc::mbar_wait(&acc_empty[slot], aph ^ 1u);
        tc::tc_fence_after();
        const uint32_t dtm = tmem + slot * (uint32_t)kDtBN;
        uint32_t acc = 0;
        for (int ks = ks0; ks < ks1; ++ks, ++it) {
          const uint32_t st = it % kMgTcStages, ph = (it / kMgTcStages) & 1u;
          tc::mbar_wait(&full[st], ph);
          tc::tc_fence_after();
          const uint32_t sa = tc::smem_u32(stages + st * kDtStageBytes), sb = sa + 2 * kDtABytes;
          const uint64_t da = tc::make_smem_desc(sa, lbo_a, 128u), db = tc::make_smem_desc(sb, lbo_b, 128u);  // per stage; see dense_tc.cuh
#pragma unroll
          for (int pass = 0; pass < 3; ++pass) {
            const uint32_t ao = pass == 2 ? (uint32_t)kDtABytes : 0u, bo = pass == 1 ? (uint32_t)kDtBBytes : 0u;
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
    }
  } else {
    // ===== epilogue warps: group 0 (warps 2-5) takes this CTA's even items, group 1 (warps 6-9) the odd ones
    const uint32_t group = (uint32_t)(warp - 2) >> 2;
    const MgGemm& g = a.g;
    const bool vec = (((uintptr_t)g.C | (uintptr_t)g.part | (uintptr_t)g.gate) & 15) == 0 && (g.ldc & 3) == 0 &&
                     (g.ldg & 3) == 0 && (g.N & 3) == 0 && !(g.map.mode == 3 && a.S == 1);
    const bool cmap = g.map.mode == 3 && a.S == 1;
    uint32_t tl = 0;
    for (int t = blockIdx.x; t < nitems; t += gridDim.x) {
      const int z = t / per_split, r = t - z * per_split, mt = r / a.NT, nt = r - mt * a.NT;
      int ks0, ks1;
      mgtc_krange(a, z, mt, nt, ks0, ks1);
      if (ks1 <= ks0) continue;
      const uint32_t mine = tl & 1u, aph = (tl >> 1) & 1u;
      ++tl;
      if (mine != group) continue;
      const uint32_t slot = group;
      tc::mbar_wait(&acc_full[slot], aph);
      tc::tc_fence_after();
      const uint32_t tb = tmem + slot * (uint32_t)kDtBN + ((uint32_t)((warp & 3) * 32) << 16);
      const int row = mt * 128 + (warp & 3) * 32 + lane;
      const bool rok = row < g.M;
      const int jend = min(kDtBN, g.N - nt * kDtBN);  // valid columns of this tile (uniform)
      float* crow = a.S > 1 ? g.part + ((int64_t)z * g.M + row) * g.N : g.C + (int64_t)row * g.ldc;
      const float* grow = g.gate ? g.gate + (int64_t)row * g.ldg : nullptr;
      const bool has_bias = a.S == 1 && g.bias != nullptr;
      float* sb = sbias + group * kDtBN;
      if (has_bias) {  // this tile's bias columns, remapped once
        for (int n = (tid - 64) & 127; n < kDtBN; n += 128) {
          const int col = nt * kDtBN + n;
          sb[n] = col < g.N ? g.bias[g.map.mode == 1 ? mg_full_col(g.map, col) : col] : 0.f;
        }
        if (group == 0) asm volatile("bar.sync 1, 128;" ::: "memory");
        else asm volatile("bar.sync 2, 128;" ::: "memory");
      }
      const bool accum = a.S == 1 && g.accumulate;
      // 16 columns of this thread's row: epilogue + store
      auto emit = [&](const uint32_t* vv, int c) {
        const int col = nt * kDtBN + c;
        float o[16];
#pragma unroll
        for (int e = 0; e < 16; ++e) o[e] = __uint_as_float(vv[e]);
        if (a.S == 1) {
          if (has_bias) {
#pragma unroll
            for (int e = 0; e < 16; ++e) o[e] += sb[c + e];
          }
          if (g.relu) {
#pragma unroll
            for (int e = 0; e < 16; ++e) o[e] = fmaxf(o[e], 0.f);
          }
          if (grow) {
            if (vec && c + 16 <= jend) {
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const float4 gv = *reinterpret_cast<const float4*>(grow + col + 4 * q);
                if (!(gv.x > 0.f)) o[4 * q] = 0.f;
                if (!(gv.y > 0.f)) o[4 * q + 1] = 0.f;
                if (!(gv.z > 0.f)) o[4 * q + 2] = 0.f;
                if (!(gv.w > 0.f)) o[4 * q + 3] = 0.f;
              }
            } else {
#pragma unroll
              for (int e = 0; e < 16; ++e)
                if (col + e < g.N && !(grow[col + e] > 0.f)) o[e] = 0.f;
            }
          }
          if (g.mask.on && g.mask_on_b == 0) {
#pragma unroll
            for (int e = 0; e < 16; ++e)
              if (col + e < g.N && !mg_mask(g.mask, row, col + e)) o[e] = 0.f;
          }
        }
        if (vec && c + 16 <= jend) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float4* dst = reinterpret_cast<float4*>(crow + col + 4 * q);
            float4 w = make_float4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
            if (accum) {
              const float4 old = *dst;
              w.x += old.x; w.y += old.y; w.z += old.z; w.w += old.w;
            }
            *dst = w;
          }
        } else {
#pragma unroll
          for (int e = 0; e < 16; ++e)
            if (col + e < g.N) {
              float* dst = crow + (cmap ? mg_full_col(g.map, col + e) : col + e);
              *dst = accum ? *dst + o[e] : o[e];
            }
        }
      };
#pragma unroll 1
      for (int c = 0; c < jend; c += 32) {
        uint32_t v0[16], v1[16];
        const bool two = c + 16 < jend;  // uniform
        tc::tmem_ld_x16(tb + (uint32_t)c, v0);
        if (two) tc::tmem_ld_x16(tb + (uint32_t)(c + 16), v1);
        tc::tmem_ld_wait();
        if (rok) {
          emit(v0, c);
          if (two) emit(v1, c + 16);
        }
      }
      if (has_bias) {  // everyone is done with sb before the group's next tile rewrites it
        if (group == 0) asm volatile("bar.sync 1, 128;" ::: "memory");
        else asm volatile("bar.sync 2, 128;" ::: "memory");
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&acc_empty[slot]);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tc::tmem_dealloc(tmem, 2 * kDtBN);
  }
}

// C = epilogue(sum_z part[z])   (z ascending: deterministic); the epilogue of the unsplit path: bias, relu,
// relu gate, MADE mask on C, accumulate.  Masked-out weight gradients never read their partials (their tiles
// were skipped).
static __global__ void __launch_bounds__(256) mgtc_reduce_kernel(MgGemm g) {
  const int64_t tot = (int64_t)g.M * g.N;
  // four consecutive columns of one row per thread (N % 4 == 0): 128-bit reads of the partials, one division per four
  // elements — the scalar form ran at 1-2 TB/s on the 20 MB partials of the last Linear's weight gradient
  if ((g.N & 3) == 0 && ((uintptr_t)g.part & 15) == 0) {
    const int64_t tot4 = tot >> 2;
    const int n4 = g.N >> 2;
    const bool cvec = g.map.mode != 3 && (g.ldc & 3) == 0 && ((uintptr_t)g.C & 15) == 0;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < tot4; q += (int64_t)gridDim.x * blockDim.x) {
      const int gi = (int)(q / n4), gj = (int)(q - (int64_t)gi * n4) * 4;
      float r[4] = {0.f, 0.f, 0.f, 0.f};
      const float4* pp = reinterpret_cast<const float4*>(g.part) + q;
      for (int z = 0; z < g.splits; ++z) {
        const float4 t = pp[(int64_t)z * tot4];
        r[0] += t.x; r[1] += t.y; r[2] += t.z; r[3] += t.w;
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int j = gj + e;
        if (g.mask.on && g.mask_on_b == 0 && !mg_mask(g.mask, gi, j)) { r[e] = 0.f; continue; }
        if (g.bias) r[e] += g.bias[g.map.mode == 1 ? mg_full_col(g.map, j) : j];
        if (g.relu) r[e] = fmaxf(r[e], 0.f);
        if (g.gate && !(g.gate[gi * g.ldg + j] > 0.f)) r[e] = 0.f;
      }
      if (cvec) {
        float4* c = reinterpret_cast<float4*>(g.C + gi * g.ldc + gj);
        float4 w = make_float4(r[0], r[1], r[2], r[3]);
        if (g.accumulate) { const float4 o = *c; w.x += o.x; w.y += o.y; w.z += o.z; w.w += o.w; }
        *c = w;
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float* c = g.C + gi * g.ldc + (g.map.mode == 3 ? mg_full_col(g.map, gj + e) : gj + e);
          *c = g.accumulate ? *c + r[e] : r[e];
        }
      }
    }
    return;
  }
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < tot; e += (int64_t)gridDim.x * blockDim.x) {
    const int gi = (int)(e / g.N), gj = (int)(e - (int64_t)gi * g.N);
    float r = 0.f;
    if (!(g.mask.on && g.mask_on_b == 0) || mg_mask(g.mask, gi, gj)) {
      for (int z = 0; z < g.splits; ++z) r += g.part[(int64_t)z * tot + e];
      if (g.bias) r += g.bias[g.map.mode == 1 ? mg_full_col(g.map, gj) : gj];
      if (g.relu) r = fmaxf(r, 0.f);
      if (g.gate && !(g.gate[gi * g.ldg + gj] > 0.f)) r = 0.f;
    }
    float* c = g.C + gi * g.ldc + (g.map.mode == 3 ? mg_full_col(g.map, gj) : gj);
    *c = g.accumulate ? *c + r : r;
  }
}

struct MgTcScratch {
  uint8_t* a_img; size_t a_bytes;
  uint8_t* b_img; size_t b_bytes;
  float* part; size_t part_floats;
};

static inline size_t mgtc_a_bytes(int64_t M, int64_t K) {
  return (size_t)((M + 127) / 128) * (size_t)((K + kDtBK - 1) / kDtBK) * 2 * kDtABytes;
}
static inline size_t mgtc_b_bytes(int64_t N, int64_t K) {
  return (size_t)((N + kDtBN - 1) / kDtBN) * (size_t)((K + kDtBK - 1) / kDtBK) * 2 * kDtBBytes;
}

// Operand images some other kernel has already written in the engine's format (NULL: pack op(A) / op(B) here).
struct MgTcPre {
  const uint8_t* a_img;
  const uint8_t* b_img;
};

// Same contract as mg_launch; the engine itself decides whether and how far to split K.
static int mgtc_launch(MgGemm g, const MgTcScratch& sc, cudaStream_t s, MgTcPre pre = MgTcPre{nullptr, nullptr}) {
  MgTcArgs a = {};
  a.MT = (g.M + 127) / 128;
  a.NT = (g.N + kDtBN - 1) / kDtBN;
  a.KS = (int)((g.K + kDtBK - 1) / kDtBK);
  if (a.KS < 1) a.KS = 1;
  a.S = 1;
  {
    // Split K when that shortens the schedule: cost(S) = waves(S) x stages per item + the pass over the
    // partials, in units of one K stage (~0.65 us per CTA measured; the reduce streams (S+1) x M x N floats
    // at ~3 TB/s).  Weight gradients (K = batch, few output tiles), the input gradient of a wide last Linear
    // (K = its width) and tile counts just above the SM count (a two-wave tail) are the cases that split.
    const int sms = sm_count(), items = a.MT * a.NT;
    int smax = a.KS / 16;  // a split keeps at least 16 stages
    if (smax > kMgTcSplits) smax = kMgTcSplits;
    double best = 0.0;
    for (int S = 1; S <= (smax < 1 ? 1 : smax); ++S) {
      if (S > 1 && (size_t)S * g.M * g.N > sc.part_floats) break;
      const int chunk = (a.KS + S - 1) / S;
      const int Se = (a.KS + chunk - 1) / chunk;
      double cost = (double)(((int64_t)items * Se + sms - 1) / sms) * chunk;
      if (Se > 1) cost += (double)(Se + 1) * g.M * g.N * 4.0 / 3.0e12 / 0.65e-6;
      if (S == 1 || cost < best) { best = cost; a.S = Se; }
    }
  }
  a.ks_chunk = (a.KS + a.S - 1) / a.S;
  a.S = (a.KS + a.ks_chunk - 1) / a.ks_chunk;
  const size_t ab = mgtc_a_bytes(g.M, g.K), bb = mgtc_b_bytes(g.N, g.K);
  if ((!pre.a_img && ab > sc.a_bytes) || (!pre.b_img && bb > sc.b_bytes)) {
    set_error("mgtc_launch: operand images (%zu, %zu bytes) exceed the scratch (%zu, %zu)", ab, bb, sc.a_bytes, sc.b_bytes);
    return FLOWCUDA_EWORKSPACE;
  }
  const MgMask none = {};
  if (!pre.a_img) {
    const int64_t total = (int64_t)a.MT * 128 * a.KS * 4;
    const int gr = grid_for(total, 256, 8);
    if (g.acs == 1)
      mgtc_pack_kernel<128, true><<<gr, 256, 0, s>>>(g.A, g.ars, g.acs, g.M, g.K, a.KS, a.MT, none, 0, MgColMap{}, sc.a_img);
    else
      mgtc_pack_kernel<128, false><<<gr, 256, 0, s>>>(g.A, g.ars, g.acs, g.M, g.K, a.KS, a.MT, none, 0, MgColMap{}, sc.a_img);
    FC_LAUNCH_CHECK("mgtc_pack_kernel<A>");
  }
  if (!pre.b_img) {
    const int64_t total = (int64_t)a.NT * kDtBN * a.KS * 4;
    // B(k, j) = B[k*brs + j*bcs]: image row j, K index k
    const int gr = grid_for(total, 256, 8);
    const MgColMap bm = g.map.mode == 3 ? MgColMap{} : g.map;
    const int mm = g.mask.on ? g.mask_on_b : 0;
    if (g.brs == 1)
      mgtc_pack_kernel<kDtBN, true><<<gr, 256, 0, s>>>(g.B, g.bcs, g.brs, g.N, g.K, a.KS, a.NT, g.mask, mm, bm, sc.b_img);
    else
      mgtc_pack_kernel<kDtBN, false><<<gr, 256, 0, s>>>(g.B, g.bcs, g.brs, g.N, g.K, a.KS, a.NT, g.mask, mm, bm, sc.b_img);
    FC_LAUNCH_CHECK("mgtc_pack_kernel<B>");
  }
  a.a_img = pre.a_img ? pre.a_img : sc.a_img;
  a.b_img = pre.b_img ? pre.b_img : sc.b_img;
  g.splits = a.S;
  g.part = a.S > 1 ? sc.part : nullptr;
  a.g = g;
  FC_ENSURE_SMEM(mgtc_gemm_kernel, kMgTcSmem);
  const int items = a.MT * a.NT * a.S;
  const int grid = items < sm_count() ? items : sm_count();
  mgtc_gemm_kernel<<<grid, kMgTcThreads, kMgTcSmem, s>>>(a);
  FC_LAUNCH_CHECK("mgtc_gemm_kernel");
  if (a.S > 1) {
    mgtc_reduce_kernel<<<grid_for(((int64_t)g.M * g.N + 3) / 4, 256, 8), 256, 0, s>>>(g);
    FC_LAUNCH_CHECK("mgtc_reduce_kernel");
  }
  return FLOWCUDA_OK;
}

}  // namespace flowcuda
