// coupling_dense_tc.cu — rows #4/#6 (and the C5 deep stacks) on the 5th-gen tensor cores
// (FLOWCUDA_PRECISION_BF16X3): distrax.MaskedCoupling stacks with a wide MLP conditioner
// (src/maf_mnist.py:34-89: 784 -> 512 -> 512 -> 512 -> 784*25) as tcgen05.mma kind::f16 GEMMs.
//
// Every Linear is C[128 x 256 tile] = A · W with bf16 operands in error-compensated 3-pass form
// (hi·hi + hi·lo + lo·hi, fp32 accumulate in TMEM; single-pass bf16 misses the 1e-3 log_prob
// budget by 7x — profiles/r01/precision_study.txt).  Operands live in HBM as ready-made
// shared-memory tile images (tc.cuh), so a pipeline stage is two cp.async.bulk (TMA) copies:
//   A image  [m_tile][k_stage][hi|lo][4 chunks][128 rows][16 B]   (activations, 32 K per stage)
//   B image  [n_tile][k_stage][hi|lo][4 chunks][256 rows][16 B]   (weights, packed per call;
//            masks folded in: the first Linear keeps only the visible dims' rows, the last only
//            the transformed dims' columns, 10 dims x 25 spline params (+6 pad) per 256 tile)
// CTA = loader warp + MMA warp + 4 epilogue warps, 2 CTAs per SM so that one CTA's epilogue
// overlaps the other's MMAs.  Epilogues (thread = sample row = TMEM lane):
//   hidden: +bias, relu, bf16 hi/lo split, written straight into the next Linear's A image;
//   last:   the inner bijector (RQS / affine) on the tile's dims, state updated in place,
//           log-det partial per (row, tile) — the [N, 19600] conditioner output never exists.
#include "tc.cuh"
#include "transforms.cuh"

namespace flowcuda {

constexpr int kDtBN = 256, kDtBK = 32, kDtStages = 2, kDtThreads = 192;
constexpr int kDtABytes = 128 * kDtBK * 2;   // one of hi/lo of an A stage (8 KB)
constexpr int kDtBBytes = kDtBN * kDtBK * 2; // one of hi/lo of a B stage (16 KB)
constexpr int kDtStageBytes = 2 * kDtABytes + 2 * kDtBBytes;  // 48 KB

// ---- operand packing ------------------------------------------------------------------------
// v[N, D] fp32 (flow state) -> A image of the visible dims d = vis + 2j (j < nvis), K padded to KS*32.
__global__ void __launch_bounds__(256) dense_tc_pack_x_kernel(const float* __restrict__ v, int64_t N, int D, int vis,
                                                             int nvis, int KS, uint8_t* __restrict__ img) {
  const int64_t total = (int64_t)((N + 127) / 128) * 128 * (KS * 4);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = i % ((N + 127) / 128 * 128);
    const int ck = (int)(i / ((N + 127) / 128 * 128));  // global chunk index along K (8 elements each)
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int j = ck * 8 + e;
      f[e] = (row < N && j < nvis) ? v[row * D + vis + 2 * j] : 0.f;
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

struct DtPackW {
  const float* params;   // flow parameters, all layers
  int64_t per_layer;     // floats per flow layer
  int64_t w_off;         // offset of this Linear's W inside a layer
  int k_in, n_out;       // W[k_in, n_out] row-major
  int kstride, kvalid;   // image k index j <-> W row koff(l) + kstride*j, j < kvalid
  int first;             // 1: first Linear (koff = visible parity of the layer), else koff = 0
  int dpt, ppd, D;       // last Linear: dims per tile, params per dim (dpt = 0: plain columns); D = event dim
  int KS, NT;            // k stages, n tiles
  uint8_t* img;          // [layer][NT][KS][hi|lo][16 KB]
};

// grid (blocks, NT, layers)
__global__ void __launch_bounds__(256) dense_tc_pack_w_kernel(DtPackW a) {
  const int l = blockIdx.z, nt = blockIdx.y;
  const float* W = a.params + a.per_layer * l + a.w_off;
  const int vis = (l & 1) ? 0 : 1;        // visible parity of layer l (mask True at odd dims on even layers)
  const int tp = 1 - vis;                 // parity of the transformed dims
  const int koff = a.first ? vis : 0;
  const int ndims = (a.D - tp + 1) / 2;                       // transformed dims of this layer
  const int kvalid = a.first ? (a.D - vis + 1) / 2 : a.kvalid;  // visible dims feed the first Linear
  uint8_t* base = a.img + ((int64_t)l * a.NT + nt) * a.KS * 2 * (int64_t)kDtBBytes;
  const int total = a.KS * 4 * kDtBN;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int n = i % kDtBN, ck = i / kDtBN;
    int col = -1;
    if (a.dpt == 0) {
      col = nt * kDtBN + n;
      if (col >= a.n_out) col = -1;
    } else {
      const int di = n / a.ppd, pp = n - di * a.ppd;
      const int dim = nt * a.dpt + di;
      if (di < a.dpt && dim < ndims) col = (tp + 2 * dim) * a.ppd + pp;
    }
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int j = ck * 8 + e;
      f[e] = (col >= 0 && j < kvalid) ? W[(int64_t)(koff + a.kstride * j) * a.n_out + col] : 0.f;
    }
    uint32_t h[4], lo[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) tc::split_bf16x2(f[2 * e], f[2 * e + 1], h[e], lo[e]);
    uint8_t* st = base + (int64_t)(ck >> 2) * 2 * kDtBBytes + tc::img_off(kDtBN, n, ck & 3);
    *reinterpret_cast<uint4*>(st) = make_uint4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<uint4*>(st + kDtBBytes) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
  }
}

// ---- the GEMM ---------------------------------------------------------------------------------
struct DtGemm {
  const uint8_t* a_img;  // [m_tiles][KS][hi|lo][8 KB]
  const uint8_t* b_img;  // [n_tiles][KS][hi|lo][16 KB]
  const float* bias;     // this Linear's bias (inside the flow parameters)
  int KS;
  int n_out;             // valid output columns (hidden epilogue) / conditioner width (transform epilogue)
  uint8_t* out_img;      // hidden epilogue: next Linear's A image, out_KS stages
  int out_KS;
  float* v;              // transform epilogue: state [N, D], updated in place
  float* ldp;            // [N, tiles] log-det partials
  int64_t N;
  int D, tp, ndims, dpt, ppd, tiles;
  float lo, hi;
};

constexpr int kEpiHidden = 0, kEpiRqs8 = 1, kEpiAffine = 2;

template <int EPI, bool INV>
__global__ void __launch_bounds__(kDtThreads, 2) dense_tc_gemm_kernel(DtGemm a) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* stages = smem;                                                     // [2][48 KB]
  float* sbias = reinterpret_cast<float*>(smem + kDtStages * kDtStageBytes);  // [256]
  uint64_t* full = reinterpret_cast<uint64_t*>(sbias + kDtBN);               // [2]
  uint64_t* empty = full + kDtStages;                                         // [2]
  uint64_t* dfull = empty + kDtStages;                                        // [1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(dfull + 1);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int mt = blockIdx.x, nt = blockIdx.y;
  const int KS = a.KS;

  if (warp == 1) tc::tmem_alloc(tmem_slot, kDtBN);
  if (tid == 0) {
    for (int s = 0; s < kDtStages; ++s) {
      tc::mbar_init(&full[s], 1);
      tc::mbar_init(&empty[s], 1);
    }
    tc::mbar_init(dfull, 1);
    tc::fence_mbar_init();
  }
  if (warp >= 2) {  // bias of this tile's columns
    const int c = tid - 64;
    for (int n = c; n < kDtBN; n += 128) {
      float b = 0.f;
      if (EPI == kEpiHidden) {
        const int col = nt * kDtBN + n;
        if (col < a.n_out) b = a.bias[col];
      } else {
        const int di = n / a.ppd, pp = n - di * a.ppd, dim = nt * a.dpt + di;
        if (di < a.dpt && dim < a.ndims) b = a.bias[(a.tp + 2 * dim) * a.ppd + pp];
      }
      sbias[n] = b;
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {  // ===== TMA loader
      const uint8_t* ag = a.a_img + (int64_t)mt * KS * 2 * kDtABytes;
      const uint8_t* bg = a.b_img + (int64_t)nt * KS * 2 * kDtBBytes;
      for (int ks = 0; ks < KS; ++ks) {
        const int st = ks % kDtStages, use = ks / kDtStages;
        if (use > 0) tc::mbar_wait(&empty[st], (uint32_t)((use - 1) & 1));
        uint8_t* dst = stages + st * kDtStageBytes;
        tc::mbar_arrive_expect_tx(&full[st], kDtStageBytes);
        tc::tma_bulk_g2s(dst, ag + (int64_t)ks * 2 * kDtABytes, 2 * kDtABytes, &full[st]);
        tc::tma_bulk_g2s(dst + 2 * kDtABytes, bg + (int64_t)ks * 2 * kDtBBytes, 2 * kDtBBytes, &full[st]);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {  // ===== MMA issuer: 3 passes x 2 K steps per stage
      constexpr uint32_t idesc = tc::make_idesc(tc::kFmtBF16, 128, kDtBN);
      constexpr uint32_t lbo_a = 16u * 128u, lbo_b = (uint32_t)(kDtBN >> 3) * 128u;
      uint32_t acc = 0;
      for (int ks = 0; ks < KS; ++ks) {
        const int st = ks % kDtStages, use = ks / kDtStages;
        tc::mbar_wait(&full[st], (uint32_t)(use & 1));
        tc::tc_fence_after();
        const uint32_t sa = tc::smem_u32(stages + st * kDtStageBytes), sb = sa + 2 * kDtABytes;
#pragma unroll
        for (int pass = 0; pass < 3; ++pass) {
          const uint32_t ab = sa + (pass == 2 ? kDtABytes : 0), bb = sb + (pass == 1 ? kDtBBytes : 0);
#pragma unroll
          for (int s = 0; s < kDtBK / 16; ++s) {
            tc::mma_ss<tc::kFmtBF16>(tmem, tc::make_smem_desc(ab + s * 2u * lbo_a, lbo_a, 128u),
                                     tc::make_smem_desc(bb + s * 2u * lbo_b, lbo_b, 128u), idesc, acc);
            acc = 1;
          }
        }
        tc::mma_commit(&empty[st]);
      }
      tc::mma_commit(dfull);
    }
  } else {
    // ===== epilogue warps: thread = row = TMEM lane
    const int row = (warp & 3) * 32 + lane;
    const int64_t grow = (int64_t)mt * 128 + row;
    const uint32_t tb = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    tc::mbar_wait(dfull, 0);
    tc::tc_fence_after();
    if (EPI == kEpiHidden) {
      uint8_t* og = a.out_img + (int64_t)mt * a.out_KS * 2 * kDtABytes;
#pragma unroll 1
      for (int c = 0; c < kDtBN; c += 16) {
        uint32_t vv[16];
        tc::tmem_ld_x16(tb + (uint32_t)c, vv);
        tc::tmem_ld_wait();
        uint32_t h[8], l[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float f0 = fmaxf(__uint_as_float(vv[2 * e]) + sbias[c + 2 * e], 0.f);
          const float f1 = fmaxf(__uint_as_float(vv[2 * e + 1]) + sbias[c + 2 * e + 1], 0.f);
          tc::split_bf16x2(f0, f1, h[e], l[e]);
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
    } else if (EPI == kEpiRqs8) {
      float ldsum = 0.f;
#pragma unroll 1
      for (int di = 0; di < 10; ++di) {
        const int dim = nt * 10 + di;
        if (dim >= a.ndims) break;  // uniform across the CTA
        uint32_t vv[25];
        tc::tmem_ld_x16(tb + (uint32_t)(25 * di), vv);
        tc::tmem_ld_x8(tb + (uint32_t)(25 * di + 16), vv + 16);
        tc::tmem_ld_x1(tb + (uint32_t)(25 * di + 24), vv + 24);
        tc::tmem_ld_wait();
        if (grow < a.N) {
          float o[25];
#pragma unroll
          for (int i = 0; i < 25; ++i) o[i] = __uint_as_float(vv[i]) + sbias[25 * di + i];
          const int d = a.tp + 2 * dim;
          const float xin = a.v[grow * a.D + d];
          float nv, l1;
          rqs_apply<INV, 8>([&](int i) { return o[i]; }, 8, a.lo, a.hi, xin, nv, l1);
          a.v[grow * a.D + d] = nv;
          ldsum += l1;
        }
      }
      if (grow < a.N) a.ldp[grow * a.tiles + nt] = ldsum;
    } else {
      float ldsum = 0.f;
#pragma unroll 1
      for (int c = 0; c < kDtBN; c += 32) {
        if (nt * 128 + c / 2 >= a.ndims) break;
        uint32_t vv[32];
        tc::tmem_ld_x32(tb + (uint32_t)c, vv);
        tc::tmem_ld_wait();
        if (grow < a.N) {
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            const int dim = nt * 128 + c / 2 + e;
            if (dim < a.ndims) {
              const int d = a.tp + 2 * dim;
              float nv, l1;
              affine_apply<INV>(__uint_as_float(vv[2 * e]) + sbias[c + 2 * e],
                                __uint_as_float(vv[2 * e + 1]) + sbias[c + 2 * e + 1], a.v[grow * a.D + d], nv, l1);
              a.v[grow * a.D + d] = nv;
              ldsum += l1;
            }
          }
        }
      }
      if (grow < a.N) a.ldp[grow * a.tiles + nt] = ldsum;
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tc::tmem_dealloc(tmem, kDtBN);
  }
}

// ---- log-det bookkeeping (fixed summation order: deterministic) ----------------------------------
__global__ void __launch_bounds__(256) dense_tc_ld_accum_kernel(const float* __restrict__ ldp, int tiles, int used,
                                                               float* __restrict__ ld, int64_t N, int first) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= N) return;
  float s = first ? 0.f : ld[r];
  for (int t = 0; t < used; ++t) s += ldp[r * tiles + t];
  ld[r] = s;
}

__global__ void __launch_bounds__(256) dense_tc_finish_kernel(const float* __restrict__ v, const float* __restrict__ ld,
                                                             int64_t N, int D, int with_base, int base,
                                                             float* __restrict__ out_lp, float* __restrict__ out_ld) {
  int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;  // one warp per row
  int lane = threadIdx.x & 31;
  if (r >= N) return;
  float s = 0.f;
  if (with_base)
    for (int d = lane; d < D; d += 32) s += base_term(base, v[r * D + d]);
  s = warp_sum(s);
  if (lane == 0) {
    if (out_ld) out_ld[r] = ld[r];
    if (out_lp) out_lp[r] = ld[r] + s;
  }
}

// ---- host side ---------------------------------------------------------------------------------------
struct DtLayout {
  int nlin, sizes[FLOWCUDA_MAX_HIDDEN_LAYERS + 2];
  int64_t w_off[FLOWCUDA_MAX_HIDDEN_LAYERS + 1], b_off[FLOWCUDA_MAX_HIDDEN_LAYERS + 1], per_layer;
  int ppd, dpt, tiles, out_full;
  int KS[FLOWCUDA_MAX_HIDDEN_LAYERS + 1], NT[FLOWCUDA_MAX_HIDDEN_LAYERS + 1];
  int64_t MT;
  size_t off_v, off_ld, off_ldp, off_x, off_h[2], off_w[FLOWCUDA_MAX_HIDDEN_LAYERS + 1], total;
};

static size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }

static bool dt_layout(const flowcuda_coupling_spec* sp, int64_t n, DtLayout& L) {
  const int D = sp->event_dim, nh = sp->num_hidden;
  if (sp->transform == FLOWCUDA_TRANSFORM_RQS) {
    if (sp->num_bins != 8) return false;
    L.ppd = 25; L.dpt = 10;
  } else if (sp->transform == FLOWCUDA_TRANSFORM_AFFINE) {
    L.ppd = 2; L.dpt = 128;
  } else {
    return false;
  }
  if (nh < 1) return false;
  L.nlin = nh + 1;
  L.out_full = D * L.ppd;
  L.sizes[0] = D;
  for (int i = 0; i < nh; ++i) L.sizes[i + 1] = sp->hidden[i];
  L.sizes[nh + 1] = L.out_full;
  int64_t o = 0;
  for (int i = 0; i < L.nlin; ++i) {
    L.w_off[i] = o; o += (int64_t)L.sizes[i] * L.sizes[i + 1];
    L.b_off[i] = o; o += L.sizes[i + 1];
  }
  L.per_layer = o;
  const int ndmax = (D + 1) / 2;
  L.tiles = (ndmax + L.dpt - 1) / L.dpt;
  L.MT = (n + 127) / 128;
  for (int i = 0; i < L.nlin; ++i) {
    const int kin = i == 0 ? ndmax : L.sizes[i];
    L.KS[i] = (kin + kDtBK - 1) / kDtBK;
    L.NT[i] = i < nh ? (L.sizes[i + 1] + kDtBN - 1) / kDtBN : L.tiles;
  }
  size_t off = 0;
  L.off_v = off; off = al256(off + (size_t)n * D * 4);
  L.off_ld = off; off = al256(off + (size_t)n * 4);
  L.off_ldp = off; off = al256(off + (size_t)n * L.tiles * 4);
  L.off_x = off; off = al256(off + (size_t)L.MT * L.KS[0] * 2 * kDtABytes);
  int ksh = 1;
  for (int i = 1; i < L.nlin; ++i) ksh = L.KS[i] > ksh ? L.KS[i] : ksh;
  for (int j = 0; j < 2; ++j) { L.off_h[j] = off; off = al256(off + (size_t)L.MT * ksh * 2 * kDtABytes); }
  for (int i = 0; i < L.nlin; ++i) {
    L.off_w[i] = off;
    off = al256(off + (size_t)sp->num_layers * L.NT[i] * L.KS[i] * 2 * kDtBBytes);
  }
  L.total = off + 256;
  return true;
}

bool coupling_dense_tc_supported(const flowcuda_coupling_spec* sp) {
  DtLayout L;
  return sp->event_dim >= 4 && dt_layout(sp, 128, L);
}

size_t coupling_dense_tc_workspace_bytes(const flowcuda_coupling_spec* sp, int64_t n) {
  DtLayout L;
  if (!dt_layout(sp, n, L)) return 0;
  return L.total;
}

template <int EPI, bool INV>
static int dt_launch_gemm(const DtGemm& g, int64_t MT, int NT, cudaStream_t s) {
  const size_t smem = (size_t)kDtStages * kDtStageBytes + kDtBN * 4 + 5 * 8 + 16;
  auto kern = dense_tc_gemm_kernel<EPI, INV>;
  FC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<dim3((unsigned)MT, (unsigned)NT), kDtThreads, smem, s>>>(g);
  FC_LAUNCH_CHECK("dense_tc_gemm_kernel");
  return FLOWCUDA_OK;
}

int coupling_dense_tc_run(const flowcuda_coupling_spec* sp, const float* params, const float* in, float* out_v,
                          float* out_lp, float* out_ld, int64_t n, int layer_begin, int layer_end, int with_base,
                          bool inverse, void* workspace, size_t workspace_bytes, cudaStream_t s) {
  DtLayout L;
  if (!dt_layout(sp, n, L)) {
    set_error("coupling(bf16x3): needs RQS with 8 bins or per-dimension affine and >= 1 hidden layer");
    return FLOWCUDA_EUNSUPPORTED;
  }
  if (!workspace || workspace_bytes < L.total) {
    set_error("coupling(bf16x3): workspace %zu < %zu bytes", workspace_bytes, L.total);
    return FLOWCUDA_EWORKSPACE;
  }
  uint8_t* ws = (uint8_t*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  const int D = sp->event_dim, nh = sp->num_hidden;
  float* v = (float*)(ws + L.off_v);
  float* ld = (float*)(ws + L.off_ld);
  float* ldp = (float*)(ws + L.off_ldp);
  FC_CUDA(cudaMemcpyAsync(v, in, (size_t)n * D * sizeof(float), cudaMemcpyDeviceToDevice, s));
  // pack every Linear of every layer (masks folded in)
  for (int i = 0; i < L.nlin; ++i) {
    DtPackW p;
    p.params = params; p.per_layer = L.per_layer; p.w_off = L.w_off[i];
    p.k_in = L.sizes[i]; p.n_out = L.sizes[i + 1];
    p.first = i == 0; p.kstride = i == 0 ? 2 : 1; p.kvalid = L.sizes[i];
    p.dpt = i == nh ? L.dpt : 0; p.ppd = L.ppd; p.D = D;
    p.KS = L.KS[i]; p.NT = L.NT[i]; p.img = ws + L.off_w[i];
    int blocks = (L.KS[i] * 4 * kDtBN + 255) / 256;
    dense_tc_pack_w_kernel<<<dim3(blocks, L.NT[i], sp->num_layers), 256, 0, s>>>(p);
    FC_LAUNCH_CHECK("dense_tc_pack_w_kernel");
  }
  const int nl = layer_end - layer_begin;
  for (int it = 0; it < nl; ++it) {
    const int l = inverse ? layer_begin + it : layer_end - 1 - it;
    const int vis = (l & 1) ? 0 : 1, tp = 1 - vis;
    const int ndims = (D - tp + 1) / 2, nvis = (D - vis + 1) / 2;
    const float* lp = params + L.per_layer * l;
    {
      int64_t total = L.MT * 128 * L.KS[0] * 4;
      int blocks = (int)((total + 255) / 256);
      dense_tc_pack_x_kernel<<<blocks, 256, 0, s>>>(v, n, D, vis, nvis, L.KS[0], ws + L.off_x);
      FC_LAUNCH_CHECK("dense_tc_pack_x_kernel");
    }
    const uint8_t* aimg = ws + L.off_x;
    for (int i = 0; i < L.nlin; ++i) {
      DtGemm g;
      g.a_img = aimg;
      g.b_img = ws + L.off_w[i] + (size_t)l * L.NT[i] * L.KS[i] * 2 * kDtBBytes;
      g.bias = lp + L.b_off[i];
      g.KS = L.KS[i];
      g.n_out = L.sizes[i + 1];
      g.out_img = nullptr; g.out_KS = 0;
      g.v = v; g.ldp = ldp; g.N = n; g.D = D; g.tp = tp; g.ndims = ndims; g.dpt = L.dpt; g.ppd = L.ppd;
      g.tiles = L.tiles; g.lo = sp->range_min; g.hi = sp->range_max;
      int rc;
      if (i < nh) {
        g.out_img = ws + L.off_h[i & 1];
        g.out_KS = L.KS[i + 1];
        rc = dt_launch_gemm<kEpiHidden, false>(g, L.MT, L.NT[i], s);
        aimg = g.out_img;
      } else {
        const int used = (ndims + L.dpt - 1) / L.dpt;
        if (sp->transform == FLOWCUDA_TRANSFORM_RQS)
          rc = inverse ? dt_launch_gemm<kEpiRqs8, true>(g, L.MT, used, s) : dt_launch_gemm<kEpiRqs8, false>(g, L.MT, used, s);
        else
          rc = inverse ? dt_launch_gemm<kEpiAffine, true>(g, L.MT, used, s) : dt_launch_gemm<kEpiAffine, false>(g, L.MT, used, s);
        if (rc) return rc;
        dense_tc_ld_accum_kernel<<<(int)((n + 255) / 256), 256, 0, s>>>(ldp, L.tiles, used, ld, n, it == 0);
        FC_LAUNCH_CHECK("dense_tc_ld_accum_kernel");
      }
      if (rc) return rc;
    }
  }
  if (nl == 0) FC_CUDA(cudaMemsetAsync(ld, 0, (size_t)n * sizeof(float), s));
  if (out_v) FC_CUDA(cudaMemcpyAsync(out_v, v, (size_t)n * D * sizeof(float), cudaMemcpyDeviceToDevice, s));
  if (out_lp || out_ld) {
    dense_tc_finish_kernel<<<(int)((n * 32 + 255) / 256), 256, 0, s>>>(v, ld, n, D, with_base, sp->base, out_lp, out_ld);
    FC_LAUNCH_CHECK("dense_tc_finish_kernel");
  }
  return FLOWCUDA_OK;
}

}  // namespace flowcuda
