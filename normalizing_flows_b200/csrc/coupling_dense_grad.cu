// coupling_dense_grad.cu — row #7 for row #4 (and the C5 affine stacks): value_and_grad of
// loss = -mean(Transformed.log_prob(x)) for distrax.MaskedCoupling flows of any event size — the training
// step of src/maf_mnist.py:134-166 (conditioner Flatten -> MLP(512,512,512, relu, activate_final) ->
// Linear(784*25) -> Reshape, inner bijector RationalQuadraticSpline, maf_mnist.py:34-89).
// Hand-written FP32 reverse mode over stored activations; every product goes through the generic
// deterministic SGEMM (mg_sgemm.cuh), the spline adjoints through rqs_grad.cuh.
//   forward per flow layer (stored): v (layer input), h_1..h_nh, o = conditioner output [N, D*P]
//   backward: per (sample, transformed dim) the inner bijector's adjoint gives dL/dx and dL/do; dims that
//   pass through get dL/do = 0 and receive the conditioner's input gradient instead.
#include "mgtc.cuh"
#include "rqs_grad.cuh"

namespace flowcuda {

struct CgLayout {
  int D, nh, nlin, P, L, sizes[FLOWCUDA_MAX_HIDDEN_LAYERS + 2];
  int64_t w_off[FLOWCUDA_MAX_HIDDEN_LAYERS + 1], b_off[FLOWCUDA_MAX_HIDDEN_LAYERS + 1], per_layer;
  size_t act_h[FLOWCUDA_MAX_HIDDEN_LAYERS + 1];  // float offsets of h_i / o inside one layer's activation block
  size_t per_layer_act;
  size_t off_state, off_g, off_tmp, off_tmp2, off_ld, off_lp, off_dO, off_dh[2], off_sum, off_part, off_act, total;  // bytes
  int hmax;
  int tc;  // tensor-core engine (spec->precision == BF16X3)
  size_t off_aimg, off_bimg, a_bytes, b_bytes, part_floats;
  int img;  // tensor-core engine + spline transform: the bijector adjoint writes dO as operand images (no fp32 dO)
  // Tensor-core engine: the weight-gradient jobs (dW_i, db_i — nothing downstream of them inside the step) run on an
  // internal side stream next to the input-gradient chain.  They get their own operand/partial scratch, and the buffers
  // both streams touch are doubled by flow-layer parity (masked input, one dH per Linear).
  size_t off_mx2, off_dhx[2][FLOWCUDA_MAX_HIDDEN_LAYERS + 1], off_aimg2, off_bimg2, off_part2, a2_bytes, b2_bytes;
  int fwd_dt;  // ... and the forward pass runs on the density path's GEMMs with fused epilogues (coupling_dense_tc.cu)
  int64_t ldo_dt;
  size_t off_dt, dt_bytes;
  size_t off_dOa, off_dOb, off_colp, dOa_bytes, dOb_bytes;
};
constexpr int kCgColSlices = 64;  // row slices of cg_transform_bwd_img_kernel = partial column sums per column

static size_t cg_al(size_t x) { return (x + 255) & ~(size_t)255; }

static bool cg_layout(const flowcuda_coupling_spec* sp, int64_t n, CgLayout& L) {
  if (!sp) return false;
  if (sp->transform == FLOWCUDA_TRANSFORM_RQS) {
    if (sp->num_bins != 8 && sp->num_bins != 12) return false;
    L.P = 3 * sp->num_bins + 1;
  } else if (sp->transform == FLOWCUDA_TRANSFORM_AFFINE) {
    L.P = 2;
  } else {
    return false;
  }
  L.D = sp->event_dim; L.nh = sp->num_hidden; L.nlin = L.nh + 1; L.L = sp->num_layers;
  if (L.nh < 1 || L.nh > FLOWCUDA_MAX_HIDDEN_LAYERS) return false;
  L.sizes[0] = L.D;
  L.hmax = 1;
  for (int i = 0; i < L.nh; ++i) {
    L.sizes[i + 1] = sp->hidden[i];
    if (sp->hidden[i] > L.hmax) L.hmax = sp->hidden[i];
  }
  L.sizes[L.nh + 1] = L.D * L.P;
  int64_t o = 0;
  for (int i = 0; i < L.nlin; ++i) {
    L.w_off[i] = o; o += (int64_t)L.sizes[i] * L.sizes[i + 1];
    L.b_off[i] = o; o += L.sizes[i + 1];
  }
  L.per_layer = o;
  L.tc = sp->precision == FLOWCUDA_PRECISION_BF16X3;
  L.img = L.tc && sp->transform == FLOWCUDA_TRANSFORM_RQS;
  L.fwd_dt = L.img && sp->num_bins == 8 && coupling_dense_tc_supported(sp);
  L.ldo_dt = L.fwd_dt ? coupling_dense_tc_store_ldo(sp) : 0;
  auto al64 = [](size_t x) { return (x + 63) & ~(size_t)63; };  // 256-byte aligned blocks: the epilogues store 16-byte vectors
  size_t a = al64((size_t)n * L.D);  // v
  for (int i = 0; i < L.nlin; ++i) {
    size_t w = L.sizes[i + 1];
    if (i == L.nh && L.fwd_dt && (size_t)L.ldo_dt > w) w = L.ldo_dt;
    L.act_h[i] = a; a = al64(a + (size_t)n * w);
  }
  L.per_layer_act = a;
  size_t off = 0;
  L.off_state = off; off = cg_al(off + (size_t)n * L.D * 4);
  L.off_g = off; off = cg_al(off + (size_t)n * L.D * 4);
  L.off_tmp = off; off = cg_al(off + (size_t)n * L.D * 4);
  L.off_tmp2 = off; off = cg_al(off + (size_t)n * L.D * 4);
  L.off_ld = off; off = cg_al(off + (size_t)n * 4);
  L.off_lp = off; off = cg_al(off + (size_t)n * 4);
  L.off_dO = off;
  if (!L.img) off = cg_al(off + (size_t)n * L.D * L.P * 4);
  L.off_dOa = L.off_dOb = L.off_colp = off; L.dOa_bytes = L.dOb_bytes = 0;
  if (L.img) {
    const int64_t ncmax = (int64_t)((L.D + 1) / 2) * L.P;
    L.dOa_bytes = mgtc_a_bytes(n, ncmax);
    L.dOb_bytes = mgtc_b_bytes(ncmax, n);
    L.off_dOa = off; off = cg_al(off + L.dOa_bytes);
    L.off_dOb = off; off = cg_al(off + L.dOb_bytes);
    L.off_colp = off; off = cg_al(off + (size_t)kCgColSlices * ncmax * sizeof(double));
  }
  for (int j = 0; j < 2; ++j) { L.off_dh[j] = off; off = cg_al(off + (size_t)n * L.hmax * 4); }
  L.off_sum = off; off = cg_al(off + 1024 * 8);
  size_t mx = 0;
  for (int i = 0; i < L.nlin; ++i) {
    const size_t e = (size_t)L.sizes[i] * L.sizes[i + 1];
    if (e > mx) mx = e;
  }
  L.part_floats = mx * (L.tc ? kMgTcSplits : kMgSplits);
  {
    size_t wmax = 0;  // the column sums borrow the buffer: kMgColSlices doubles per column
    for (int i = 0; i <= L.nlin; ++i) if ((size_t)L.sizes[i] > wmax) wmax = L.sizes[i];
    if (L.part_floats < 2 * kMgColSlices * wmax) L.part_floats = 2 * kMgColSlices * wmax;
  }
  L.off_part = off; off = cg_al(off + L.part_floats * 4);
  L.off_aimg = L.off_bimg = off; L.a_bytes = L.b_bytes = 0;
  if (L.tc) {
    auto upd = [&](int64_t am, int64_t ak, int64_t bn, int64_t bk) {
      if (mgtc_a_bytes(am, ak) > L.a_bytes) L.a_bytes = mgtc_a_bytes(am, ak);
      if (mgtc_b_bytes(bn, bk) > L.b_bytes) L.b_bytes = mgtc_b_bytes(bn, bk);
    };
    for (int i = 0; i < L.nlin; ++i) {
      upd(n, L.sizes[i], L.sizes[i + 1], L.sizes[i]);      // forward
      upd(L.sizes[i], n, L.sizes[i + 1], n);               // weight gradient
      upd(n, L.sizes[i + 1], L.sizes[i], L.sizes[i + 1]);  // input gradient
    }
    L.off_aimg = off; off = cg_al(off + L.a_bytes);
    L.off_bimg = off; off = cg_al(off + L.b_bytes);
  }
  L.off_mx2 = L.off_tmp; L.off_aimg2 = L.off_aimg; L.off_bimg2 = L.off_bimg; L.off_part2 = L.off_part;
  L.a2_bytes = L.a_bytes; L.b2_bytes = L.b_bytes;
  for (int p = 0; p < 2; ++p)
    for (int i = 0; i <= L.nh; ++i) L.off_dhx[p][i] = L.off_dh[i & 1];
  if (L.tc) {
    L.off_mx2 = off; off = cg_al(off + (size_t)n * L.D * 4);
    for (int p = 0; p < 2; ++p)
      for (int i = 1; i <= L.nh; ++i) { L.off_dhx[p][i] = off; off = cg_al(off + (size_t)n * L.sizes[i] * 4); }
    L.a2_bytes = L.b2_bytes = 0;
    for (int i = 0; i < L.nlin; ++i) {  // the side stream only runs dW_i = in_iᵀ · dout_i (K = batch)
      const size_t ab = mgtc_a_bytes(L.sizes[i], n), bb = mgtc_b_bytes(L.sizes[i + 1], n);
      if (ab > L.a2_bytes) L.a2_bytes = ab;
      if (!(L.img && i == L.nh) && bb > L.b2_bytes) L.b2_bytes = bb;
    }
    L.off_aimg2 = off; off = cg_al(off + L.a2_bytes);
    L.off_bimg2 = off; off = cg_al(off + L.b2_bytes);
    L.off_part2 = off; off = cg_al(off + L.part_floats * 4);
  }
  L.off_dt = off; L.dt_bytes = 0;
  if (L.fwd_dt) { L.dt_bytes = coupling_dense_tc_workspace_bytes(sp, n) + 256; off = cg_al(off + L.dt_bytes); }
  L.off_act = off; off = cg_al(off + L.per_layer_act * 4 * (size_t)L.L);
  L.total = off + 256;
  return true;
}

// mx = v with the transformed dims (parity tp) zeroed: the conditioner's input
__global__ void __launch_bounds__(256) cg_mask_input_kernel(const float* __restrict__ v, float* __restrict__ mx, int64_t N, int D,
                                                           int tp) {
  const int64_t tot = N * D;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x) {
    const int d = (int)(i % D);
    mx[i] = ((d & 1) == tp) ? 0.f : v[i];
  }
}

// forward bijector on the transformed dims, one warp per row: out = y (pass-through dims copied),
// ld[r] (+)= sum of the log-dets in lane order, then a fixed shuffle tree (deterministic)
template <int K>
__global__ void __launch_bounds__(256) cg_transform_fwd_kernel(const float* __restrict__ v, const float* __restrict__ o,
                                                              float* __restrict__ out, float* __restrict__ ld, int64_t N, int D,
                                                              int P, int tp, float lo, float hi, int first, int64_t ldo) {
  const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (r >= N) return;
  float s = 0.f;
  if (ldo) {  // compact conditioner output: lanes walk the transformed dims only (no idle half)
    const int nd = (D - tp + 1) / 2;
    for (int k = lane; k < nd; k += 32) {
      const int d = tp + 2 * k, dp = d ^ 1;
      const float* oo = o + r * ldo + (int64_t)k * P;
      float y, l1;
      if (K > 0) {
        rqs_apply<false, K>([&](int i) { return oo[i]; }, K, lo, hi, v[r * D + d], y, l1);
      } else {
        affine_apply<false>(oo[0], oo[1], v[r * D + d], y, l1);
      }
      s += l1;
      out[r * D + d] = y;
      if (dp < D) out[r * D + dp] = v[r * D + dp];
      if (tp == 1 && (D & 1) && k == nd - 1) out[r * D + D - 1] = v[r * D + D - 1];
    }
    s = warp_sum(s);
    if (lane == 0) ld[r] = first ? s : ld[r] + s;
    return;
  }
  for (int d = lane; d < D; d += 32) {
    const float x = v[r * D + d];
    float y = x;
    if ((d & 1) == tp) {
      // ldo != 0: compact conditioner output, only the transformed dims' blocks, row stride ldo
      const float* oo = ldo ? o + r * ldo + (int64_t)(d >> 1) * P : o + (r * D + d) * (int64_t)P;
      float l1;
      if (K > 0) {
        rqs_apply<false, K>([&](int i) { return oo[i]; }, K, lo, hi, x, y, l1);
      } else {
        affine_apply<false>(oo[0], oo[1], x, y, l1);
      }
      s += l1;
    }
    out[r * D + d] = y;
  }
  s = warp_sum(s);
  if (lane == 0) ld[r] = first ? s : ld[r] + s;
}

// lp = base(z) + ld ; g = d(-mean lp)/dz
__global__ void __launch_bounds__(256) cg_finish_kernel(const float* __restrict__ z, const float* __restrict__ ld,
                                                       float* __restrict__ lp, float* __restrict__ g, int64_t N, int D, int base) {
  const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (r >= N) return;
  const float invn = 1.f / (float)N;
  float s = 0.f;
  for (int d = lane; d < D; d += 32) {
    const float zz = z[r * D + d];
    s += base_term(base, zz);
    g[r * D + d] = base == FLOWCUDA_BASE_NORMAL ? zz * invn : 0.f;
  }
  s = warp_sum(s);
  if (lane == 0) lp[r] = ld[r] + s;
}

// adjoint of the bijector: g (dL/d layer output) -> gin (dL/d layer input, bijector path only) and dO
template <int K>
__global__ void __launch_bounds__(128) cg_transform_bwd_kernel(const float* __restrict__ v, const float* __restrict__ o,
                                                              const float* __restrict__ g, float* __restrict__ gin,
                                                              float* __restrict__ dO, int64_t N, int D, int P, int tp, float lo,
                                                              float hi, int64_t ldo) {
  const int64_t tot = N * D;
  const float ldbar = -1.f / (float)N;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x) {
    const int d = (int)(i % D);
    const int64_t r = i / D;
    float* ob = ldo ? dO + r * ldo + (int64_t)(d >> 1) * P : dO + i * P;
    if ((d & 1) != tp) {
      gin[i] = g[i];
      if (!ldo)
        for (int p = 0; p < P; ++p) ob[p] = 0.f;
      continue;
    }
    const float* oo = ldo ? o + r * ldo + (int64_t)(d >> 1) * P : o + i * P;
    float y, l1, xbar;
    if (K > 0) {
      float ov[3 * (K > 0 ? K : 1) + 1], obar[3 * (K > 0 ? K : 1) + 1];
#pragma unroll
      for (int p = 0; p < 3 * K + 1; ++p) ov[p] = oo[p];
      rqs_fwd_bwd<(K > 0 ? K : 1)>(ov, lo, hi, v[i], g[i], ldbar, y, l1, xbar, obar);
#pragma unroll
      for (int p = 0; p < 3 * K + 1; ++p) ob[p] = obar[p];
    } else {
      float ov[2] = {oo[0], oo[1]}, obar[2];
      affine_fwd_bwd(ov, v[i], g[i], ldbar, y, l1, xbar, obar);
      ob[0] = obar[0];
      ob[1] = obar[1];
    }
    gin[i] = xbar;
  }
}

// The same adjoint on the COMPACT layout (tensor-core engine): item i = (sample r, k-th transformed dim) owns
// the contiguous block o[i*PP .. i*PP+PP), so a warp's 32 items are one contiguous run.  The run is staged
// through shared memory in both directions: global loads and stores are full 128-byte lines, each lane then
// reads / writes its PP parameters at stride PP (odd: bank-conflict free).  Every lane works (the full-layout
// kernel idles the pass-through half); it also forwards the gradient of its pass-through neighbour dim.
template <int K>
__global__ void __launch_bounds__(128) cg_transform_bwd_compact_kernel(const float* __restrict__ v, const float* __restrict__ o,
                                                                      const float* __restrict__ g, float* __restrict__ gin,
                                                                      float* __restrict__ dO, int64_t N, int D, int tp, float lo,
                                                                      float hi) {
  constexpr int PP = K > 0 ? 3 * K + 1 : 2;
  __shared__ __align__(16) float sm[4][32 * PP];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nd = (D - tp + 1) / 2;
  const int64_t items = N * nd, nwarps = (items + 31) / 32;
  const float ldbar = -1.f / (float)N;
  float* buf = sm[warp];
  for (int64_t w = (int64_t)blockIdx.x * 4 + warp; w < nwarps; w += (int64_t)gridDim.x * 4) {
    const int64_t i0 = w * 32;
    const int cnt = (int)min((int64_t)32, items - i0);
    const float* src = o + i0 * PP;
    for (int e = lane; e < cnt * PP; e += 32) buf[e] = src[e];
    __syncwarp();
    const int64_t i = i0 + lane;
    float obar[PP];
    if (lane < cnt) {
      const int64_t r = i / nd;
      const int k = (int)(i - r * nd), d = tp + 2 * k;
      const int64_t xi = r * D + d;
      float ov[PP];
#pragma unroll
      for (int p = 0; p < PP; ++p) ov[p] = buf[lane * PP + p];
      float y, l1, xbar;
      if (K > 0) rqs_fwd_bwd<(K > 0 ? K : 1)>(ov, lo, hi, v[xi], g[xi], ldbar, y, l1, xbar, obar);
      else affine_fwd_bwd(ov, v[xi], g[xi], ldbar, y, l1, xbar, obar);
      gin[xi] = xbar;
      const int dp = d ^ 1;  // the pass-through neighbour
      if (dp < D) gin[r * D + dp] = g[r * D + dp];
      if (tp == 1 && (D & 1) && k == nd - 1) gin[r * D + D - 1] = g[r * D + D - 1];  // odd D: the last dim has no pair
    }
    __syncwarp();
    if (lane < cnt) {
#pragma unroll
      for (int p = 0; p < PP; ++p) buf[lane * PP + p] = obar[p];
    }
    __syncwarp();
    float* dst = dO + i0 * PP;
    for (int e = lane; e < cnt * PP; e += 32) dst[e] = buf[e];
    __syncwarp();
  }
}

// The tensor-core training step's form of that adjoint (spline transforms): dO never exists as an fp32 matrix.
// The kernel writes it straight into the two operand images the reverse-mode GEMMs consume —
//   A image  [m tile of 128 rows][k stage of 32 columns][hi|lo]   for  d h = dO · Wᵀ        (K = conditioner columns)
//   B image  [n tile of 256 columns][k stage of 32 rows][hi|lo]   for  dW  = hᵀ · dO        (K = batch)
// (the bf16 hi/lo split mgtc_pack_kernel would have made of the same values: bit-identical operands, two HBM passes
// over dO fewer per flow layer) — and sums its columns for the bias gradient on the way (fixed order: eight rows in
// float, then a double per (CTA, column) across the CTA's row blocks, then the slices in order in stage 2).
// CTA = 8 rows x 32 dims (thread = one (sample, transformed dim) item, as above); the block's 8 x 32·PP values are
// staged in shared memory, so that both images are written in full 128-byte runs: eight consecutive rows of one
// 8-column chunk (A), eight consecutive columns of one 8-row chunk (B).  Row stride = 4 (mod 32) words: the A
// pass's 128-bit reads of eight rows are conflict-free.  Pads (rows up to the next multiple of 128, columns up to
// the next image tile) are written as zeros by the same passes, so no memset of the images is needed.
struct CgBwdImg {
  const float* v; const float* o; const float* g; float* gin;
  int64_t N, ldo;
  int D, tp, nd, ncomp;   // nd transformed dims, ncomp = nd * PP conditioner columns
  int opad;               // 0: o holds the dims back to back (column k*PP); else log2 of the column stride per dim (5: 32 floats)
  float lo, hi;
  uint8_t* a_img; int KSa;
  uint8_t* b_img; int KSb, NTb;
  double* colpart;        // [gridDim.y][ncomp]
  int rblocks;            // 8-row blocks, pads included
  const float* bias;      // o holds raw accumulators: + bias[(tp + 2k)*PP + p] (the last Linear's full bias); NULL: o is final
};

template <int K>
__global__ void __launch_bounds__(256) cg_transform_bwd_img_kernel(CgBwdImg a) {
  constexpr int PP = 3 * K + 1, CW = 32 * PP, RS = CW + 36 - (CW % 32), NJ = (CW + 255) / 256;
  static_assert(CW % 8 == 0 && RS % 32 == 4, "chunk alignment / bank layout");
  __shared__ __align__(16) float buf[8][RS];
  constexpr int NL = K == 8 ? 32 : (CW + 31) / 32;  // loads per lane that cover a row's run of this task's columns (K = 8: up to 32 floats per dim)
  __shared__ float sbe[NL * 32];               // bias of run element e (0 where o already holds the final values)
  __shared__ unsigned short sdst[NL * 32];     // run element e -> column inside the task (0xffff: pad of the tile layout)
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int c = blockIdx.x, k0 = c * 32, k = k0 + lane;
  const int per = (a.rblocks + gridDim.y - 1) / gridDim.y;
  const int rb0 = blockIdx.y * per, rb1 = min(a.rblocks, rb0 + per);
  const float ldbar = -1.f / (float)a.N;
  const int64_t cb = (int64_t)c * CW;  // first conditioner column of this task
  auto ocol = [&](int kk) { return a.opad ? kk << a.opad : kk * PP; };
  double csum[NJ];
#pragma unroll
  for (int j = 0; j < NJ; ++j) csum[j] = 0.0;
  // the task's columns inside a row of o: one contiguous run [q0, q1) (pads of the tile layout included); where each
  // element of the run goes, and its bias, is the same for every row: tabulated once
  const int klast = min(k0 + 31, a.nd - 1);
  const int q0 = k0 < a.nd ? ocol(k0) : 0, q1 = k0 < a.nd ? ocol(klast) + PP : 0;
  for (int e = tid; e < NL * 32; e += 256) {
    int d = 0xffff;
    float b = 0.f;
    if (e < q1 - q0) {
      const int q = q0 + e;
      if (!a.opad) d = e;
      else if ((q & ((1 << a.opad) - 1)) < PP) d = (q >> a.opad) * PP + (q & ((1 << a.opad) - 1)) - (int)cb;
      if (d != 0xffff && a.bias != nullptr) b = a.bias[(a.tp + 2 * (k0 + d / PP)) * PP + d % PP];
    }
    sdst[e] = (unsigned short)d;
    sbe[e] = b;
  }
  __syncthreads();
  for (int rb = rb0; rb < rb1; ++rb) {
    const int64_t r = (int64_t)rb * 8 + warp;
    float* row = buf[warp];
    const bool valid = r < a.N && k < a.nd;
    float xv = 0.f, gv = 0.f, gp = 0.f;
    if (valid) {  // issued before the staging loads complete
      const int64_t xi = r * a.D + a.tp + 2 * k;
      xv = a.v[xi];
      gv = a.g[xi];
      if (((a.tp + 2 * k) ^ 1) < a.D) gp = a.g[r * a.D + ((a.tp + 2 * k) ^ 1)];
    }
    if (r < a.N) {
      const float* __restrict__ src = a.o + r * a.ldo + q0;
      float t[NL];
#pragma unroll
      for (int i = 0; i < NL; ++i) {  // all loads first: one round trip per row block, not one per dim
        const int e = lane + 32 * i;
        t[i] = e < q1 - q0 ? src[e] : 0.f;
      }
      int dd[NL];  // table reads first: behind the stores below they would be ordered one by one (same address space)
#pragma unroll
      for (int i = 0; i < NL; ++i) {
        dd[i] = sdst[lane + 32 * i];
        t[i] += sbe[lane + 32 * i];
      }
#pragma unroll
      for (int i = 0; i < NL; ++i)
        if (dd[i] != 0xffff) row[dd[i]] = t[i];
    }
    __syncwarp();
    float obar[PP];
    if (valid) {
      const int d = a.tp + 2 * k;
      const int64_t xi = r * a.D + d;
      float ov[PP];
#pragma unroll
      for (int p = 0; p < PP; ++p) ov[p] = row[lane * PP + p];
      float y, l1, xbar;
      rqs_fwd_bwd<K, true>(ov, a.lo, a.hi, xv, gv, ldbar, y, l1, xbar, obar);
      a.gin[xi] = xbar;
      if ((d ^ 1) < a.D) a.gin[r * a.D + (d ^ 1)] = gp;  // the pass-through neighbour
      if (a.tp == 1 && (a.D & 1) && k == a.nd - 1) a.gin[r * a.D + a.D - 1] = a.g[r * a.D + a.D - 1];
    } else {
#pragma unroll
      for (int p = 0; p < PP; ++p) obar[p] = 0.f;
    }
    __syncwarp();
#pragma unroll
    for (int p = 0; p < PP; ++p) row[lane * PP + p] = obar[p];
    __syncthreads();
    // A image: (8-column chunk, row) items, the eight rows of a chunk adjacent
    {
      const int64_t r0 = (int64_t)rb * 8;
      const int64_t mt = r0 >> 7;
      for (int idx = tid; idx < CW; idx += 256) {
        const int row8 = idx & 7, cc = idx >> 3;
        const int64_t gc = (cb >> 3) + cc;  // global chunk of 8 columns
        if (gc >= (int64_t)a.KSa * 4) continue;
        const float4 f0 = *reinterpret_cast<const float4*>(&buf[row8][cc * 8]);
        const float4 f1 = *reinterpret_cast<const float4*>(&buf[row8][cc * 8 + 4]);
        uint32_t h[4], l[4];
        tc::split_bf16x2(f0.x, f0.y, h[0], l[0]);
        tc::split_bf16x2(f0.z, f0.w, h[1], l[1]);
        tc::split_bf16x2(f1.x, f1.y, h[2], l[2]);
        tc::split_bf16x2(f1.z, f1.w, h[3], l[3]);
        uint8_t* st = a.a_img + ((mt * a.KSa + (gc >> 2)) * 2) * (int64_t)kDtABytes +
                      tc::img_off(128, (int)((r0 + row8) & 127), (int)(gc & 3));
        *reinterpret_cast<uint4*>(st) = make_uint4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<uint4*>(st + kDtABytes) = make_uint4(l[0], l[1], l[2], l[3]);
      }
      // B image + column sums: one column per item, its eight rows are one 8-element K chunk
      const bool brow = r0 < (int64_t)a.KSb * 32;
      const int ks = (int)(r0 >> 5), c4 = (int)((r0 & 31) >> 3);
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        const int jj = tid + 256 * j;
        if (jj >= CW) break;
        const int64_t col = cb + jj;
        if (col >= (int64_t)a.NTb * kDtBN) continue;
        float f[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = buf[e][jj];
        csum[j] += (double)(((f[0] + f[1]) + (f[2] + f[3])) + ((f[4] + f[5]) + (f[6] + f[7])));
        if (!brow) continue;
        uint32_t h[4], l[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) tc::split_bf16x2(f[2 * e], f[2 * e + 1], h[e], l[e]);
        uint8_t* st = a.b_img + (((col >> 8) * a.KSb + ks) * 2) * (int64_t)kDtBBytes + tc::img_off(kDtBN, (int)(col & 255), c4);
        *reinterpret_cast<uint4*>(st) = make_uint4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<uint4*>(st + kDtBBytes) = make_uint4(l[0], l[1], l[2], l[3]);
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    const int jj = tid + 256 * j;
    const int64_t col = cb + jj;
    if (jj < CW && col < a.ncomp) a.colpart[(int64_t)blockIdx.y * a.ncomp + col] = csum[j];
  }
}

// gin[:, visible dims] += dmx[:, visible dims]
__global__ void __launch_bounds__(256) cg_add_visible_kernel(const float* __restrict__ dmx, float* __restrict__ gin, int64_t N,
                                                            int D, int tp) {
  const int64_t tot = N * D;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x)
    if ((int)(i % D & 1) != tp) gin[i] += dmx[i];
}


template <int K>
static int cg_run(const flowcuda_coupling_spec* sp, const CgLayout& L, const float* params, const float* x, int64_t n,
                  float* loss, float* grad, uint8_t* ws, cudaStream_t s, flowcuda_layer_grad_fn on_layer, void* user) {
  const int D = L.D, P = L.P, N = (int)n, nh = L.nh;
  float* S = (float*)(ws + L.off_state);
  float* G = (float*)(ws + L.off_g);
  float* MX = (float*)(ws + L.off_tmp);
  float* DMX = (float*)(ws + L.off_tmp2);
  float* LD = (float*)(ws + L.off_ld);
  float* LP = (float*)(ws + L.off_lp);
  float* dO = (float*)(ws + L.off_dO);
  float* part = (float*)(ws + L.off_part);
  float* act = (float*)(ws + L.off_act);
  const int splits = n >= 1024 ? kMgSplits : 1;
  const int eb = grid_for((int64_t)n * D, 256, 16), rb = (int)((n * 32 + 255) / 256);
  auto V = [&](int l) { return act + (size_t)l * L.per_layer_act; };
  auto H = [&](int l, int i) { return V(l) + L.act_h[i]; };  // i < nh: h_{i+1}; i == nh: o
  int rc;
  const MgTcScratch sc = {ws + L.off_aimg, L.a_bytes, ws + L.off_bimg, L.b_bytes, part, L.part_floats};
  auto mg_launch = [&](const MgGemm& g, cudaStream_t st) { return L.tc ? mgtc_launch(g, sc, st) : flowcuda::mg_launch(g, st); };
  if (L.fwd_dt) {
    // forward on the density path: per layer four GEMMs whose epilogues apply relu / the spline, write the next
    // operand image AND leave the reverse mode's operands behind (fp32 h_i, raw conditioner accumulators, layer inputs)
    DtTrainStores tr = {};
    tr.act = act; tr.per_layer_act = L.per_layer_act; tr.ldo = L.ldo_dt;
    for (int i = 0; i < L.nlin; ++i) tr.act_h[i] = L.act_h[i];
    if ((rc = coupling_dense_tc_train_forward(sp, params, x, n, tr, S, LD, ws + L.off_dt, L.dt_bytes, s))) return rc;
  } else {
  FC_CUDA(cudaMemcpyAsync(S, x, (size_t)n * D * 4, cudaMemcpyDeviceToDevice, s));
  for (int it = 0; it < L.L; ++it) {  // data -> base: layer L-1 first (Inverse(Chain(layers)))
    const int l = L.L - 1 - it, tp = l & 1;
    const float* Pl = params + L.per_layer * l;
    FC_CUDA(cudaMemcpyAsync(V(l), S, (size_t)n * D * 4, cudaMemcpyDeviceToDevice, s));
    cg_mask_input_kernel<<<eb, 256, 0, s>>>(S, MX, n, D, tp);
    const float* in = MX;
    const int ncomp = ((D - tp + 1) / 2) * P;  // conditioner columns of the dims this layer transforms
    const int64_t ldo = L.tc ? ncomp : 0;       // tensor-core engine: only those columns exist (compact o / dO)
    const MgColMap cm = {0, P, 2 * P, tp * P};
    for (int i = 0; i < L.nlin; ++i) {
      const bool last = i == nh && L.tc;
      const int nout = last ? ncomp : L.sizes[i + 1];
      MgGemm g = mg_gemm(in, L.sizes[i], 1, Pl + L.w_off[i], L.sizes[i + 1], 1, H(l, i), nout, N, nout, L.sizes[i]);
      g.bias = Pl + L.b_off[i];
      g.relu = i < nh;
      if (last) { g.map = cm; g.map.mode = 1; }
      if ((rc = mg_launch(g, s))) return rc;
      in = H(l, i);
    }
    cg_transform_fwd_kernel<K><<<rb, 256, 0, s>>>(V(l), H(l, nh), S, LD, n, D, P, tp, sp->range_min, sp->range_max, it == 0, ldo);
    FC_LAUNCH_CHECK("coupling_nll_grad forward");
  }
  }
  cg_finish_kernel<<<rb, 256, 0, s>>>(S, LD, LP, G, n, D, sp->base);
  FC_LAUNCH_CHECK("cg_finish_kernel");
  if ((rc = flowcuda_sum_scaled(LP, n, -1.0 / (double)n, loss, ws + L.off_sum, s))) return rc;
  // Backward.  Main stream: bijector adjoint -> input-gradient GEMMs (the dependency chain).  Side stream: per Linear the
  // weight gradient (operand packs, split-K GEMM, fixed-order reduce) and the bias gradient — off the chain, so they fill
  // the SMs the chain's small GEMMs (64 CTAs) leave idle.  Same kernels, same arguments, same summation orders as the
  // single-stream schedule: the results are bit-identical.
  SideCtx* sd = nullptr;
  static const bool overlap_env = side_env_enabled("FLOWCUDA_TRAIN_OVERLAP");
  static const bool hook_split = side_env_enabled("FLOWCUDA_HOOK_SPLIT");  // 0: one report per flow layer
  const bool overlap = L.tc && overlap_env;
  std::unique_lock<std::recursive_mutex> side_lock(side_mutex(), std::defer_lock);
  if (overlap) {
    side_lock.lock();
    if ((rc = side_ctx_get(&sd))) return rc;
  }
  cudaStream_t q = overlap ? sd->q : s;
  const MgTcScratch sc2 = {ws + L.off_aimg2, L.a2_bytes, ws + L.off_bimg2, L.b2_bytes, (float*)(ws + L.off_part2), L.part_floats};
  float* part_q = overlap ? (float*)(ws + L.off_part2) : part;
  float* MXp[2] = {MX, overlap ? (float*)(ws + L.off_mx2) : MX};
  for (int l = 0; l < L.L; ++l) {  // the layer applied last first
    const int tp = l & 1, lp = l & 1;
    const float* Pl = params + L.per_layer * l;
    float* Gl = grad + L.per_layer * l;
    const int ncomp = ((D - tp + 1) / 2) * P;
    const int64_t ldo = L.tc ? ncomp : 0;
    const MgColMap cm = {0, P, 2 * P, tp * P};
    const int nd = ncomp / P;
    const int KSa = (ncomp + kDtBK - 1) / kDtBK, KSb = (int)((n + kDtBK - 1) / kDtBK), NTb = (ncomp + kDtBN - 1) / kDtBN;
    int cslices = 1;
    float* MXl = MXp[lp];
    if (overlap) {
      // parity-lp buffers (masked input, dH) were last read by the side jobs of layer l-2; dO / its column sums by the
      // last-Linear job of layer l-1
      if (l >= 2) FC_CUDA(cudaStreamWaitEvent(s, sd->ev_s[lp], 0));
      if (l >= 1) FC_CUDA(cudaStreamWaitEvent(s, sd->ev_last, 0));
    }
    if constexpr (K > 0) {
      if (L.img) {
        CgBwdImg bi = {};
        bi.v = V(l); bi.o = H(l, nh); bi.g = G; bi.gin = S; bi.N = n; bi.ldo = ldo; bi.D = D; bi.tp = tp; bi.nd = nd;
        bi.ncomp = ncomp; bi.opad = 0; bi.lo = sp->range_min; bi.hi = sp->range_max;
        if (L.fwd_dt) { bi.opad = 5; bi.ldo = L.ldo_dt; bi.bias = Pl + L.b_off[nh]; }  // 2^5 floats per dim
        bi.a_img = ws + L.off_dOa; bi.KSa = KSa; bi.b_img = ws + L.off_dOb; bi.KSb = KSb; bi.NTb = NTb;
        bi.colpart = (double*)(ws + L.off_colp);
        bi.rblocks = (int)(((n + 127) / 128) * 16);
        cslices = bi.rblocks < kCgColSlices ? bi.rblocks : kCgColSlices;
        constexpr int CW = 32 * (3 * K + 1);
        const int64_t cols = (int64_t)KSa * kDtBK > (int64_t)NTb * kDtBN ? (int64_t)KSa * kDtBK : (int64_t)NTb * kDtBN;
        cg_transform_bwd_img_kernel<K><<<dim3((unsigned)((cols + CW - 1) / CW), cslices), 256, 0, s>>>(bi);
      }
    }
    if (L.img) {
    } else if (ldo)
      cg_transform_bwd_compact_kernel<K><<<grid_for(((int64_t)n * (ncomp / P) + 31) / 32, 4, 16), 128, 0, s>>>(
          V(l), H(l, nh), G, S, dO, n, D, tp, sp->range_min, sp->range_max);
    else
      cg_transform_bwd_kernel<K><<<grid_for((int64_t)n * D, 128, 32), 128, 0, s>>>(V(l), H(l, nh), G, S, dO, n, D, P, tp,
                                                                                  sp->range_min, sp->range_max, ldo);
    FC_LAUNCH_CHECK("cg_transform_bwd_kernel");
    cg_mask_input_kernel<<<eb, 256, 0, s>>>(V(l), MXl, n, D, tp);
    const float* dout = dO;
    for (int i = nh; i >= 0; --i) {
      const bool last = i == nh && L.tc;  // compact dO: the other columns' gradients are exactly zero
      const bool pre = last && L.img;     // ... and it exists only as the two operand images
      const int kin = L.sizes[i], nfull = L.sizes[i + 1], nout = last ? ncomp : nfull;
      const float* in = i == 0 ? MXl : H(l, i - 1);
      if (overlap) {  // dout (and, for i = nh, the masked input) is now enqueued on the main stream
        FC_CUDA(cudaEventRecord(sd->ev_m, s));
        FC_CUDA(cudaStreamWaitEvent(q, sd->ev_m, 0));
      }
      if (on_layer && hook_split && i == nh - 1) {
        // The last Linear's weights and bias — nine tenths of the layer's parameters — are final on q here: its
        // gradient jobs were enqueued first, and q has just been ordered behind the input-gradient GEMM that was the
        // last reader of its weights.  Reported now, its all-reduce / update runs under the rest of the layer instead
        // of after it (the last layer's bucket is otherwise the exposed tail of the whole step).
        on_layer(user, l, L.per_layer * l + L.w_off[nh], L.per_layer - L.w_off[nh], (void*)q);
      }
      // dW_i = inᵀ · dout (split-K, fixed-order reduce), db_i = column sums                                [side stream]
      if (last) FC_CUDA(cudaMemsetAsync(Gl + L.w_off[i], 0, ((size_t)kin * nfull + nfull) * sizeof(float), q));
      MgColMap bm = cm;
      bm.mode = last ? 3 : 0;
      if (pre) {  // the partial column sums came with the images
        mg_colsum_stage2<<<(nout + 255) / 256, 256, 0, q>>>((const double*)(ws + L.off_colp), cslices, nout, Gl + L.b_off[i], bm);
        FC_LAUNCH_CHECK("mg_colsum_stage2");
      }
      MgGemm g = mg_gemm(in, 1, kin, dout, nout, 1, Gl + L.w_off[i], nfull, kin, nout, n);
      g.splits = splits; g.part = part_q;
      if (last) { g.map = cm; g.map.mode = 3; }
      if (L.tc) { if ((rc = mgtc_launch(g, overlap ? sc2 : sc, q, MgTcPre{nullptr, pre ? ws + L.off_dOb : nullptr}))) return rc; }
      else if ((rc = flowcuda::mg_launch(g, q))) return rc;
      if (!pre && (rc = mg_colsum(dout, n, nout, Gl + L.b_off[i], part_q, q, bm))) return rc;
      if (overlap && i == nh) FC_CUDA(cudaEventRecord(sd->ev_last, q));
      // d in = dout · W_iᵀ (relu gate of the stored activation; the masked input has no gate)             [main stream]
      float* din = i == 0 ? DMX : (float*)(ws + L.off_dhx[overlap ? lp : 0][i]);
      g = mg_gemm(dout, nout, 1, Pl + L.w_off[i], 1, nfull, din, kin, N, kin, nout);
      if (last) { g.map = cm; g.map.mode = 2; }
      if (i > 0) { g.gate = H(l, i - 1); g.ldg = kin; }
      if (pre) { if ((rc = mgtc_launch(g, sc, s, MgTcPre{ws + L.off_dOa, nullptr}))) return rc; }
      else if ((rc = mg_launch(g, s))) return rc;
      dout = din;
    }
    if (overlap) FC_CUDA(cudaEventRecord(sd->ev_s[lp], q));
    if (on_layer) {
      // The rest of this layer's gradient slice is complete on q (every dW / db job of the layer has been enqueued there).  q is
      // also made to follow the layer's input-gradient chain, the last reader of the layer's PARAMETERS in this call:
      // the callee may update them in place from the hook (optimiser step per layer, under the remaining layers).
      if (overlap) {
        FC_CUDA(cudaEventRecord(sd->ev_m, s));
        FC_CUDA(cudaStreamWaitEvent(q, sd->ev_m, 0));
      }
      if (hook_split) on_layer(user, l, L.per_layer * l, L.w_off[nh], (void*)q);  // the hidden Linears' slice; the last Linear's went out above
      else on_layer(user, l, L.per_layer * l, L.per_layer, (void*)q);
    }
    cg_add_visible_kernel<<<eb, 256, 0, s>>>(DMX, S, n, D, tp);
    FC_LAUNCH_CHECK("coupling_nll_grad backward");
    float* t = G; G = S; S = t;  // S held dL/d(layer input): it is the next layer's output gradient
  }
  if (overlap) {  // join: everything the side stream did is ordered before whatever the caller enqueues next
    FC_CUDA(cudaEventRecord(sd->ev_m, q));
    FC_CUDA(cudaStreamWaitEvent(s, sd->ev_m, 0));
  }
  return FLOWCUDA_OK;
}

bool coupling_dense_grad_supported(const flowcuda_coupling_spec* sp) {
  CgLayout L;
  return cg_layout(sp, 1, L);
}

size_t coupling_dense_grad_workspace_bytes(const flowcuda_coupling_spec* sp, int64_t n) {
  CgLayout L;
  if (!cg_layout(sp, n, L)) return 0;
  return L.total;
}

int coupling_dense_grad_run(const flowcuda_coupling_spec* sp, const float* params, const float* x, int64_t n, float* loss,
                            float* grad, void* workspace, size_t wsb, cudaStream_t s, flowcuda_layer_grad_fn on_layer, void* user) {
  CgLayout L;
  if (!cg_layout(sp, n, L)) {
    set_error("coupling_nll_grad: needs RQS with 8 or 12 bins or the per-dimension affine transform");
    return FLOWCUDA_EUNSUPPORTED;
  }
  if (!workspace || wsb < L.total) {
    set_error("coupling_nll_grad: workspace %zu < %zu bytes", wsb, L.total);
    return FLOWCUDA_EWORKSPACE;
  }
  FC_REQUIRE(n < (1ll << 31) / (L.D * L.P > 0 ? 1 : 1), "coupling_nll_grad: n too large");
  uint8_t* ws = (uint8_t*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  if (sp->transform == FLOWCUDA_TRANSFORM_AFFINE) return cg_run<0>(sp, L, params, x, n, loss, grad, ws, s, on_layer, user);
  if (sp->num_bins == 8) return cg_run<8>(sp, L, params, x, n, loss, grad, ws, s, on_layer, user);
  return cg_run<12>(sp, L, params, x, n, loss, grad, ws, s, on_layer, user);
}

}  // namespace flowcuda
