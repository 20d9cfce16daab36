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
//            the transformed dims' columns; a 256-column tile is two self-contained halves of 5 dims x 25 spline
//            params (+3 pad) each, so that the tail of the tile list can be split into half tiles)
// CTA = loader warp + MMA warp + 4 epilogue warps, 2 CTAs per SM so that one CTA's epilogue
// overlaps the other's MMAs.  Epilogues (thread = sample row = TMEM lane):
//   hidden: +bias, relu, bf16 hi/lo split, written straight into the next Linear's A image;
//   last:   the inner bijector (RQS / affine) on the tile's dims, state updated in place,
//           log-det partial per (row, tile) — the [N, 19600] conditioner output never exists.
#include "dense_tc_chain.cuh"

namespace flowcuda {

// ---- weight packing ----------------------------------------------------------------------------
struct DtPackW {
  const float* params;   // flow parameters, all layers
  int64_t per_layer;     // floats per flow layer
  int64_t w_off;         // offset of this Linear's W inside a layer
  int k_in, n_out;       // W[k_in, n_out] row-major
  int kstride, kvalid;   // image k index j <-> W row koff(l) + kstride*j, j < kvalid
  int first;             // 1: first Linear (koff = visible parity of the layer), else koff = 0
  int dpt, ppd, D;       // last Linear: dims per tile, params per dim (dpt = 0: plain columns); D = event dim
  int KS, NT;            // k stages, n tiles
  uint8_t* img;          // [layer][NT][KS][hi|lo][tr rows x 64 B]
  int tr;                // rows per n tile of the image: 256 (dense_tc_gemm_kernel)
  uint8_t* img128;       // hidden Linears: the same weights once more in 128-row tiles (dense_tc_chain_kernel), or NULL
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
  const int TR = a.tr;
  const int64_t half = (int64_t)TR * kDtBK * 2;  // bytes of one of hi/lo of a stage
  uint8_t* base = a.img + ((int64_t)l * a.NT + nt) * a.KS * 2 * half;
  const int total = a.KS * 4 * TR;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int n = i % TR, ck = i / TR;
    int col = -1;
    if (a.dpt == 0) {
      col = nt * TR + n;
      if (col >= a.n_out) col = -1;
    } else {
      int di, pp;  // tile column -> (dim of the tile, parameter): two self-contained 128-column halves (dt_col_to_dim)
      if (dt_col_to_dim(n, a.dpt, a.ppd, di, pp)) {
        const int dim = nt * a.dpt + di;
        if (dim < ndims) col = (tp + 2 * dim) * a.ppd + pp;
      }
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
    uint8_t* st = base + (int64_t)(ck >> 2) * 2 * half + tc::img_off(TR, n, ck & 3);
    *reinterpret_cast<uint4*>(st) = make_uint4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<uint4*>(st + half) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    if (a.img128 != nullptr) {  // tile (2 nt + n / 128) of twice as many 128-row tiles, same K chunk
      const int64_t half2 = (int64_t)128 * kDtBK * 2;
      uint8_t* s2 = a.img128 + (((int64_t)l * 2 * a.NT + 2 * nt + (n >> 7)) * a.KS + (ck >> 2)) * 2 * half2 + tc::img_off(128, n & 127, ck & 3);
      *reinterpret_cast<uint4*>(s2) = make_uint4(h[0], h[1], h[2], h[3]);
      *reinterpret_cast<uint4*>(s2 + half2) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
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
  size_t off_wc[FLOWCUDA_MAX_HIDDEN_LAYERS + 1];  // hidden Linears once more, in 128-row tiles (dense_tc_chain.cuh)
  int chain_ok;                                   // every hidden Linear is 512 wide: the chain kernel can run
};

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
  L.MT = 2 * ((n + 255) / 256);  // 128-row tiles, even: a CTA pair works on 256 rows (dense_tc2.cuh)
  for (int i = 0; i < L.nlin; ++i) {
    const int kin = i == 0 ? ndmax : L.sizes[i];
    L.KS[i] = (kin + kDtBK - 1) / kDtBK;
    L.NT[i] = i < nh ? (L.sizes[i + 1] + kDtBN - 1) / kDtBN : L.tiles;
  }
  size_t off = 0;
  // the weight images come first: their offsets do not depend on n, so a workspace packed once
  // (flowcuda_*_pack_weights) serves every later batch size that fits it
  for (int i = 0; i < L.nlin; ++i) {
    L.off_w[i] = off;
    off = dt_al256(off + (size_t)sp->num_layers * L.NT[i] * L.KS[i] * 2 * kDtBBytes);
  }
  // the hidden Linears' weights (3 MB per flow layer at the maf_mnist shape) are kept in BOTH tile heights, so that a
  // workspace packed once serves calls of any batch size, whichever kernel they pick
  L.chain_ok = nh >= 1 && nh <= kDcMaxLinears;
  for (int i = 0; i < nh && L.chain_ok; ++i) L.chain_ok = L.sizes[i + 1] == 512;
  for (int i = 0; i < L.nlin; ++i) {
    L.off_wc[i] = off;
    if (L.chain_ok && i < nh) off = dt_al256(off + (size_t)sp->num_layers * L.NT[i] * L.KS[i] * 2 * kDtBBytes);
  }
  L.off_v = off; off = dt_al256(off + (size_t)n * D * 4);
  L.off_ld = off; off = dt_al256(off + (size_t)n * 4);
  // four log-det partials per (row, tile) — one per 64-column quarter — and one slab per flow layer (summed once, by the finish kernel)
  L.off_ldp = off; off = dt_al256(off + (size_t)sp->num_layers * n * L.tiles * 4 * 4);
  L.off_x = off; off = dt_al256(off + (size_t)L.MT * L.KS[0] * 2 * kDtABytes);
  int ksh = 1;
  for (int i = 1; i < L.nlin; ++i) ksh = L.KS[i] > ksh ? L.KS[i] : ksh;
  for (int j = 0; j < 2; ++j) { L.off_h[j] = off; off = dt_al256(off + (size_t)L.MT * ksh * 2 * kDtABytes); }
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

static int dt_run(const flowcuda_coupling_spec* sp, const float* params, const float* in, float* out_v,
                  float* out_lp, float* out_ld, int64_t n, int layer_begin, int layer_end, int with_base,
                  bool inverse, void* workspace, size_t workspace_bytes, cudaStream_t s, int mode, const DtTrainStores* tr) {
  // mode 0: pack the weights, then run; 1: the workspace already holds this flow's weight images; 2: pack only
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
  if (mode != 2) FC_CUDA(cudaMemcpyAsync(v, in, (size_t)n * D * sizeof(float), cudaMemcpyDeviceToDevice, s));
  // the hidden Linears as one cluster kernel (dense_tc_chain.cuh) when every one is 512 wide and the batch is one wave
  // of clusters; its weight images (128-row tiles) are packed beside the 256-row ones whenever the shapes allow it
  const bool chain_pack = L.chain_ok && dt_chain_possible() && dt2_mode() == 1;
  const bool chain = chain_pack && dt_chain_wanted(L.MT);
  // pack every Linear of every layer (masks folded in)
  for (int i = 0; i < L.nlin && mode != 1; ++i) {
    DtPackW p;
    p.tr = kDtBN;
    p.params = params; p.per_layer = L.per_layer; p.w_off = L.w_off[i];
    p.k_in = L.sizes[i]; p.n_out = L.sizes[i + 1];
    p.first = i == 0; p.kstride = i == 0 ? 2 : 1; p.kvalid = L.sizes[i];
    p.dpt = i == nh ? L.dpt : 0; p.ppd = L.ppd; p.D = D;
    p.KS = L.KS[i]; p.NT = L.NT[i]; p.img = ws + L.off_w[i];
    p.img128 = chain_pack && i < nh ? ws + L.off_wc[i] : nullptr;  // both tile heights from the one pass over the weights
    int blocks = (L.KS[i] * 4 * kDtBN + 255) / 256;
    dense_tc_pack_w_kernel<<<dim3(blocks, L.NT[i], sp->num_layers), 256, 0, s>>>(p);
    FC_LAUNCH_CHECK("dense_tc_pack_w_kernel");
  }
  if (mode == 2) return FLOWCUDA_OK;
  const int nl = layer_end - layer_begin;
  // the finish kernel sums every slot of the processed layers' slabs; tiles a layer does not use (odd event sizes: the
  // two parities differ by one dim) must read as zero
  if (nl > 0) FC_CUDA(cudaMemsetAsync(ldp, 0, (size_t)nl * n * 4 * L.tiles * sizeof(float), s));
  int used2 = 0;
  for (int it = 0; it < nl; ++it) {
    const int l = inverse ? layer_begin + it : layer_end - 1 - it;
    const int vis = (l & 1) ? 0 : 1, tp = 1 - vis;
    const int ndims = (D - tp + 1) / 2, nvis = (D - vis + 1) / 2;
    const float* lp = params + L.per_layer * l;
    // the first layer's activation image comes from the state; every later one is written by the previous layer's
    // transform epilogue (its transformed dims are this layer's visible dims).  Odd event sizes: the two parities
    // differ by one dim, whose column the epilogue never writes - keep the packing pass there.
    const bool fuse_x = (D % 2) == 0;
    if (tr) FC_CUDA(cudaMemcpyAsync(tr->act + (size_t)l * tr->per_layer_act, v, (size_t)n * D * sizeof(float), cudaMemcpyDeviceToDevice, s));
    if (it == 0 || !fuse_x) {
      int64_t total = L.MT * 128 * L.KS[0] * 4;
      int blocks = (int)((total + 255) / 256);
      FC_CUDA(launch_chain(dense_tc_pack_x_kernel, dim3(blocks), dim3(256), 0, s, (const float*)v, n, D, vis, 2, nvis, L.KS[0],
                           ws + L.off_x));
    }
    const uint8_t* aimg = ws + L.off_x;
    int first_lin = 0;
    if (chain) {
      DtChain c = {};
      c.nlin = nh; c.N = n; c.a_img0 = aimg;
      for (int i = 0; i < nh; ++i) {
        c.KS[i] = L.KS[i];
        c.b_img[i] = ws + L.off_wc[i] + (size_t)l * L.NT[i] * L.KS[i] * 2 * kDtBBytes;
        c.bias[i] = lp + L.b_off[i];
        c.out_img[i] = ws + L.off_h[i & 1];
        c.out_KS[i] = L.KS[i + 1];
        c.store_h[i] = tr ? tr->act + (size_t)l * tr->per_layer_act + tr->act_h[i] : nullptr;
      }
      int rc = dt_launch_chain(c, L.MT, s);
      if (rc) return rc;
      aimg = c.out_img[nh - 1];
      first_lin = nh;
    }
    for (int i = first_lin; i < L.nlin; ++i) {
      DtGemm g = {};
      g.a_img = aimg;
      g.b_img = ws + L.off_w[i] + (size_t)l * L.NT[i] * L.KS[i] * 2 * kDtBBytes;
      g.bias = lp + L.b_off[i];
      g.KS = L.KS[i];
      g.n_out = L.sizes[i + 1];
      g.out_img = nullptr; g.out_KS = 0;
      g.v = v; g.ldp = ldp + (size_t)it * n * 4 * L.tiles; g.N = n; g.D = D; g.tp = tp; g.dstride = 2; g.ndims = ndims; g.dpt = L.dpt; g.ppd = L.ppd;
      g.tiles = 4 * L.tiles; g.lo = sp->range_min; g.hi = sp->range_max;
      for (int i = 0; i < 9; ++i) g.koff[i] = sp->range_min + (float)i * kRqsMinBin;
      int rc;
      // CTA pairs for the RQS transform Linear (its single-CTA form outruns the L2 feed) and for hidden Linears of
      // batches beyond the chain kernel's one wave (density / sampling passes; the training forward keeps the kernels
      // it was measured with); single CTAs for the affine output Linear (dense_tc2.cuh, dt2_mode)
      const bool pairs = i == nh ? sp->transform == FLOWCUDA_TRANSFORM_RQS : (tr == nullptr && L.MT * 4 > sm_count());
      const Dt2Launch dl = {L.MT, (int64_t)sp->num_layers * L.NT[i], (int64_t)l * L.NT[i], pairs ? 2 : 1};
      if (tr) {
        float* blk = tr->act + (size_t)l * tr->per_layer_act + tr->act_h[i];
        if (i < nh) { g.store_h = blk; g.store_ld = L.sizes[i + 1]; }
        else { g.store_o = blk; g.store_ldo = tr->ldo; }
      }
      if (i < nh) {
        g.out_img = ws + L.off_h[i & 1];
        g.out_KS = L.KS[i + 1];
        rc = dt2_launch_gemm<kEpiHidden, false>(g, dl, L.MT, L.NT[i], s);
        aimg = g.out_img;
      } else {
        const int used = (ndims + L.dpt - 1) / L.dpt;
        if (fuse_x && it + 1 < nl) {
          g.out_img = ws + L.off_x;
          g.out_KS = L.KS[0];
          g.next_rev = 0;
        }
        if (sp->transform == FLOWCUDA_TRANSFORM_RQS && tr && !inverse)
          rc = dt_launch_gemm<kEpiRqs8Store, false>(g, L.MT, used, s);  // training forward: single-CTA kernel with the accumulator store
        else if (sp->transform == FLOWCUDA_TRANSFORM_RQS)
          rc = inverse ? dt2_launch_gemm<kEpiRqs8, true>(g, dl, L.MT, used, s) : dt2_launch_gemm<kEpiRqs8, false>(g, dl, L.MT, used, s);
        else
          rc = inverse ? dt2_launch_gemm<kEpiAffine, true>(g, dl, L.MT, used, s) : dt2_launch_gemm<kEpiAffine, false>(g, dl, L.MT, used, s);
        if (rc) return rc;
        used2 = 4 * L.tiles;
      }
      if (rc) return rc;
    }
  }
  (void)ld;
  if (out_lp || out_ld) {
    FC_CUDA(launch_chain(dense_tc_finish_kernel, dim3((unsigned)((n * 32 + 255) / 256)), dim3(256), 0, s, (const float*)v,
                         (const float*)ldp, nl, 4 * L.tiles, used2, n, D, with_base, (int)sp->base, out_lp, out_ld));
  }
  if (out_v) FC_CUDA(cudaMemcpyAsync(out_v, v, (size_t)n * D * sizeof(float), cudaMemcpyDeviceToDevice, s));
  return FLOWCUDA_OK;
}

int coupling_dense_tc_run(const flowcuda_coupling_spec* sp, const float* params, const float* in, float* out_v,
                          float* out_lp, float* out_ld, int64_t n, int layer_begin, int layer_end, int with_base,
                          bool inverse, void* workspace, size_t workspace_bytes, cudaStream_t s, int mode) {
  return dt_run(sp, params, in, out_v, out_lp, out_ld, n, layer_begin, layer_end, with_base, inverse, workspace,
                workspace_bytes, s, mode, nullptr);
}

int64_t coupling_dense_tc_store_ldo(const flowcuda_coupling_spec* sp) {
  DtLayout L;
  if (!dt_layout(sp, 128, L)) return 0;
  return (int64_t)((sp->event_dim + 1) / 2) * 32;  // 32 floats per transformed dim (dense_tc.cuh: store_o)
}

int coupling_dense_tc_train_forward(const flowcuda_coupling_spec* sp, const float* params, const float* x, int64_t n,
                                    const DtTrainStores& tr, float* out_z, float* out_ld, void* workspace,
                                    size_t workspace_bytes, cudaStream_t s) {
  return dt_run(sp, params, x, out_z, nullptr, out_ld, n, 0, sp->num_layers, 0, false, workspace, workspace_bytes, s, 0, &tr);
}

}  // namespace flowcuda

#ifdef FLOWCUDA_TRACE
// diagnostics build only (tools/dt_trace.py): point the GEMM kernels of this translation unit at a device buffer of
// 8 counters per CTA
extern "C" int flowcuda_debug_set_dt_trace(unsigned long long* device_buf) {
  return cudaMemcpyToSymbol(flowcuda::g_dt_trace, &device_buf, sizeof(device_buf)) == cudaSuccess ? 0 : -2;
}
#endif
