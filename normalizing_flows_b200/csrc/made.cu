// made.cu — row #8: MADE-conditioned masked autoregressive affine flow (MAF), the component
// BASELINE.json's north_star names for the D=784 configs.  The reference repository has no
// MADE code (src/maf_mwe.py:36-88 is commented out and is not MADE); the semantics follow
// Germain et al. 2015 / Papamakarios et al. 2017 as restated in SURVEY App. A.7 and
// oracle/flows_oracle.py:510-606 (PARITY UNPINNED):
//   input degrees 1..D, hidden degrees max(1, ceil(k (D-1) / (H+1))), k = 1..H,
//   hidden mask deg_in <= deg_out, output mask deg_in < deg_out, relu hidden activations,
//   output columns [shift_0..shift_{D-1}, log_scale_0..log_scale_{D-1}],
//   odd layers see the dimensions in reversed order.
//
// Density pass (one conditioner evaluation per layer) = three masked GEMMs on the tensor cores
// (dense_tc.cuh): the masks and the reversal are folded into the packed weight images, and since a
// masked Linear is block lower-triangular in degree order an output tile only walks the K stages
// that can be non-zero.  The last GEMM's epilogue applies u = (x - shift)·exp(-log_scale) in place
// and emits the log-det partial, so the [N, 2D] conditioner output never reaches HBM.
//
// Sampling is inherently sequential over dimensions.  made_sample_kernel is a persistent kernel:
// a CTA owns S samples for the whole stack and walks q = 0..D-1; all pre-activations live in
// registers and are updated INCREMENTALLY (rank-1 updates as x_q, then the hidden units whose
// degree has just completed, become known), so a layer costs one conditioner evaluation's worth of
// FMAs instead of D of them.
#include "dense_tc2.cuh"

namespace flowcuda {

// degree of hidden unit k (0-based) of a width-H layer, event dim D
__host__ __device__ __forceinline__ int made_hdeg(int k, int H, int D) {
  const int d = (int)(((int64_t)(k + 1) * (D - 1) + H) / (H + 1));  // ceil((k+1)(D-1)/(H+1))
  return d < 1 ? 1 : d;
}
// smallest hidden unit index whose degree is >= m (H if none)
__host__ __device__ __forceinline__ int made_kmin(int m, int H, int D) {
  if (m <= 1) return 0;
  if (D <= 1) return H;
  const int64_t k = ((int64_t)(m - 1) * (H + 1)) / (D - 1);
  return k > H ? H : (int)k;
}

struct MadeLayout {
  int D, nh, nlin, sizes[FLOWCUDA_MAX_HIDDEN_LAYERS + 2];
  int64_t w_off[FLOWCUDA_MAX_HIDDEN_LAYERS + 1], b_off[FLOWCUDA_MAX_HIDDEN_LAYERS + 1], per_layer;
  int KS[FLOWCUDA_MAX_HIDDEN_LAYERS + 1], NT[FLOWCUDA_MAX_HIDDEN_LAYERS + 1], tiles;
  int64_t MT;
  size_t off_v, off_ld, off_ldp, off_x, off_h[2], off_w[FLOWCUDA_MAX_HIDDEN_LAYERS + 1], total;
};

static bool made_layout(const flowcuda_made_spec* sp, int64_t n, MadeLayout& L) {
  const int D = sp->event_dim, nh = sp->num_hidden;
  L.D = D; L.nh = nh; L.nlin = nh + 1;
  L.sizes[0] = D;
  for (int i = 0; i < nh; ++i) L.sizes[i + 1] = sp->hidden[i];
  L.sizes[nh + 1] = 2 * D;
  int64_t o = 0;
  for (int i = 0; i < L.nlin; ++i) {
    L.w_off[i] = o; o += (int64_t)L.sizes[i] * L.sizes[i + 1];
    L.b_off[i] = o; o += L.sizes[i + 1];
  }
  L.per_layer = o;
  L.tiles = (D + 127) / 128;
  L.MT = 2 * ((n + 255) / 256);  // 128-row tiles, even: a CTA pair works on 256 rows (dense_tc2.cuh)
  for (int i = 0; i < L.nlin; ++i) {
    L.KS[i] = (L.sizes[i] + kDtBK - 1) / kDtBK;
    L.NT[i] = i < nh ? (L.sizes[i + 1] + kDtBN - 1) / kDtBN : L.tiles;
    if (L.KS[i] > 255) return false;  // ks_tile entries are bytes
  }
  size_t off = 0;
  // the weight images come first: their offsets do not depend on n, so a workspace packed once
  // (flowcuda_*_pack_weights) serves every later batch size that fits it
  for (int i = 0; i < L.nlin; ++i) {
    L.off_w[i] = off;
    off = dt_al256(off + (size_t)sp->num_layers * L.NT[i] * L.KS[i] * 2 * kDtBBytes);
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

// ---- weight packing: masks and the reversal folded in ---------------------------------------------
struct MadePackW {
  const float* params;
  int64_t per_layer, w_off;
  int k_in, n_out;   // W[k_in, n_out] row-major
  int H_in, H_out;   // widths for the degree formula; H_in = 0: inputs (degree j+1); H_out = 0: the output Linear
  int D, KS, NT;
  uint8_t* img;      // [layer][NT][KS][hi|lo][16 KB]
};

// grid (blocks, NT, layers)
__global__ void __launch_bounds__(256) made_tc_pack_w_kernel(MadePackW a) {
  const int l = blockIdx.z, nt = blockIdx.y;
  const float* W = a.params + a.per_layer * l + a.w_off;
  const bool rev = (l & 1) != 0;
  uint8_t* base = a.img + ((int64_t)l * a.NT + nt) * a.KS * 2 * (int64_t)kDtBBytes;
  const int total = a.KS * 4 * kDtBN;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int n = i % kDtBN, ck = i / kDtBN;
    int col = -1, dout = 0;
    if (a.H_out > 0) {  // hidden Linear: plain columns, mask deg_in <= deg_out
      col = nt * kDtBN + n;
      if (col >= a.n_out) col = -1;
      else dout = made_hdeg(col, a.H_out, a.D);
    } else {  // output Linear: tile column n = (state dim d = 128 nt + n/2, shift | log_scale), mask deg_in < deg_out
      const int d = nt * 128 + (n >> 1), pp = n & 1;
      if (d < a.D) {
        const int q = rev ? a.D - 1 - d : d;
        col = pp * a.D + q;
        dout = q;  // deg_in < q+1  <=>  deg_in <= q
      }
    }
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int j = ck * 8 + e;
      float w = 0.f;
      if (col >= 0 && j < a.k_in) {
        const int din = a.H_in > 0 ? made_hdeg(j, a.H_in, a.D) : j + 1;
        if (din <= dout) w = W[(int64_t)j * a.n_out + col];
      }
      f[e] = w;
    }
    uint32_t h[4], lo[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) tc::split_bf16x2(f[2 * e], f[2 * e + 1], h[e], lo[e]);
    uint8_t* st = base + (int64_t)(ck >> 2) * 2 * kDtBBytes + tc::img_off(kDtBN, n, ck & 3);
    *reinterpret_cast<uint4*>(st) = make_uint4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<uint4*>(st + kDtBBytes) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
  }
}

// K stages tile nt of Linear i needs in a layer of the given parity (the rest of its rows are masked out)
static int made_ks_needed(const MadeLayout& L, int i, int nt, bool rev) {
  const int D = L.D;
  int max_dout;  // inputs with degree <= max_dout can reach this tile
  if (i < L.nh) {
    const int H = L.sizes[i + 1];
    int c = nt * kDtBN + kDtBN - 1;
    if (c > H - 1) c = H - 1;
    max_dout = made_hdeg(c, H, D);
  } else {
    int qmax;
    if (rev) qmax = D - 1 - nt * 128;
    else { qmax = nt * 128 + 127; if (qmax > D - 1) qmax = D - 1; }
    max_dout = qmax;  // deg_in < qmax + 1
  }
  const int kin = L.sizes[i];
  int cnt;
  if (i == 0) cnt = max_dout < kin ? max_dout : kin;  // degrees 1..D
  else cnt = made_kmin(max_dout + 1, kin, D);         // hidden units with degree <= max_dout
  int ks = (cnt + kDtBK - 1) / kDtBK;
  if (ks < 1) ks = 1;
  if (ks > L.KS[i]) ks = L.KS[i];
  return ks;
}

static int made_check(const flowcuda_made_spec* sp) {
  FC_REQUIRE(sp, "made: NULL spec");
  FC_REQUIRE(sp->event_dim >= 1, "made: event_dim=%d must be >= 1", sp->event_dim);
  FC_REQUIRE(sp->num_layers >= 1 && sp->num_layers <= 1024, "made: num_layers=%d", sp->num_layers);
  FC_REQUIRE(sp->num_hidden >= 1 && sp->num_hidden <= FLOWCUDA_MAX_HIDDEN_LAYERS, "made: num_hidden=%d outside [1,%d]",
             sp->num_hidden, FLOWCUDA_MAX_HIDDEN_LAYERS);
  for (int i = 0; i < sp->num_hidden; ++i) FC_REQUIRE(sp->hidden[i] >= 1, "made: hidden[%d]=%d", i, sp->hidden[i]);
  FC_REQUIRE(sp->base == FLOWCUDA_BASE_NORMAL || sp->base == FLOWCUDA_BASE_UNIFORM01, "made: unknown base %d", sp->base);
  return FLOWCUDA_OK;
}

static int made_log_prob_run(const flowcuda_made_spec* sp, const float* params, const float* x, float* out_lp,
                             float* out_u, float* out_ld, int64_t n, int with_base, void* workspace,
                             size_t workspace_bytes, cudaStream_t s, int mode = 0) {
  // mode 0: pack the weights, then run; 1: the workspace already holds this flow's weight images; 2: pack only
  MadeLayout L;
  if (!made_layout(sp, n, L)) {
    set_error("made: layer too wide for the packed K-stage table");
    return FLOWCUDA_EUNSUPPORTED;
  }
  if (!workspace || workspace_bytes < L.total) {
    set_error("made_log_prob: workspace %zu < %zu bytes", workspace_bytes, L.total);
    return FLOWCUDA_EWORKSPACE;
  }
  uint8_t* ws = (uint8_t*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  const int D = L.D, nh = L.nh;
  float* v = (float*)(ws + L.off_v);
  float* ld = (float*)(ws + L.off_ld);
  float* ldp = (float*)(ws + L.off_ldp);
  if (mode != 2) FC_CUDA(cudaMemcpyAsync(v, x, (size_t)n * D * sizeof(float), cudaMemcpyDeviceToDevice, s));
  for (int i = 0; i < L.nlin && mode != 1; ++i) {
    MadePackW p;
    p.params = params; p.per_layer = L.per_layer; p.w_off = L.w_off[i];
    p.k_in = L.sizes[i]; p.n_out = L.sizes[i + 1];
    p.H_in = i == 0 ? 0 : L.sizes[i];
    p.H_out = i < nh ? L.sizes[i + 1] : 0;
    p.D = D; p.KS = L.KS[i]; p.NT = L.NT[i]; p.img = ws + L.off_w[i];
    int blocks = (L.KS[i] * 4 * kDtBN + 255) / 256;
    made_tc_pack_w_kernel<<<dim3(blocks, L.NT[i], sp->num_layers), 256, 0, s>>>(p);
    FC_LAUNCH_CHECK("made_tc_pack_w_kernel");
  }
  if (mode == 2) return FLOWCUDA_OK;
  // the finish kernel sums every partial slot of every layer's slab; four-warp epilogue groups fill only two of a
  // tile's four quarter slots
  FC_CUDA(cudaMemsetAsync(ldp, 0, (size_t)sp->num_layers * n * 4 * L.tiles * sizeof(float), s));
  for (int it = 0; it < sp->num_layers; ++it) {
    const int l = sp->num_layers - 1 - it;  // data -> base: last layer first
    const bool rev = (l & 1) != 0;
    const float* lp = params + L.per_layer * l;
    // the previous layer's epilogue has already written this layer's activation image when it can
    const bool fuse_x = (D % 16) == 0;
    if (it == 0 || !fuse_x) {
      int64_t total = L.MT * 128 * L.KS[0] * 4;
      int blocks = (int)((total + 255) / 256);
      FC_CUDA(launch_chain(dense_tc_pack_x_kernel, dim3(blocks), dim3(256), 0, s, (const float*)v, n, D, rev ? D - 1 : 0,
                           rev ? -1 : 1, D, L.KS[0], ws + L.off_x));
    }
    const uint8_t* aimg = ws + L.off_x;
    for (int i = 0; i < L.nlin; ++i) {
      DtGemm g = {};
      g.a_img = aimg;
      g.b_img = ws + L.off_w[i] + (size_t)l * L.NT[i] * L.KS[i] * 2 * kDtBBytes;
      g.bias = lp + L.b_off[i];
      g.KS = L.KS[i];
      g.n_out = L.sizes[i + 1];
      g.v = v; g.ldp = ldp + (size_t)it * n * 4 * L.tiles; g.N = n; g.D = D; g.tp = 0; g.dstride = 1; g.ndims = D; g.dpt = 128; g.ppd = 2;
      g.tiles = 4 * L.tiles; g.made = 1; g.rev = rev ? 1 : 0;
      for (int t = 0; t < L.NT[i] && t < kDtMaxKsTiles; ++t) g.ks_tile[t] = (unsigned char)made_ks_needed(L, i, t, rev);
      int rc;
      const Dt2Launch dl = {L.MT, (int64_t)sp->num_layers * L.NT[i], (int64_t)l * L.NT[i], 1};  // single CTAs: measured faster here
      if (i < nh) {
        g.out_img = ws + L.off_h[i & 1];
        g.out_KS = L.KS[i + 1];
        rc = dt2_launch_gemm<kEpiHidden, false>(g, dl, L.MT, L.NT[i], s);
        aimg = g.out_img;
      } else {
        if (fuse_x && it + 1 < sp->num_layers) {
          g.out_img = ws + L.off_x;
          g.out_KS = L.KS[0];
          g.next_rev = ((l - 1) & 1) != 0;
        }
        rc = dt2_launch_gemm<kEpiAffine, true>(g, dl, L.MT, L.tiles, s);
        if (rc) return rc;
      }
      if (rc) return rc;
    }
  }
  (void)ld;
  if (out_lp || out_ld) {
    FC_CUDA(launch_chain(dense_tc_finish_kernel, dim3((unsigned)((n * 32 + 255) / 256)), dim3(256), 0, s, (const float*)v,
                         (const float*)ldp, (int)sp->num_layers, 4 * L.tiles, 4 * L.tiles, n, D, with_base, (int)sp->base, out_lp,
                         out_ld));
  }
  if (out_u) FC_CUDA(cudaMemcpyAsync(out_u, v, (size_t)n * D * sizeof(float), cudaMemcpyDeviceToDevice, s));
  return FLOWCUDA_OK;
}

// ---- sequential sampler -------------------------------------------------------------------------
struct MadeSample {
  const float* params;
  const float* z;
  float* x;
  int64_t N, per_layer;
  int64_t w_off[3], b_off[3];
  int D, H1, H2, L;
};

__device__ __forceinline__ void cp_async4(float* dst_smem, const float* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(tc::smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// S samples per CTA, T threads; thread t owns hidden columns t + T*c (c < CH) of both hidden layers and
// the (shift, log_scale) accumulators of ordering positions t + T*c (c < CD).
// Step q needs row q of W1 and the W2 / W3 rows of the hidden units whose degree completes at q; which
// rows is known in advance, so every thread streams ITS elements of those rows into a private
// shared-memory ring kMsPF steps ahead (cp.async), hiding the L2 latency behind the previous steps.
constexpr int kMsRing = 4, kMsPF = 3;

template <int S, int T, int CH, int CD>
__global__ void __launch_bounds__(T, S <= 4 ? 2 : 1) made_sample_kernel(MadeSample a) {
  extern __shared__ __align__(16) float sm[];
  const int D = a.D, H1 = a.H1, H2 = a.H2;
  constexpr int NC = 2 * CH + 2 * CD;  // ring components per thread and step
  float* zbuf = sm;            // [D][S] state, indexed by STATE dim
  float* hb1 = zbuf + D * S;   // [H1][S] finished hidden-1 activations
  float* hb2 = hb1 + H1 * S;   // [H2][S]
  float* ring = hb2 + H2 * S;  // [kMsRing][NC][T]
  // km1[m] / km2[m]: first hidden unit whose degree is >= m (the 64-bit division of made_kmin is far too
  // slow for the per-step path)
  int* km1 = reinterpret_cast<int*>(ring + kMsRing * NC * T);  // [D + 2]
  int* km2 = km1 + D + 2;
  const int tid = threadIdx.x;
  for (int m = tid; m < D + 2; m += T) {
    km1[m] = made_kmin(m, H1, D);
    km2[m] = made_kmin(m, H2, D);
  }
  const int64_t n0 = (int64_t)blockIdx.x * S;
  for (int i = tid; i < D * S; i += T) {
    const int s = i / D, d = i - s * D;
    const int64_t r = n0 + s;
    zbuf[d * S + s] = r < a.N ? a.z[r * D + d] : 0.f;
  }
  int deg1[CH], deg2[CH];
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    const int k = tid + T * c;
    deg1[c] = k < H1 ? made_hdeg(k, H1, D) : 0;
    deg2[c] = k < H2 ? made_hdeg(k, H2, D) : 0;
  }
  __syncthreads();

  for (int l = 0; l < a.L; ++l) {
    const float* P = a.params + a.per_layer * l;
    const float* __restrict__ W1 = P + a.w_off[0];
    const float* __restrict__ W2 = P + a.w_off[1];
    const float* __restrict__ W3 = P + a.w_off[2];
    const bool rev = (l & 1) != 0;
    // this thread's weights for step qq -> ring slot qq % kMsRing (always commits a group, possibly empty)
    auto issue = [&](int qq) {
      if (qq < D - 1) {
        float* rb = ring + (qq % kMsRing) * (NC * T) + tid;
        const int m = qq + 1;
#pragma unroll
        for (int c = 0; c < CH; ++c)
          if (tid + T * c < H1 && deg1[c] >= m) cp_async4(rb + c * T, W1 + (int64_t)qq * H1 + tid + T * c);
        const int k1 = km1[m];
        if (km1[m + 1] > k1) {
#pragma unroll
          for (int c = 0; c < CH; ++c)
            if (deg2[c] >= m) cp_async4(rb + (CH + c) * T, W2 + (int64_t)k1 * H2 + tid + T * c);
        }
        const int k2 = km2[m];
        if (km2[m + 1] > k2) {
#pragma unroll
          for (int c = 0; c < CD; ++c) {
            const int qd = tid + T * c;
            if (qd > qq && qd < D) {
              cp_async4(rb + (2 * CH + 2 * c) * T, W3 + (int64_t)k2 * 2 * D + qd);
              cp_async4(rb + (2 * CH + 2 * c + 1) * T, W3 + (int64_t)k2 * 2 * D + D + qd);
            }
          }
        }
      }
      cp_async_commit();
    };
#pragma unroll
    for (int i = 0; i < kMsPF; ++i) issue(i);

    float a1[CH][S], a2[CH][S], om[CD][S], oa[CD][S];
#pragma unroll
    for (int c = 0; c < CH; ++c) {
      const int k = tid + T * c;
      const float b1 = k < H1 ? P[a.b_off[0] + k] : 0.f, b2 = k < H2 ? P[a.b_off[1] + k] : 0.f;
#pragma unroll
      for (int s = 0; s < S; ++s) { a1[c][s] = b1; a2[c][s] = b2; }
    }
#pragma unroll
    for (int c = 0; c < CD; ++c) {
      const int q = tid + T * c;
      const float bm = q < D ? P[a.b_off[2] + q] : 0.f, ba = q < D ? P[a.b_off[2] + D + q] : 0.f;
#pragma unroll
      for (int s = 0; s < S; ++s) { om[c][s] = bm; oa[c][s] = ba; }
    }

    for (int q = 0; q < D; ++q) {
      const int m = q + 1;  // the degree that completes at this step
      const int pos = rev ? D - 1 - q : q;
      issue(q + kMsPF);
      // x_q = u_q * exp(log_scale_q) + shift_q  (owner of position q)
#pragma unroll
      for (int c = 0; c < CD; ++c) {
        if (tid + T * c == q) {
#pragma unroll
          for (int s = 0; s < S; ++s) zbuf[pos * S + s] = fmaf(zbuf[pos * S + s], expf(oa[c][s]), om[c][s]);
        }
      }
      cp_async_wait<kMsPF>();  // this thread's ring entries of step q have landed
      __syncthreads();
      if (q == D - 1) break;  // nothing downstream of the last position
      const float* rb = ring + (q % kMsRing) * (NC * T) + tid;
      float xs[S];
#pragma unroll
      for (int s = 0; s < S; s += 4) {
        const float4 t4 = *reinterpret_cast<const float4*>(zbuf + pos * S + s);
        xs[s] = t4.x; xs[s + 1] = t4.y; xs[s + 2] = t4.z; xs[s + 3] = t4.w;
      }
      // hidden 1: rank-1 update with row q of W1 (mask: degree q+1 <= deg1)
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        if (deg1[c] >= m) {
          const float w = rb[c * T];
#pragma unroll
          for (int s = 0; s < S; ++s) a1[c][s] = fmaf(xs[s], w, a1[c][s]);
          if (deg1[c] == m) {
#pragma unroll
            for (int s = 0; s < S; ++s) hb1[(tid + T * c) * S + s] = fmaxf(a1[c][s], 0.f);
          }
        }
      }
      // hidden 2: the hidden-1 units of degree m are final now
      const int k1lo = km1[m], k1hi = km1[m + 1];
      if (k1hi > k1lo) {
        __syncthreads();
        for (int ks = k1lo; ks < k1hi; ++ks) {
          float hs[S];
#pragma unroll
          for (int s = 0; s < S; s += 4) {
            const float4 t4 = *reinterpret_cast<const float4*>(hb1 + ks * S + s);
            hs[s] = t4.x; hs[s + 1] = t4.y; hs[s + 2] = t4.z; hs[s + 3] = t4.w;
          }
#pragma unroll
          for (int c = 0; c < CH; ++c) {
            if (deg2[c] >= m) {
              const float w = ks == k1lo ? rb[(CH + c) * T] : W2[(int64_t)ks * H2 + tid + T * c];
#pragma unroll
              for (int s = 0; s < S; ++s) a2[c][s] = fmaf(hs[s], w, a2[c][s]);
            }
          }
        }
      }
      const int k2lo = km2[m], k2hi = km2[m + 1];
      if (k2hi > k2lo) {
#pragma unroll
        for (int c = 0; c < CH; ++c) {
          if (deg2[c] == m) {
#pragma unroll
            for (int s = 0; s < S; ++s) hb2[(tid + T * c) * S + s] = fmaxf(a2[c][s], 0.f);
          }
        }
        __syncthreads();
        // outputs: positions q' > q see the hidden-2 units of degree m
        for (int js = k2lo; js < k2hi; ++js) {
          float gs[S];
#pragma unroll
          for (int s = 0; s < S; s += 4) {
            const float4 t4 = *reinterpret_cast<const float4*>(hb2 + js * S + s);
            gs[s] = t4.x; gs[s + 1] = t4.y; gs[s + 2] = t4.z; gs[s + 3] = t4.w;
          }
#pragma unroll
          for (int c = 0; c < CD; ++c) {
            const int qq = tid + T * c;
            if (qq > q && qq < D) {
              float wm, wa;
              if (js == k2lo) {
                wm = rb[(2 * CH + 2 * c) * T];
                wa = rb[(2 * CH + 2 * c + 1) * T];
              } else {
                wm = W3[(int64_t)js * 2 * D + qq];
                wa = W3[(int64_t)js * 2 * D + D + qq];
              }
#pragma unroll
              for (int s = 0; s < S; ++s) {
                om[c][s] = fmaf(gs[s], wm, om[c][s]);
                oa[c][s] = fmaf(gs[s], wa, oa[c][s]);
              }
            }
          }
        }
      }
    }
    cp_async_wait<0>();
    __syncthreads();
  }
  for (int i = tid; i < D * S; i += T) {
    const int s = i / D, d = i - s * D;
    const int64_t r = n0 + s;
    if (r < a.N) a.x[r * D + d] = zbuf[d * S + s];
  }
}

constexpr int kMsS = 16, kMsT = 512;

static int made_sample_run(const flowcuda_made_spec* sp, const float* params, const float* z, float* x, int64_t n,
                           cudaStream_t s) {
  MadeLayout L;
  made_layout(sp, 128, L);
  if (sp->num_hidden != 2) {
    set_error("made_sample: the sequential sampler is built for two hidden layers (got %d)", sp->num_hidden);
    return FLOWCUDA_EUNSUPPORTED;
  }
  const int D = sp->event_dim, H1 = sp->hidden[0], H2 = sp->hidden[1];
  if (H1 > kMsT || H2 > kMsT || D > 2 * kMsT) {
    set_error("made_sample: needs hidden <= %d and event_dim <= %d", kMsT, 2 * kMsT);
    return FLOWCUDA_EUNSUPPORTED;
  }
  MadeSample a;
  a.params = params; a.z = z; a.x = x; a.N = n; a.per_layer = L.per_layer;
  for (int i = 0; i < 3; ++i) { a.w_off[i] = L.w_off[i]; a.b_off[i] = L.b_off[i]; }
  a.D = D; a.H1 = H1; a.H2 = H2; a.L = sp->num_layers;
  const size_t smem = ((size_t)(D + H1 + H2) * kMsS + (size_t)kMsRing * (2 * 1 + 2 * 2) * kMsT + 2 * (size_t)(D + 2)) * sizeof(float);
  auto kern = made_sample_kernel<kMsS, kMsT, 1, 2>;
  FC_ENSURE_SMEM(kern, smem);
  const int64_t blocks = (n + kMsS - 1) / kMsS;
  kern<<<(unsigned)blocks, kMsT, smem, s>>>(a);
  FC_LAUNCH_CHECK("made_sample_kernel");
  return FLOWCUDA_OK;
}

}  // namespace flowcuda

using namespace flowcuda;

extern "C" {

int64_t flowcuda_made_param_count(const flowcuda_made_spec* spec) {
  if (made_check(spec)) return -1;
  MadeLayout L;
  made_layout(spec, 128, L);
  return L.per_layer * spec->num_layers;
}

size_t flowcuda_made_workspace_bytes(const flowcuda_made_spec* spec, int64_t n) {
  if (made_check(spec) || n < 0) return 0;
  MadeLayout L;
  if (!made_layout(spec, n, L)) return 0;
  return L.total;
}

int flowcuda_made_log_prob(const flowcuda_made_spec* spec, const float* params, const float* x, float* log_prob,
                           float* u, float* log_det, int64_t n, int with_base, void* workspace,
                           size_t workspace_bytes, void* stream) {
  int rc = made_check(spec);
  if (rc) return rc;
  FC_REQUIRE(n >= 0, "made_log_prob: negative n");
  if (n == 0) return FLOWCUDA_OK;
  FC_REQUIRE(params && x, "made_log_prob: NULL params or input");
  FC_REQUIRE(log_prob || u || log_det, "made_log_prob: no output requested");
  return made_log_prob_run(spec, params, x, log_prob, u, log_det, n, with_base, workspace, workspace_bytes,
                           (cudaStream_t)stream);
}

int flowcuda_made_pack_weights(const flowcuda_made_spec* spec, const float* params, void* workspace, size_t workspace_bytes,
                               void* stream) {
  int rc = made_check(spec);
  if (rc) return rc;
  FC_REQUIRE(params && workspace, "made_pack_weights: NULL params or workspace");
  return made_log_prob_run(spec, params, nullptr, nullptr, nullptr, nullptr, 0, 0, workspace, workspace_bytes,
                           (cudaStream_t)stream, 2);
}

int flowcuda_made_log_prob_packed(const flowcuda_made_spec* spec, const float* params, const float* x, float* log_prob,
                                  float* u, float* log_det, int64_t n, int with_base, void* workspace,
                                  size_t workspace_bytes, void* stream) {
  int rc = made_check(spec);
  if (rc) return rc;
  FC_REQUIRE(n >= 0, "made_log_prob_packed: negative n");
  if (n == 0) return FLOWCUDA_OK;
  FC_REQUIRE(params && x, "made_log_prob_packed: NULL params or input");
  FC_REQUIRE(log_prob || u || log_det, "made_log_prob_packed: no output requested");
  return made_log_prob_run(spec, params, x, log_prob, u, log_det, n, with_base, workspace, workspace_bytes,
                           (cudaStream_t)stream, 1);
}

int flowcuda_made_sample(const flowcuda_made_spec* spec, const float* params, const float* z, float* x, int64_t n,
                         void* stream) {
  int rc = made_check(spec);
  if (rc) return rc;
  FC_REQUIRE(n >= 0, "made_sample: negative n");
  if (n == 0) return FLOWCUDA_OK;
  FC_REQUIRE(params && z && x, "made_sample: NULL params, z or x");
  return made_sample_run(spec, params, z, x, n, (cudaStream_t)stream);
}

}  // extern "C"

#ifdef FLOWCUDA_TRACE
extern "C" int flowcuda_debug_set_dt_trace_made(unsigned long long* device_buf) {
  return cudaMemcpyToSymbol(flowcuda::g_dt_trace, &device_buf, sizeof(device_buf)) == cudaSuccess ? 0 : -2;
}
#endif
