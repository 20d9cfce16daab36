// made_grad.cu — rows #7/#8: value_and_grad of loss = -mean(log_prob(x)) for the MADE-conditioned MAF
// (the "train samples/s (MAF-MNIST)" step of BASELINE.json; jax.value_and_grad(loss_fn) of
// src/maf_mnist.py:134-166 applied to the north_star's MADE flow; oracle: nll_and_grads over
// oracle/flows_oracle.py:576-590).  Hand-written reverse mode, FP32 throughout:
//   forward  (per layer, stored): xp (layer input in the layer's ordering), h1, h2, o = [mu, alpha]
//   backward (per layer): dmu = -du·e^-alpha, dalpha = -du·u + 1/N, dx = du·e^-alpha + dh1·W1ᵀ,
//            dW_i = mask_i ⊙ (inᵀ·dout), db_i = colsum(dout), relu masks from the stored activations.
// Every product runs through ONE generic tiled SGEMM (strided operands, fixed-order K loop, no split-K
// atomics), so gradients are bit-reproducible run to run.  spec->precision selects the engine behind that
// descriptor: FP32 CUDA cores (mg_sgemm.cuh) or tcgen05 bf16x3 (mgtc.cuh); the stored activations are FP32
// either way.
#include "mgtc.cuh"

namespace flowcuda {

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
// flip_out: u is written with the dims reversed — the next layer's ordering (consecutive layers alternate)
__global__ void __launch_bounds__(256) mg_affine_fwd_kernel(const float* __restrict__ xp, const float* __restrict__ o,
                                                           float* __restrict__ u, float* __restrict__ ld, int64_t N, int D,
                                                           int first, int flip_out) {
  const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (r >= N) return;
  float s = 0.f;
  for (int q = lane; q < D; q += 32) {
    const float mu = o[r * 2 * D + q], al = o[r * 2 * D + D + q];
    u[r * D + (flip_out ? D - 1 - q : q)] = (xp[r * D + q] - mu) * expf(-al);
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

// backward affine: du = gin (read with the dims reversed when flip_in: gin is in the previous layer's ordering)
// -> dx_direct = du·e^-alpha, dO = [dmu, dalpha]
__global__ void __launch_bounds__(256) mg_affine_bwd_kernel(const float* __restrict__ xp, const float* __restrict__ o,
                                                           const float* __restrict__ gin, int flip_in, float* __restrict__ dx,
                                                           float* __restrict__ dO, int64_t N, int D) {
  const int64_t tot = N * D;
  const float invn = 1.f / (float)N;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / D; const int q = (int)(i - r * D);
    const float mu = o[r * 2 * D + q], al = o[r * 2 * D + D + q];
    const float e = expf(-al), g = gin[r * D + (flip_in ? D - 1 - q : q)];
    const float uu = (xp[i] - mu) * e;
    dO[r * 2 * D + q] = -g * e;
    dO[r * 2 * D + D + q] = fmaf(-g, uu, invn);  // d(-mean ld)/dalpha = +1/N
    dx[i] = g * e;
  }
}



// ---- host side ------------------------------------------------------------------------------------
struct MgLayout {
  int D, H1, H2, L;
  int64_t w1, b1, w2, b2, w3, b3, per_layer;
  size_t per_layer_act;  // floats of stored activations per layer
  size_t off_state, off_tmp, off_ld, off_lp, off_dO, off_dH2, off_dH1, off_sum, off_deg, off_part, off_act, total;  // bytes
  int tc;                       // tensor-core engine
  size_t off_aimg, off_bimg, a_bytes, b_bytes, part_floats;
  // tensor-core engine: the weight-gradient jobs run on the library's side stream (common.cuh) with their own scratch;
  // dO / dH2 / dH1 are doubled by flow-layer parity
  size_t off_dO2, off_dH2b, off_dH1b, off_aimg2, off_bimg2, off_part2;
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
    L.tc = sp->precision == FLOWCUDA_PRECISION_BF16X3;
    L.part_floats = mx * (L.tc ? kMgTcSplits : kMgSplits);
    size_t wmax = 2 * (size_t)L.D;  // the column sums borrow the buffer: kMgColSlices doubles per column
    if ((size_t)L.H1 > wmax) wmax = L.H1;
    if ((size_t)L.H2 > wmax) wmax = L.H2;
    if (L.part_floats < 2 * kMgColSlices * wmax) L.part_floats = 2 * kMgColSlices * wmax;
    L.off_part = off; off = mg_al(off + L.part_floats * 4);
  }
  L.off_aimg = L.off_bimg = off; L.a_bytes = L.b_bytes = 0;
  if (L.tc) {
    const int64_t D = L.D, H1 = L.H1, H2 = L.H2;
    // A operands: activations / their gradients [n, K], and transposed activations [units, n]
    const int64_t am[] = {n, n, n, n, n, n, H2, H1, D}, ak[] = {D, H1, H2, 2 * D, H2, H1, n, n, n};
    // B operands: weights both ways, and transposed output gradients [cols, n]
    const int64_t bn[] = {H1, H2, 2 * D, H2, H1, D, 2 * D, H2, H1}, bk[] = {D, H1, H2, 2 * D, H2, H1, n, n, n};
    for (int i = 0; i < 9; ++i) {
      if (mgtc_a_bytes(am[i], ak[i]) > L.a_bytes) L.a_bytes = mgtc_a_bytes(am[i], ak[i]);
      if (mgtc_b_bytes(bn[i], bk[i]) > L.b_bytes) L.b_bytes = mgtc_b_bytes(bn[i], bk[i]);
    }
    L.off_aimg = off; off = mg_al(off + L.a_bytes);
    L.off_bimg = off; off = mg_al(off + L.b_bytes);
  }
  L.off_dO2 = L.off_dO; L.off_dH2b = L.off_dH2; L.off_dH1b = L.off_dH1;
  L.off_aimg2 = L.off_aimg; L.off_bimg2 = L.off_bimg; L.off_part2 = L.off_part;
  if (L.tc) {
    L.off_dO2 = off; off = mg_al(off + (size_t)n * 2 * L.D * 4);
    L.off_dH2b = off; off = mg_al(off + (size_t)n * L.H2 * 4);
    L.off_dH1b = off; off = mg_al(off + (size_t)n * L.H1 * 4);
    L.off_aimg2 = off; off = mg_al(off + L.a_bytes);
    L.off_bimg2 = off; off = mg_al(off + L.b_bytes);
    L.off_part2 = off; off = mg_al(off + L.part_floats * 4);
  }
  L.off_act = off; off = mg_al(off + L.per_layer_act * 4 * (size_t)L.L);
  L.total = off + 256;
  return true;
}


// fwd_only: stop after the density pass and copy log_prob[N] (and, if asked, the base-space point u[N, D]) out —
// flowcuda_made_log_prob_fp32, the FP32 partner of the tensor-core density pass (made.cu)
static int mg_run(const flowcuda_made_spec* sp, const float* params, const float* x, int64_t n, float* loss, float* grad,
                  void* workspace, size_t wsb, cudaStream_t s, bool fwd_only = false, float* out_lp = nullptr,
                  float* out_u = nullptr) {
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
  const MgTcScratch sc = {ws + L.off_aimg, L.a_bytes, ws + L.off_bimg, L.b_bytes, part, L.part_floats};
  auto mg_launch = [&](const MgGemm& g, cudaStream_t st) { return L.tc ? mgtc_launch(g, sc, st) : flowcuda::mg_launch(g, st); };
  FC_CUDA(cudaMemcpyAsync(S, x, (size_t)n * D * 4, cudaMemcpyDeviceToDevice, s));
  // ---- forward, data -> base: layer L-1 first
  for (int it = 0; it < L.L; ++it) {
    const int l = L.L - 1 - it, rev = l & 1;
    const float* P = params + L.per_layer * l;
    if (it == 0) mg_copy_perm_kernel<<<eb, 256, 0, s>>>(S, XP(l), n, D, rev, 0);  // later layers: written by the affine kernel
    MgGemm g = mg_gemm(XP(l), D, 1, P + L.w1, H1, 1, HA(l), H1, N, H1, D);
    g.bias = P + L.b1; g.relu = 1; g.mask = m0; g.mask_on_b = 1;
    if ((rc = mg_launch(g, s))) return rc;
    g = mg_gemm(HA(l), H1, 1, P + L.w2, H2, 1, HB(l), H2, N, H2, H1);
    g.bias = P + L.b2; g.relu = 1; g.mask = m1; g.mask_on_b = 1;
    if ((rc = mg_launch(g, s))) return rc;
    g = mg_gemm(HB(l), H2, 1, P + L.w3, 2 * D, 1, OO(l), 2 * D, N, 2 * D, H2);
    g.bias = P + L.b3; g.mask = m2; g.mask_on_b = 1;
    if ((rc = mg_launch(g, s))) return rc;
    // u goes straight into the next layer's input block, in that layer's ordering (the reverse of this one);
    // layer 0 sees the natural ordering, so the base point lands in S as is
    mg_affine_fwd_kernel<<<rb, 256, 0, s>>>(XP(l), OO(l), l > 0 ? XP(l - 1) : S, LD, n, D, it == 0, l > 0);
    FC_LAUNCH_CHECK("made_nll_grad forward");
  }
  mg_finish_kernel<<<rb, 256, 0, s>>>(S, LD, LP, T, n, D);  // T = d loss / d base point (natural order)
  FC_LAUNCH_CHECK("mg_finish_kernel");
  if (fwd_only) {
    if (out_lp) FC_CUDA(cudaMemcpyAsync(out_lp, LP, (size_t)n * 4, cudaMemcpyDeviceToDevice, s));
    if (out_u) FC_CUDA(cudaMemcpyAsync(out_u, S, (size_t)n * D * 4, cudaMemcpyDeviceToDevice, s));
    return FLOWCUDA_OK;
  }
  if ((rc = flowcuda_sum_scaled(LP, n, -1.0 / (double)n, loss, ws + L.off_sum, s))) return rc;
  // ---- backward: layer 0 first.  Main stream: affine adjoint -> input-gradient GEMMs (the dependency chain); the
  // weight / bias gradients — nothing in the step depends on them — go to the side stream (tensor-core engine), as in
  // coupling_dense_grad.cu: same kernels and summation orders, bit-identical results.
  SideCtx* sd = nullptr;
  static const bool overlap_env = side_env_enabled("FLOWCUDA_TRAIN_OVERLAP");
  const bool overlap = L.tc && overlap_env;
  std::unique_lock<std::recursive_mutex> side_lock(side_mutex(), std::defer_lock);
  if (overlap) {
    side_lock.lock();
    if ((rc = side_ctx_get(&sd))) return rc;
  }
  cudaStream_t q = overlap ? sd->q : s;
  const MgTcScratch sc2 = {ws + L.off_aimg2, L.a_bytes, ws + L.off_bimg2, L.b_bytes, (float*)(ws + L.off_part2), L.part_floats};
  float* part_q = overlap ? (float*)(ws + L.off_part2) : part;
  auto side_launch = [&](const MgGemm& g) { return L.tc ? mgtc_launch(g, overlap ? sc2 : sc, q) : flowcuda::mg_launch(g, q); };
  auto fork = [&]() -> int {  // what the main stream has enqueued so far is ordered before the next side job
    if (!overlap) return FLOWCUDA_OK;
    FC_CUDA(cudaEventRecord(sd->ev_m, s));
    FC_CUDA(cudaStreamWaitEvent(q, sd->ev_m, 0));
    return FLOWCUDA_OK;
  };
  float* const S0 = S; float* const T0 = T;
  for (int l = 0; l < L.L; ++l) {
    const float* P = params + L.per_layer * l;
    float* G = grad + L.per_layer * l;
    const int lp = l & 1;
    float* dOl = overlap && lp ? (float*)(ws + L.off_dO2) : dO;
    float* dH2l = overlap && lp ? (float*)(ws + L.off_dH2b) : dH2;
    float* dH1l = overlap && lp ? (float*)(ws + L.off_dH1b) : dH1;
    if (overlap && l >= 2) FC_CUDA(cudaStreamWaitEvent(s, sd->ev_s[lp], 0));  // the side jobs of layer l-2 read these buffers
    // the gradient travels in the ordering of the layer that produced it: T (natural = layer 0's) first, then
    // the two buffers alternate and every layer reads its input gradient reversed
    float* gin = (l & 1) ? S0 : T0;
    float* S = (l & 1) ? T0 : S0;
    mg_affine_bwd_kernel<<<eb, 256, 0, s>>>(XP(l), OO(l), gin, l > 0, S, dOl, n, D);
    FC_LAUNCH_CHECK("mg_affine_bwd_kernel");
    // output Linear
    if ((rc = fork())) return rc;
    MgGemm g = mg_gemm(HB(l), 1, H2, dOl, 2 * D, 1, G + L.w3, 2 * D, H2, 2 * D, n);
    g.mask = m2; g.splits = splits; g.part = part_q;
    if ((rc = side_launch(g))) return rc;
    if ((rc = mg_colsum(dOl, n, 2 * D, G + L.b3, part_q, q))) return rc;
    g = mg_gemm(dOl, 2 * D, 1, P + L.w3, 1, 2 * D, dH2l, H2, N, H2, 2 * D);
    g.gate = HB(l); g.ldg = H2; g.mask = m2; g.mask_on_b = 2;
    if ((rc = mg_launch(g, s))) return rc;
    // hidden 2
    if ((rc = fork())) return rc;
    g = mg_gemm(HA(l), 1, H1, dH2l, H2, 1, G + L.w2, H2, H1, H2, n);
    g.mask = m1; g.splits = splits; g.part = part_q;
    if ((rc = side_launch(g))) return rc;
    if ((rc = mg_colsum(dH2l, n, H2, G + L.b2, part_q, q))) return rc;
    g = mg_gemm(dH2l, H2, 1, P + L.w2, 1, H2, dH1l, H1, N, H1, H2);
    g.gate = HA(l); g.ldg = H1; g.mask = m1; g.mask_on_b = 2;
    if ((rc = mg_launch(g, s))) return rc;
    // hidden 1
    if ((rc = fork())) return rc;
    g = mg_gemm(XP(l), 1, D, dH1l, H1, 1, G + L.w1, H1, D, H1, n);
    g.mask = m0; g.splits = splits; g.part = part_q;
    if ((rc = side_launch(g))) return rc;
    if ((rc = mg_colsum(dH1l, n, H1, G + L.b1, part_q, q))) return rc;
    if (overlap) FC_CUDA(cudaEventRecord(sd->ev_s[lp], q));
    g = mg_gemm(dH1l, H1, 1, P + L.w1, 1, H1, S, D, N, D, H1);
    g.accumulate = 1; g.mask = m0; g.mask_on_b = 2;
    if ((rc = mg_launch(g, s))) return rc;
    FC_LAUNCH_CHECK("made_nll_grad backward");
  }
  if (overlap) {  // join
    FC_CUDA(cudaEventRecord(sd->ev_m, q));
    FC_CUDA(cudaStreamWaitEvent(s, sd->ev_m, 0));
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

size_t flowcuda_made_log_prob_fp32_workspace_bytes(const flowcuda_made_spec* spec, int64_t n) {
  if (!spec) return 0;
  flowcuda_made_spec sp = *spec;
  sp.precision = FLOWCUDA_PRECISION_FP32;
  return flowcuda_made_nll_grad_workspace_bytes(&sp, n);
}

int flowcuda_made_log_prob_fp32(const flowcuda_made_spec* spec, const float* params, const float* x, float* log_prob, float* u,
                                int64_t n, void* workspace, size_t workspace_bytes, void* stream) {
  FC_REQUIRE(spec && params && x && (log_prob || u), "made_log_prob_fp32: NULL pointer or no output requested");
  FC_REQUIRE(n >= 1 && n < (1ll << 31), "made_log_prob_fp32: n=%lld outside [1, 2^31)", (long long)n);
  FC_REQUIRE(spec->event_dim >= 1 && spec->num_layers >= 1, "made_log_prob_fp32: bad spec");
  FC_REQUIRE(spec->base == FLOWCUDA_BASE_NORMAL, "made_log_prob_fp32: Normal base only (as made_nll_grad)");
  flowcuda_made_spec sp = *spec;
  sp.precision = FLOWCUDA_PRECISION_FP32;  // every product on the FP32 CUDA-core engine (mg_sgemm.cuh)
  return mg_run(&sp, params, x, n, nullptr, nullptr, workspace, workspace_bytes, (cudaStream_t)stream, true, log_prob, u);
}

int flowcuda_made_nll_grad(const flowcuda_made_spec* spec, const float* params, const float* x, int64_t n, float* loss,
                           float* grad, void* workspace, size_t workspace_bytes, void* stream) {
  FC_REQUIRE(spec && params && x && loss && grad, "made_nll_grad: NULL pointer");
  FC_REQUIRE(n >= 1 && n < (1ll << 31), "made_nll_grad: n=%lld outside [1, 2^31)", (long long)n);
  FC_REQUIRE(spec->event_dim >= 1 && spec->num_layers >= 1, "made_nll_grad: bad spec");
  FC_REQUIRE(spec->event_dim <= 32768 && spec->hidden[0] <= 32768 && spec->hidden[1] <= 32768, "made_nll_grad: sizes above 32768");
  FC_REQUIRE(spec->precision == FLOWCUDA_PRECISION_FP32 || spec->precision == FLOWCUDA_PRECISION_BF16X3,
             "made_nll_grad: precision %d (FP32 = 0 or BF16X3 = 3)", spec->precision);
  return mg_run(spec, params, x, n, loss, grad, workspace, workspace_bytes, (cudaStream_t)stream);
}

}  // extern "C"
