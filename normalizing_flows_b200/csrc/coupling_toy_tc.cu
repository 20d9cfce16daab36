// coupling_toy_tc.cu — rows #2/#3/#6 at event_dim = 2 on the 5th-gen tensor cores
// (FLOWCUDA_PRECISION_3XTF32): the whole distrax MaskedCoupling stack of
// src/real_nvp_toy_densities.py:74-110 / src/maf_toy_densities.py:73-111 in ONE persistent
// kernel, the two hidden Linears of every conditioner as tcgen05.mma in error-compensated 3-pass form
// (hi·hi + hi·lo + lo·hi, fp32 accumulate in TMEM).
//
// Operand format: fp16 hi/lo pairs under kind::f16.  fp16 carries the same 11 significant bits as tf32, so
// the three passes give the same ~2^-21 relative products as a tf32 split — but an operand element takes
// 16 bits, which (a) halves the A operand's TMEM footprint: a tile needs 32 (A hi) + 32 (A lo) + 64 (D)
// = 128 columns, so FOUR tiles are resident in the 512 columns instead of two and every sample warpgroup
// owns its TMEM slot outright (no slot hand-off between warpgroups: in the tf32 version 30 % of the sample
// warps' time was spent waiting for the partner to drain the shared slot); (b) doubles the MMA K per
// instruction (12 MMAs per GEMM instead of 24); (c) halves the weight images in shared memory.
// fp16's narrow exponent is handled by scaling: each layer's W2 / W3 are multiplied by a power of two that
// puts max|W| in [2^13, 2^14) before the split (undone exactly in the epilogue's FMA); activations are relu
// outputs, saturated at 65504 (fp16 max — far outside anything a flow's hidden layer produces) by the packing
// conversion itself, which also applies the relu (F2FP.SATFINITE.RELU); lo parts that
// fall into fp16's subnormal range lose bits below 2^-25 ABSOLUTE, which is under fp32's own rounding of
// the sums they enter.
//
// CTA = 4 sample warpgroups + 4 MMA-issuer warps (one per warpgroup) + 1 weight-loader warp, one CTA per SM.
// A warpgroup owns a 128-sample tile: thread = sample = TMEM lane.  Per flow layer
//   L1  (2->64, only the visible coordinate contributes): CUDA cores, result split into fp16
//       hi/lo and written straight into TMEM as the A operand (tcgen05.st) — activations never
//       touch shared or global memory;
//   L2  (64->64): 12 MMAs M128 N64 K16, A from TMEM, W2 (hi/lo tile images) from shared memory;
//   epilogue 2: tcgen05.ld, scale, +bias, relu, split, tcgen05.st back as the next A operand;
//   L3  (64->needed columns only, padded to N3): 12 MMAs M128 N{48,32,16};
//   epilogue 3: tcgen05.ld of the 3K+1 spline (or 2 affine) parameters of the transformed
//       coordinate, inner bijector, log-det accumulation.
// The four warpgroups' tiles interleave on the tensor pipe (one computes its epilogue while the
// others' MMAs run).  Weights of a layer are one contiguous pre-packed image fetched by a single
// cp.async.bulk (TMA) into a double-buffered slot and shared by T tile quads whose per-sample
// state (x0, x1, log-det) waits in shared memory, so a chunk of T*512 samples reads each layer's
// weights once.  HBM traffic: x in, log_prob (and/or z) out.
#include <cuda_fp16.h>

#include "tc.cuh"
#include "transforms.cuh"

namespace flowcuda {

// warps 0-15: WG0-3 (WG g owns TMEM slot g), 16-19: their MMA issuers; issuer 0 also fetches the weight images.
// 20 warps = 5 per scheduler: a 21st (a separate loader warp) put six on one scheduler's 16 K-register file and
// capped the kernel at 80 registers per thread (spills in the epilogues); 640 threads allow 96.
constexpr int kTcThreads = 4 * 128 + 128;
constexpr uint32_t kFmtF16 = 0;            // kind::f16 operand format code of fp16 (bf16 = 1)
constexpr int kTcMaxT = 8;       // tile pairs per chunk

template <int N3>
struct ToyTcImg {  // per-layer packed image (bytes)
  static constexpr int w2 = 64 * 64 * 2;                  // one of hi / lo (fp16)
  static constexpr int w3 = N3 * 64 * 2;
  static constexpr int off_w2hi = 0, off_w2lo = w2, off_w3hi = 2 * w2, off_w3lo = 2 * w2 + w3;
  static constexpr int wbytes = 2 * w2 + 2 * w3;          // the TMA-fetched part
  static constexpr int small_floats = 192 + N3 + 4;       // w1(visible row)[64] b1[64] b2[64] b3[N3] {1/s2, 1/s3, log2e/s3, -}
  static constexpr int bytes = (wbytes + small_floats * 4 + 127) & ~127;
};

__host__ __device__ constexpr int toy_need(int TR, int KT) { return TR == FLOWCUDA_TRANSFORM_RQS ? 3 * KT + 1 : 2; }
__host__ __device__ constexpr int toy_n3(int TR, int KT) { return (toy_need(TR, KT) + 15) & ~15; }

struct ToyTcArgs {
  const uint8_t* img;  // [num_layers][Img::bytes]
  const float* in;     // [N,2]
  float* out_v;        // [N,2] or NULL
  float* out_lp;       // [N] or NULL
  int64_t n;
  int layer_begin, layer_end;
  int with_base, base;
  float lo, hi;
  int T;               // tile pairs per chunk
  float koff[13];      // RQS: lo + i * min_bin_size (read from the constant bank by rqs_apply_mufu)
};

// params (haiku order) -> per-layer image.  One CTA per flow layer.
template <int TR, int KT>
__global__ void __launch_bounds__(256) toy_tc_pack_kernel(const float* __restrict__ params, int h1, int h2, int of,
                                                          uint8_t* __restrict__ img) {
  constexpr int NEED = toy_need(TR, KT), N3 = toy_n3(TR, KT);
  using I = ToyTcImg<N3>;
  const int l = blockIdx.x;
  const size_t per_layer = (size_t)2 * h1 + h1 + (size_t)h1 * h2 + h2 + (size_t)h2 * of + of;
  const float* p = params + per_layer * l;
  const float* pW1 = p;
  const float* pb1 = pW1 + 2 * h1;
  const float* pW2 = pb1 + h1;
  const float* pb2 = pW2 + (size_t)h1 * h2;
  const float* pW3 = pb2 + h2;
  const float* pb3 = pW3 + (size_t)h2 * of;
  uint8_t* out = img + (size_t)l * I::bytes;
  const int t = l & 1;  // transformed coordinate; the visible one is 1 - t
  const int col0 = TR == FLOWCUDA_TRANSFORM_AFFINE_SCALAR ? 0 : t * NEED;
  // power-of-two scales that put max|W| in [2^13, 2^14): the fp16 lo parts stay in the normal range
  __shared__ int smax[2];
  if (threadIdx.x < 2) smax[threadIdx.x] = 0;
  __syncthreads();
  float m2 = 0.f, m3 = 0.f;
  for (int i = threadIdx.x; i < h1 * h2; i += blockDim.x) m2 = fmaxf(m2, fabsf(pW2[i]));
  for (int i = threadIdx.x; i < h2 * NEED; i += blockDim.x) m3 = fmaxf(m3, fabsf(pW3[(size_t)(i / NEED) * of + col0 + i % NEED]));
  atomicMax(&smax[0], __float_as_int(m2));  // non-negative floats order like their bit patterns
  atomicMax(&smax[1], __float_as_int(m3));
  __syncthreads();
  m2 = __int_as_float(smax[0]);
  m3 = __int_as_float(smax[1]);
  const int e2 = (m2 > 0.f && isfinite(m2)) ? ilogbf(m2) : 13, e3 = (m3 > 0.f && isfinite(m3)) ? ilogbf(m3) : 13;
  const float s2 = ldexpf(1.f, 13 - e2), s3 = ldexpf(1.f, 13 - e3);
  auto split8 = [](const float* w, uint4& hi, uint4& lo) {
    __half2 h[4], l[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      h[e] = __floats2half2_rn(w[2 * e], w[2 * e + 1]);  // element 2e in the low half
      l[e] = __floats2half2_rn(w[2 * e] - __low2float(h[e]), w[2 * e + 1] - __high2float(h[e]));
    }
    hi = *reinterpret_cast<uint4*>(h);
    lo = *reinterpret_cast<uint4*>(l);
  };
  for (int i = threadIdx.x; i < 64 * 8; i += blockDim.x) {  // W2: rows n (out), 8 K chunks of 8
    const int n = i & 63, kc = i >> 6;
    float w[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int k = kc * 8 + e;
      w[e] = (k < h1 && n < h2) ? pW2[(size_t)k * h2 + n] * s2 : 0.f;
    }
    uint4 hi, lo;
    split8(w, hi, lo);
    const uint32_t off = tc::img_off(64, n, kc);
    *reinterpret_cast<uint4*>(out + I::off_w2hi + off) = hi;
    *reinterpret_cast<uint4*>(out + I::off_w2lo + off) = lo;
  }
  for (int i = threadIdx.x; i < N3 * 8; i += blockDim.x) {  // W3: needed columns only
    const int n = i % N3, kc = i / N3;
    float w[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int k = kc * 8 + e;
      w[e] = (k < h2 && n < NEED) ? pW3[(size_t)k * of + col0 + n] * s3 : 0.f;
    }
    uint4 hi, lo;
    split8(w, hi, lo);
    const uint32_t off = tc::img_off(N3, n, kc);
    *reinterpret_cast<uint4*>(out + I::off_w3hi + off) = hi;
    *reinterpret_cast<uint4*>(out + I::off_w3lo + off) = lo;
  }
  float* sm = reinterpret_cast<float*>(out + I::wbytes);
  for (int i = threadIdx.x; i < I::small_floats; i += blockDim.x) {
    float v = 0.f;
    if (i < 64) v = i < h1 ? pW1[(1 - t) * h1 + i] : 0.f;
    else if (i < 128) v = (i - 64) < h1 ? pb1[i - 64] : 0.f;
    else if (i < 192) v = (i - 128) < h2 ? pb2[i - 128] : 0.f;
    // RQS: the 2K bin logits are delivered in log2 units (rqs_apply_mufu), so their bias and unscale carry log2(e)
    else if (i < 192 + N3) v = (i - 192) < NEED ? pb3[col0 + i - 192] * ((TR == FLOWCUDA_TRANSFORM_RQS && (i - 192) < 2 * KT) ? kLog2e : 1.f) : 0.f;
    else if (i == 192 + N3) v = 1.f / s2;
    else if (i == 192 + N3 + 1) v = 1.f / s3;
    else if (i == 192 + N3 + 2) v = kLog2e / s3;
    sm[i] = v;
  }
}

// load NEED consecutive TMEM columns of this thread's lane into o[]
template <int NEED, bool WAIT = true>
__device__ __forceinline__ void tmem_load_cols(uint32_t taddr, uint32_t* o) {
  constexpr int c32 = NEED >= 32 ? 32 : 0;
  constexpr int c16 = (NEED - c32) >= 16 ? 16 : 0;
  constexpr int c8 = (NEED - c32 - c16) >= 8 ? 8 : 0;
  constexpr int c4 = (NEED - c32 - c16 - c8) >= 4 ? 4 : 0;
  constexpr int c2 = (NEED - c32 - c16 - c8 - c4) >= 2 ? 2 : 0;
  constexpr int c1 = NEED - c32 - c16 - c8 - c4 - c2;
  static_assert(NEED < 64 && c1 <= 1, "unsupported column count");
  if (c32) tc::tmem_ld_x32(taddr, o);
  if (c16) tc::tmem_ld_x16(taddr + c32, o + c32);
  if (c8) tc::tmem_ld_x8(taddr + c32 + c16, o + c32 + c16);
  if (c4) tc::tmem_ld_x4(taddr + c32 + c16 + c8, o + c32 + c16 + c8);
  if (c2) tc::tmem_ld_x2(taddr + c32 + c16 + c8 + c4, o + c32 + c16 + c8 + c4);
  if (c1) tc::tmem_ld_x1(taddr + c32 + c16 + c8 + c4 + c2, o + c32 + c16 + c8 + c4 + c2);
  if (WAIT) tc::tmem_ld_wait();
}

// relu + fp16 hi/lo split of two PRE-activations in six instructions (element 0 in the low half).
// hi = the value with its low 13 mantissa bits cleared (11 significant bits: exactly an fp16 value inside the
// normal range), lo = rn(x - hi).  The relu and the clamp at fp16's largest finite value ride on the packing
// conversion (F2FP.SATFINITE.RELU): for x < 0 the truncated hi is <= 0 and x - hi <= 0, so both halves become 0.
__device__ __forceinline__ void relu_split_f16x2(float x0, float x1, uint32_t& hi, uint32_t& lo) {
  const float h0 = __uint_as_float(__float_as_uint(x0) & 0xFFFFE000u), h1 = __uint_as_float(__float_as_uint(x1) & 0xFFFFE000u);
  asm("cvt.rn.relu.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(hi) : "f"(h1), "f"(h0));
  asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(lo) : "f"(x1 - h1), "f"(x0 - h0));
}

// The same on a packed pair (one 64-bit register pair holding two fp32): the pre-activation FMA and the
// x - hi subtraction (as hi * -1 + x) are single f32x2 instructions, so a pair costs FFMA2 + 2 LOP3 + FFMA2 + 2 F2FP.
__device__ __forceinline__ void relu_split_f16x2(f2_t x, f2_t minus1, uint32_t& hi, uint32_t& lo) {
  const f2_t h = x & 0xFFFFE000FFFFE000ull;
  const f2_t l = f2_fma(h, minus1, x);
  float h0, h1, l0, l1;
  f2_unpack(h, h0, h1);
  f2_unpack(l, l0, l1);
  asm("cvt.rn.relu.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(hi) : "f"(h1), "f"(h0));
  asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(lo) : "f"(l1), "f"(l0));
}

template <int TR, int KT, bool INV>
__global__ void __launch_bounds__(kTcThreads, 1) coupling_toy_tc_kernel(ToyTcArgs a) {
  constexpr int NEED = toy_need(TR, KT), N3 = toy_n3(TR, KT);
  using I = ToyTcImg<N3>;
  constexpr int SF = I::small_floats;
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nl = a.layer_end - a.layer_begin;
  const int T = a.T;
  // ---- shared memory carve-up
  uint8_t* wbuf = smem;                                             // [2][I::wbytes]
  float* smalls = reinterpret_cast<float*>(wbuf + 2 * I::wbytes);   // [nl][SF]
  float* sx0 = smalls + ((nl * SF + 3) & ~3);                       // [T*256] x 3
  float* sx1 = sx0 + T * 512;
  float* sld = sx1 + T * 512;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sld + T * 512);
  uint64_t* a_ready = bars;       // [4] count 128: warpgroup g's A operand is in TMEM
  uint64_t* d_full = bars + 4;    // [4] count 1: the MMAs of warpgroup g's request completed
  uint64_t* w_full = bars + 8;    // [2] count 1 + tx: weight image landed
  uint64_t* w_free = bars + 10;   // [2] count 4: every issuer's MMAs reading the weight slot completed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);

  if (warp == 16) tc::tmem_alloc(tmem_slot, 512);
  if (tid == 0) {
    for (int g = 0; g < 4; ++g) {
      tc::mbar_init(&a_ready[g], 128);
      tc::mbar_init(&d_full[g], 1);
    }
    for (int b = 0; b < 2; ++b) {
      tc::mbar_init(&w_full[b], 1);
      tc::mbar_init(&w_free[b], 4);
    }
    tc::fence_mbar_init();
  }
  for (int i = tid; i < nl * SF; i += kTcThreads) {
    const int it = i / SF;
    const int l = INV ? a.layer_begin + it : a.layer_end - 1 - it;
    smalls[i] = reinterpret_cast<const float*>(a.img + (size_t)l * I::bytes + I::wbytes)[i - it * SF];
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  const int64_t chunk_samples = (int64_t)T * 512;
  const int64_t nchunks = (a.n + chunk_samples - 1) / chunk_samples;
  const int my_chunks = blockIdx.x < nchunks ? (int)((nchunks - 1 - blockIdx.x) / gridDim.x + 1) : 0;
  const int Q = my_chunks * nl;  // layer steps this CTA runs

  if (warp >= 16) {
    // ===== MMA issuers: warp 16+g serves warpgroup g (blocking waits, no polling) =====
    if (lane == 0) {
      constexpr uint32_t idesc2 = tc::make_idesc(kFmtF16, 128, 64);
      constexpr uint32_t idesc3 = tc::make_idesc(kFmtF16, 128, N3);
      const int g = warp - 16;  // warpgroup / TMEM slot served
      const uint32_t tb = tmem + (uint32_t)g * 128u;  // A hi [0,32) A lo [32,64) D [64,128)
      uint32_t r = 0;
      // weight images: one cp.async.bulk per layer step into slot q&1, issued by issuer 0 as soon as every issuer's
      // MMAs on that slot (two steps earlier) have completed — probed without blocking between its own MMA batches,
      // waited for only when the step that needs the image begins
      int q_loaded = 0;
      auto load_next = [&](bool block) {
        if (q_loaded >= Q) return;
        const int it = q_loaded % nl, b = q_loaded & 1;
        if (q_loaded >= 2) {
          const uint32_t par = (uint32_t)(((q_loaded >> 1) - 1) & 1);
          if (block) tc::mbar_wait(&w_free[b], par);
          else if (!tc::mbar_test_wait(&w_free[b], par)) return;
        }
        const int l = INV ? a.layer_begin + it : a.layer_end - 1 - it;
        tc::mbar_arrive_expect_tx(&w_full[b], I::wbytes);
        tc::tma_bulk_g2s(wbuf + b * I::wbytes, a.img + (size_t)l * I::bytes, I::wbytes, &w_full[b]);
        ++q_loaded;
      };
      if (g == 0) {
        load_next(true);
        load_next(true);
      }
      for (int q = 0; q < Q; ++q) {
        const int b = q & 1;
        const uint32_t wb = tc::smem_u32(wbuf + b * I::wbytes);
        for (int t = 0; t < T; ++t) {
#pragma unroll 1
          for (int p = 0; p < 2; ++p, ++r) {
            if (g == 0) {
              while (q_loaded <= q) load_next(true);
              load_next(false);
            }
            tc::mbar_wait(&a_ready[g], r & 1);
            if (t == 0 && p == 0) tc::mbar_wait(&w_full[b], (uint32_t)((q >> 1) & 1));
            tc::tc_fence_after();
            const uint32_t bh = wb + (p ? I::off_w3hi : I::off_w2hi), bl = wb + (p ? I::off_w3lo : I::off_w2lo);
            const uint32_t lbo = (uint32_t)((p ? N3 : 64) >> 3) * 128u;
            const uint32_t idesc = p ? idesc3 : idesc2;
            uint32_t acc = 0;
#pragma unroll 1
            for (int pass = 0; pass < 3; ++pass) {
              const uint32_t at = tb + (pass == 2 ? 32u : 0u);
              // descriptor of K step s = descriptor of step 0 + (s * 2 * LBO) >> 4 in the address field
              const uint64_t d0 = tc::make_smem_desc(pass == 1 ? bl : bh, lbo, 128u);
              const uint64_t dstep = (uint64_t)((2u * lbo) >> 4);
#pragma unroll
              for (int s = 0; s < 4; ++s) {  // K = 16 per instruction: 8 TMEM columns of fp16 pairs, 2 smem chunks
                // (template tag selects the kind::f16 instruction; the fp16 format code is in idesc)
                tc::mma_ts<tc::kFmtBF16>(tb + 64u, at + 8u * s, d0 + dstep * (uint64_t)s, idesc, acc);
                acc = 1;
              }
            }
            tc::mma_commit(&d_full[g]);
          }
        }
        tc::mma_commit(&w_free[b]);  // this issuer's MMAs on weight slot b are done (count 4: all issuers)
      }
    }
  } else {
    // ===== sample warpgroups =====
    const int g = warp >> 2, sl = g, row = tid & 127;
    const uint32_t tb = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)g * 128u;
    for (int ci = 0; ci < my_chunks; ++ci) {
      const int64_t base = ((int64_t)blockIdx.x + (int64_t)ci * gridDim.x) * chunk_samples;
      for (int t = 0; t < T; ++t) {
        const int slot = t * 512 + g * 128 + row;
        const int64_t idx = base + slot;
        float2 v = idx < a.n ? ld_stream_f2(reinterpret_cast<const float2*>(a.in) + idx) : make_float2(0.f, 0.f);
        sx0[slot] = v.x;
        sx1[slot] = v.y;
        sld[slot] = 0.f;
      }
      // Software pipeline over the (layer, tile) steps of this warpgroup: once the accumulator of step k's last
      // Linear sits in registers the whole TMEM slot is free, so step k+1's L1 is written and its MMAs are started
      // BEFORE step k's inner bijector runs — the bijector (~360 instructions per sample) hides that MMA wait.
      // Step k+1 is another tile of the same layer, or tile 0 of the next layer (whose bijector finished T-1 steps
      // ago); with a single tile per chunk (T == 1) the next step needs this one's result and the order is serial.
      auto stage1 = [&](int it, int t) {
        const int l = INV ? a.layer_begin + it : a.layer_end - 1 - it;
        const float* sm = smalls + it * SF;
        const int slot = t * 512 + g * 128 + row;
        const f2_t xm = f2_splat((l & 1) ? sx0[slot] : sx1[slot]);  // visible value
        const f2_t minus1 = f2_splat(-1.f);
        // ---- L1: Linear(2->64)+relu on where(mask, x, 0); split; A operand -> TMEM
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t hi[16], lo[16];
#pragma unroll
          for (int e4 = 0; e4 < 8; ++e4) {
            const ulonglong2 w = *reinterpret_cast<const ulonglong2*>(sm + 32 * c + 4 * e4);
            const ulonglong2 b = *reinterpret_cast<const ulonglong2*>(sm + 64 + 32 * c + 4 * e4);
            relu_split_f16x2(f2_fma(xm, w.x, b.x), minus1, hi[2 * e4], lo[2 * e4]);
            relu_split_f16x2(f2_fma(xm, w.y, b.y), minus1, hi[2 * e4 + 1], lo[2 * e4 + 1]);
          }
          tc::tmem_st_x16(tb + 16u * c, hi);
          tc::tmem_st_x16(tb + 32u + 16u * c, lo);
        }
        tc::tmem_st_wait();
        tc::tc_fence_before();
        tc::mbar_arrive(&a_ready[sl]);
      };
      const bool pipelined = T > 1;
      stage1(0, 0);
      for (int it = 0; it < nl; ++it) {
        const int l = INV ? a.layer_begin + it : a.layer_end - 1 - it;
        const int td = l & 1;  // transformed coordinate (mask = arange(2)%2 on even layers, negated on odd)
        const float* sm = smalls + it * SF;
#pragma unroll 1
        for (int t = 0; t < T; ++t) {
          const int slot = t * 512 + g * 128 + row;
          const int nt = t + 1 == T ? 0 : t + 1, nit = t + 1 == T ? it + 1 : it;  // the next step
          tc::mbar_wait(&d_full[sl], 0);
          tc::tc_fence_after();
          // ---- epilogue 2: unscale, +b2, relu, split, next A operand -> TMEM
          {
            const f2_t inv2 = f2_splat(sm[192 + N3]), minus1 = f2_splat(-1.f);
            uint32_t v[64];
            // the second half of the accumulator is fetched while the first half is processed
            tc::tmem_ld_x32(tb + 64u, v);
            tc::tmem_ld_wait();
            tc::tmem_ld_x32(tb + 96u, v + 32);
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              if (c == 1) tc::tmem_ld_wait();
              uint32_t hi[16], lo[16];
#pragma unroll
              for (int e = 0; e < 16; ++e) {
                const int k = 32 * c + 2 * e;
                const f2_t b2 = *reinterpret_cast<const f2_t*>(sm + 128 + k);
                relu_split_f16x2(f2_fma(f2_pack(__uint_as_float(v[k]), __uint_as_float(v[k + 1])), inv2, b2), minus1,
                                 hi[e], lo[e]);
              }
              tc::tmem_st_x16(tb + 16u * c, hi);
              tc::tmem_st_x16(tb + 32u + 16u * c, lo);
            }
          }
          tc::tmem_st_wait();
          tc::tc_fence_before();
          tc::mbar_arrive(&a_ready[sl]);
          tc::mbar_wait(&d_full[sl], 1);
          tc::tc_fence_after();
          // ---- epilogue 3: inner bijector on the transformed coordinate
          float o[NEED];
          {
            const float inv3 = sm[192 + N3 + 1], inv3l2 = sm[192 + N3 + 2];
            uint32_t v[NEED];
            constexpr int FIRST = NEED > 32 ? 32 : NEED;  // columns past 32 are fetched while the first 32 are unscaled
            tmem_load_cols<FIRST>(tb + 64u, v);
            if (NEED > FIRST) tmem_load_cols<(NEED > FIRST ? NEED - FIRST : 1), false>(tb + 64u + FIRST, v + FIRST);
#pragma unroll
            for (int i = 0; i < NEED; ++i) {
              if (i == FIRST) tc::tmem_ld_wait();
              o[i] = fmaf(__uint_as_float(v[i]), (TR == FLOWCUDA_TRANSFORM_RQS && i < 2 * KT) ? inv3l2 : inv3, sm[192 + i]);
            }
          }
          tc::tc_fence_before();  // the accumulator is drained before this warpgroup's next A operand / MMA
          if (pipelined && nit < nl) stage1(nit, nt);
          const float xt = td ? sx1[slot] : sx0[slot];  // value to transform
          float nv, l1;
          if (TR == FLOWCUDA_TRANSFORM_RQS) {
            rqs_apply_mufu<INV, (KT > 0 ? KT : 1)>([&](int i) { return o[i]; }, a.koff, a.lo, a.hi, xt, nv, l1);
          } else {
            affine_apply<INV>(o[0], o[1], xt, nv, l1);
          }
          if (td) sx1[slot] = nv; else sx0[slot] = nv;
          sld[slot] += l1;
          if (!pipelined && nit < nl) stage1(nit, nt);
        }
      }
      for (int t = 0; t < T; ++t) {
        const int slot = t * 512 + g * 128 + row;
        const int64_t idx = base + slot;
        if (idx < a.n) {
          const float x0 = sx0[slot], x1 = sx1[slot];
          if (a.out_v) st_stream_f2(reinterpret_cast<float2*>(a.out_v) + idx, make_float2(x0, x1));
          if (a.out_lp) {
            float lp = sld[slot];
            if (a.with_base) lp += base_term(a.base, x0) + base_term(a.base, x1);
            a.out_lp[idx] = lp;
          }
        }
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 16) {
    __syncwarp();
    tc::tmem_dealloc(tmem, 512);
  }
}

template <int TR, int KT>
static size_t toy_tc_smem(int nl, int T) {
  using I = ToyTcImg<toy_n3(TR, KT)>;
  return (size_t)2 * I::wbytes + (size_t)((nl * I::small_floats + 3) & ~3) * 4 + (size_t)3 * T * 512 * 4 + 12 * 8 + 16;
}

template <int TR, int KT, bool INV>
static int toy_tc_launch(const flowcuda_coupling_spec* sp, const float* params, ToyTcArgs a, void* ws, size_t wsb,
                         cudaStream_t s) {
  using I = ToyTcImg<toy_n3(TR, KT)>;
  const size_t need = (size_t)sp->num_layers * I::bytes + 256;
  if (!ws || wsb < need) {
    set_error("coupling(3xTF32): workspace %zu < %zu bytes", wsb, need);
    return FLOWCUDA_EWORKSPACE;
  }
  uint8_t* img = (uint8_t*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
  const int of = TR == FLOWCUDA_TRANSFORM_AFFINE_SCALAR ? 2 : 2 * toy_need(TR, KT);
  toy_tc_pack_kernel<TR, KT><<<sp->num_layers, 256, 0, s>>>(params, sp->hidden[0], sp->hidden[1], of, img);
  FC_LAUNCH_CHECK("toy_tc_pack_kernel");
  a.img = img;
  const int nl = a.layer_end - a.layer_begin;
  const int sms = sm_count();
  int T = (int)((a.n + (int64_t)512 * sms - 1) / ((int64_t)512 * sms));
  T = T < 1 ? 1 : (T > kTcMaxT ? kTcMaxT : T);
  while (T > 1 && toy_tc_smem<TR, KT>(nl, T) > 227 * 1024) --T;
  const size_t smem = toy_tc_smem<TR, KT>(nl, T);
  if (smem > 227 * 1024) {
    set_error("coupling(3xTF32): %d layers do not fit in shared memory", nl);
    return FLOWCUDA_EUNSUPPORTED;
  }
  a.T = T;
  auto kern = coupling_toy_tc_kernel<TR, KT, INV>;
  FC_ENSURE_SMEM(kern, smem);
  const int64_t nchunks = (a.n + (int64_t)T * 512 - 1) / ((int64_t)T * 512);
  const int grid = (int)(nchunks < sms ? nchunks : sms);
  kern<<<grid, kTcThreads, smem, s>>>(a);
  FC_LAUNCH_CHECK("coupling_toy_tc_kernel");
  return FLOWCUDA_OK;
}

size_t coupling_toy_tc_workspace_bytes(const flowcuda_coupling_spec* sp) {
  // the largest image (RQS, 12 bins) bounds every variant
  return (size_t)sp->num_layers * ToyTcImg<48>::bytes + 256;
}

int coupling_toy_tc_dispatch(const flowcuda_coupling_spec* sp, const float* params, const float* in, float* out_v,
                             float* out_lp, int64_t n, int layer_begin, int layer_end, int with_base, bool inverse,
                             void* ws, size_t wsb, cudaStream_t s) {
  if (sp->event_dim != 2 || sp->num_hidden != 2 || sp->hidden[0] > 64 || sp->hidden[1] > 64) return FLOWCUDA_EUNSUPPORTED;
  if (layer_end <= layer_begin) return FLOWCUDA_EUNSUPPORTED;
  ToyTcArgs a;
  a.img = nullptr; a.in = in; a.out_v = out_v; a.out_lp = out_lp; a.n = n;
  a.layer_begin = layer_begin; a.layer_end = layer_end; a.with_base = with_base; a.base = sp->base;
  a.lo = sp->range_min; a.hi = sp->range_max; a.T = 1;
  for (int i = 0; i < 13; ++i) a.koff[i] = sp->range_min + (float)i * kRqsMinBin;
  switch (sp->transform) {
    case FLOWCUDA_TRANSFORM_AFFINE_SCALAR:
      return inverse ? toy_tc_launch<FLOWCUDA_TRANSFORM_AFFINE_SCALAR, 0, true>(sp, params, a, ws, wsb, s)
                     : toy_tc_launch<FLOWCUDA_TRANSFORM_AFFINE_SCALAR, 0, false>(sp, params, a, ws, wsb, s);
    case FLOWCUDA_TRANSFORM_AFFINE:
      return inverse ? toy_tc_launch<FLOWCUDA_TRANSFORM_AFFINE, 0, true>(sp, params, a, ws, wsb, s)
                     : toy_tc_launch<FLOWCUDA_TRANSFORM_AFFINE, 0, false>(sp, params, a, ws, wsb, s);
    case FLOWCUDA_TRANSFORM_RQS:
      if (sp->num_bins == 12)
        return inverse ? toy_tc_launch<FLOWCUDA_TRANSFORM_RQS, 12, true>(sp, params, a, ws, wsb, s)
                       : toy_tc_launch<FLOWCUDA_TRANSFORM_RQS, 12, false>(sp, params, a, ws, wsb, s);
      if (sp->num_bins == 8)
        return inverse ? toy_tc_launch<FLOWCUDA_TRANSFORM_RQS, 8, true>(sp, params, a, ws, wsb, s)
                       : toy_tc_launch<FLOWCUDA_TRANSFORM_RQS, 8, false>(sp, params, a, ws, wsb, s);
      return FLOWCUDA_EUNSUPPORTED;
  }
  return FLOWCUDA_EUNSUPPORTED;
}

}  // namespace flowcuda
