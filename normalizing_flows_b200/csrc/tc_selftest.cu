// tc_selftest.cu — diagnostic entry point that exercises every tensor-core primitive of tc.cuh
// on one 128-row tile: tile-image staging by st.shared (A) and by pack kernel + TMA bulk copy
// (B), tcgen05.mma kind::tf32 / kind::f16 in 1-pass and 3-pass form, tcgen05.commit ->
// mbarrier, tcgen05.ld at aligned and unaligned column offsets.  tests/test_gpu_tc.py compares
// the result with a float64 matmul.
#include "tc.cuh"
#include "../../include/flowcuda_selftest.h"

namespace flowcuda {

// W[K,N] row-major fp32 (y = x @ W) -> B-operand tile images (rows = n, K-major), hi and lo.
// BF16 = false: tf32 words (4 per chunk); true: bf16 (8 per chunk).
template <bool BF16>
__global__ void __launch_bounds__(256) pack_b_image_kernel(const float* __restrict__ W, int K, int N, int ldw,
                                                           uint8_t* __restrict__ hi, uint8_t* __restrict__ lo) {
  constexpr int CH = BF16 ? 8 : 4;
  const int chunks = K / CH;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N * chunks; i += gridDim.x * blockDim.x) {
    const int n = i % N, kc = i / N;
    const uint32_t off = tc::img_off(N, n, kc);
    float v[CH];
#pragma unroll
    for (int e = 0; e < CH; ++e) v[e] = W[(size_t)(kc * CH + e) * ldw + n];
    if (BF16) {
      uint32_t h[4], l[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) tc::split_bf16x2(v[2 * e], v[2 * e + 1], h[e], l[e]);
      *reinterpret_cast<uint4*>(hi + off) = make_uint4(h[0], h[1], h[2], h[3]);
      *reinterpret_cast<uint4*>(lo + off) = make_uint4(l[0], l[1], l[2], l[3]);
    } else {
      float h[4], l[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) tc::split_tf32(v[e], h[e], l[e]);
      *reinterpret_cast<float4*>(hi + off) = make_float4(h[0], h[1], h[2], h[3]);
      *reinterpret_cast<float4*>(lo + off) = make_float4(l[0], l[1], l[2], l[3]);
    }
  }
}

template <bool BF16>
__global__ void __launch_bounds__(128) tc_selftest_kernel(const float* __restrict__ A, const uint8_t* __restrict__ Bhi,
                                                          const uint8_t* __restrict__ Blo, float* __restrict__ D,
                                                          float* __restrict__ D2, int N, int K, int three_pass,
                                                          int a_in_tmem, uint32_t tmem_cols) {
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr int CH = BF16 ? 8 : 4;
  constexpr int ES = BF16 ? 2 : 4;
  constexpr uint32_t FMT = BF16 ? tc::kFmtBF16 : tc::kFmtTF32;
  const uint32_t a_bytes = 128u * K * ES, b_bytes = (uint32_t)N * K * ES;
  uint8_t* a_hi = smem;
  uint8_t* a_lo = a_hi + a_bytes;
  uint8_t* b_hi = a_lo + a_bytes;
  uint8_t* b_lo = b_hi + b_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(b_lo + b_bytes);  // [0] B landed, [1] D ready
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
  const int tid = threadIdx.x, warp = tid >> 5;

  if (warp == 0) tc::tmem_alloc(tmem_slot, tmem_cols);
  if (tid == 0) {
    tc::mbar_init(&bars[0], 1);
    tc::mbar_init(&bars[1], 1);
    tc::fence_mbar_init();
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (tid == 0) {
    tc::mbar_arrive_expect_tx(&bars[0], 2 * b_bytes);
    tc::tma_bulk_g2s(b_hi, Bhi, b_bytes, &bars[0]);
    tc::tma_bulk_g2s(b_lo, Blo, b_bytes, &bars[0]);
  }
  // A: thread = row.  a_in_tmem: operand A goes to TMEM columns [256, 256+Kc) (hi) and
  // [256+Kc, 256+2Kc) (lo), lane = row, one 32-bit column per tf32 element / bf16 pair.
  const uint32_t Kc = (uint32_t)(K * ES / 4);
  const uint32_t my_lanes = tmem + ((uint32_t)(warp * 32) << 16);
  for (int kc = 0; kc < K / CH; ++kc) {
    const float* src = A + (size_t)tid * K + kc * CH;
    const uint32_t off = tc::img_off(128, tid, kc);
    if (BF16) {
      uint32_t h[4], l[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) tc::split_bf16x2(src[2 * e], src[2 * e + 1], h[e], l[e]);
      if (a_in_tmem) {
        tc::tmem_st_x4(my_lanes + 256u + 4u * kc, h);
        tc::tmem_st_x4(my_lanes + 256u + Kc + 4u * kc, l);
      } else {
        *reinterpret_cast<uint4*>(a_hi + off) = make_uint4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<uint4*>(a_lo + off) = make_uint4(l[0], l[1], l[2], l[3]);
      }
    } else {
      float h[4], l[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) tc::split_tf32(src[e], h[e], l[e]);
      if (a_in_tmem) {
        tc::tmem_st_x4(my_lanes + 256u + 4u * kc, reinterpret_cast<uint32_t*>(h));
        tc::tmem_st_x4(my_lanes + 256u + Kc + 4u * kc, reinterpret_cast<uint32_t*>(l));
      } else {
        *reinterpret_cast<float4*>(a_hi + off) = make_float4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<float4*>(a_lo + off) = make_float4(l[0], l[1], l[2], l[3]);
      }
    }
  }
  if (a_in_tmem) tc::tmem_st_wait();
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();

  if (tid == 0) {
    tc::mbar_wait(&bars[0], 0);
    tc::tc_fence_after();
    const uint32_t idesc = tc::make_idesc(FMT, 128, N);
    const int ksteps = K / (2 * CH);
    if (a_in_tmem) {
      const uint32_t lbo_b = (uint32_t)(N >> 3) * 128u;
      uint32_t acc = 0;
      for (int pass = 0; pass < 3; ++pass) {
        const uint32_t at = tmem + 256u + (pass == 2 ? Kc : 0u);
        const uint32_t b = tc::smem_u32(pass == 1 ? b_lo : b_hi);
        for (int s = 0; s < ksteps; ++s) {
          tc::mma_ts<FMT>(tmem, at + 8u * s, tc::make_smem_desc(b + s * 2u * lbo_b, lbo_b, 128u), idesc, acc);
          acc = 1;
        }
      }
    } else if (three_pass) {
      tc::mma_3pass<FMT>(tmem, tc::smem_u32(a_hi), tc::smem_u32(a_lo), 128, tc::smem_u32(b_hi), tc::smem_u32(b_lo), N,
                         ksteps, idesc, false);
    } else {
      const uint32_t lbo_a = 16u * 128u, lbo_b = (uint32_t)(N >> 3) * 128u;
      for (int s = 0; s < ksteps; ++s)
        tc::mma_ss<FMT>(tmem, tc::make_smem_desc(tc::smem_u32(a_hi) + s * 2u * lbo_a, lbo_a, 128u),
                        tc::make_smem_desc(tc::smem_u32(b_hi) + s * 2u * lbo_b, lbo_b, 128u), idesc, s > 0);
    }
    tc::mma_commit(&bars[1]);
  }
  tc::mbar_wait(&bars[1], 0);
  tc::tc_fence_after();
  const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
  for (int c = 0; c < N; c += 8) {
    uint32_t v[8];
    tc::tmem_ld_x8(lane_base + (uint32_t)c, v);
    tc::tmem_ld_wait();
#pragma unroll
    for (int e = 0; e < 8; ++e) D[(size_t)tid * N + c + e] = __uint_as_float(v[e]);
  }
  if (D2) {  // 25-column groups at unaligned offsets: x16 @25d, x8 @25d+16, x1 @25d+24
    for (int d = 0; d < N / 25; ++d) {
      uint32_t v[25];
      tc::tmem_ld_x16(lane_base + (uint32_t)(25 * d), v);
      tc::tmem_ld_x8(lane_base + (uint32_t)(25 * d + 16), v + 16);
      tc::tmem_ld_x1(lane_base + (uint32_t)(25 * d + 24), v + 24);
      tc::tmem_ld_wait();
#pragma unroll
      for (int e = 0; e < 25; ++e) D2[(size_t)tid * N + 25 * d + e] = __uint_as_float(v[e]);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, tmem_cols);
}

}  // namespace flowcuda

using namespace flowcuda;

extern "C" {

size_t flowcuda_selftest_umma_workspace_bytes(int n, int k) { return (size_t)2 * n * k * 4 + 512; }

int flowcuda_selftest_umma(const float* a, const float* w, float* d, float* d_unaligned, int n, int k, int mode,
                           void* workspace, size_t workspace_bytes, void* stream) {
  FC_REQUIRE(a && w && d && workspace, "selftest_umma: NULL pointer");
  FC_REQUIRE(mode >= 0 && mode <= 5, "selftest_umma: mode must be 0 (tf32x3), 1 (bf16x3), 2 (tf32), 3 (bf16), 4/5 (tf32x3/bf16x3, A in TMEM)");
  const bool bf16 = (mode & 1) != 0;
  FC_REQUIRE(n >= 16 && n <= 256 && n % 16 == 0, "selftest_umma: n=%d must be a multiple of 16 in [16,256]", n);
  FC_REQUIRE(k >= 16 && k % 16 == 0 && k <= 128, "selftest_umma: k=%d must be a multiple of 16 in [16,128]", k);
  FC_REQUIRE(workspace_bytes >= flowcuda_selftest_umma_workspace_bytes(n, k), "selftest_umma: workspace too small");
  cudaStream_t s = (cudaStream_t)stream;
  const int es = bf16 ? 2 : 4;
  uint8_t* hi = (uint8_t*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  uint8_t* lo = hi + (size_t)n * k * es;
  uint32_t cols = 32;
  while (cols < (uint32_t)n) cols <<= 1;
  const int a_tmem = mode >= 4;
  if (a_tmem) cols = 512;
  size_t smem = (size_t)2 * 128 * k * es + (size_t)2 * n * k * es + 64;
  if (bf16) {
    pack_b_image_kernel<true><<<8, 256, 0, s>>>(w, k, n, n, hi, lo);
    FC_LAUNCH_CHECK("pack_b_image_kernel");
    FC_ENSURE_SMEM(tc_selftest_kernel<true>, smem);
    tc_selftest_kernel<true><<<1, 128, smem, s>>>(a, hi, lo, d, d_unaligned, n, k, mode < 2 || a_tmem, a_tmem, cols);
  } else {
    pack_b_image_kernel<false><<<8, 256, 0, s>>>(w, k, n, n, hi, lo);
    FC_LAUNCH_CHECK("pack_b_image_kernel");
    FC_ENSURE_SMEM(tc_selftest_kernel<false>, smem);
    tc_selftest_kernel<false><<<1, 128, smem, s>>>(a, hi, lo, d, d_unaligned, n, k, mode < 2 || a_tmem, a_tmem, cols);
  }
  FC_LAUNCH_CHECK("tc_selftest_kernel");
  return FLOWCUDA_OK;
}

}  // extern "C"

// ---- MMA timing microbenchmark: chains of small tcgen05.mma kind::tf32 (M128, K8) ------------------
namespace flowcuda {
// out[4*i + {0,1,2,3}] = {N, nacc*1000 + n_mma (+100000 if A from smem), issue cycles, total cycles to commit arrival}
__global__ void __launch_bounds__(128) tc_mma_timing_kernel(volatile long long* out) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 160 * 1024);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 160 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (warp == 0) tc::tmem_alloc(slot, 512);
  if (tid == 0) {
    tc::mbar_init(bar, 1);
    tc::mbar_init(bar + 2, 1);
    tc::fence_mbar_init();
  }
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *slot;
  if (tid == 0) {
    const int Ns[4] = {48, 64, 128, 256};
    int rec = 0;
    uint32_t phase = 0;
    for (int a_smem = 0; a_smem < 2; ++a_smem)
      for (int ni = 0; ni < 4; ++ni)
        for (int nacc = 1; nacc <= 3; ++nacc)
          for (int nm = 24; nm <= 96; nm += 72) {
            const int N = Ns[ni];
            if (nacc * N > 384) continue;
            const uint32_t idesc = tc::make_idesc(tc::kFmtTF32, 128, N);
            const uint32_t lbo_b = (uint32_t)(N >> 3) * 128u, lbo_a = 16u * 128u;
            const uint32_t bbase = tc::smem_u32(smem), abase = tc::smem_u32(smem + 96 * 1024);
            long long t0 = clock64();
            for (int m = 0; m < nm; ++m) {
              const uint32_t d = tmem + 128u + (uint32_t)((m % nacc) * N);
              const uint32_t s = (uint32_t)(m & 7);
              if (a_smem)
                tc::mma_ss<tc::kFmtTF32>(d, tc::make_smem_desc(abase + s * 2u * lbo_a, lbo_a, 128u),
                                         tc::make_smem_desc(bbase + s * 2u * lbo_b, lbo_b, 128u), idesc, 1u);
              else
                tc::mma_ts<tc::kFmtTF32>(d, tmem + 8u * s, tc::make_smem_desc(bbase + s * 2u * lbo_b, lbo_b, 128u), idesc, 1u);
            }
            long long t1 = clock64();
            tc::mma_commit(bar);
            tc::mbar_wait(bar, phase);
            phase ^= 1;
            long long t2 = clock64();
            out[4 * rec + 0] = N;
            out[4 * rec + 1] = nacc * 1000 + nm + (a_smem ? 100000 : 0);
            out[4 * rec + 2] = t1 - t0;
            out[4 * rec + 3] = t2 - t0;
            ++rec;
          }
    // tight variant: descriptors prebuilt, 8-way unrolled, A from TMEM, one accumulator
    for (int ni = 0; ni < 5; ++ni) {
      const int Nt[5] = {16, 48, 64, 128, 256};
      const int N = Nt[ni];
      const uint32_t idesc = tc::make_idesc(tc::kFmtTF32, 128, N);
      const uint32_t lbo_b = (uint32_t)(N >> 3) * 128u;
      uint64_t bd[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) bd[k] = tc::make_smem_desc(tc::smem_u32(smem) + k * 2u * lbo_b, lbo_b, 128u);
      for (int nm = 24; nm <= 96; nm += 72) {
        long long t0 = clock64();
        for (int m = 0; m < nm; m += 8) {
#pragma unroll
          for (int k = 0; k < 8; ++k) tc::mma_ts<tc::kFmtTF32>(tmem + 128u, tmem + 8u * k, bd[k], idesc, 1u);
        }
        long long t1 = clock64();
        tc::mma_commit(bar);
        tc::mbar_wait(bar, phase);
        phase ^= 1;
        long long t2 = clock64();
        out[4 * rec + 0] = N;
        out[4 * rec + 1] = 9000 + nm;
        out[4 * rec + 2] = t1 - t0;
        out[4 * rec + 3] = t2 - t0;
        ++rec;
      }
    }
    // bf16 kind::f16 SS MMAs (K = 16) exactly as the D=784 GEMMs issue them: one descriptor pair per 32-K stage, six MMAs
    // per stage (3 passes x 2 K steps) over three stage buffers, one accumulator; code 7000 + n_mma, N = 128 / 256
    for (int ni = 0; ni < 2; ++ni) {
      const int N = ni == 0 ? 128 : 256;
      const uint32_t idesc = tc::make_idesc(tc::kFmtBF16, 128, N);
      const uint32_t lbo_a = 16u * 128u, lbo_b = (uint32_t)(N >> 3) * 128u;
      const uint32_t half_b = (uint32_t)N * 32u * 2u;  // bytes of one of hi/lo of a B stage
      const uint32_t stage = 2u * 8192u + 2u * half_b;
      for (int var = 0; var < 3; ++var)  // 0: MMAs only; 1: + tcgen05 fence and a commit per stage (to a second barrier); 2: + a satisfied try_wait per stage
      for (int nm = 96; nm <= 192; nm += 96) {
        uint32_t acc = 0;
        long long t0 = clock64();
        for (int m = 0; m < nm; m += 6) {
          if (var >= 2) tc::mbar_wait(bar, phase ^ 1u);  // the previous phase: passes at once, costs the probe
          if (var >= 1) tc::tc_fence_after();
          const uint32_t sa = tc::smem_u32(smem) + (uint32_t)((m / 6) % 3) * stage, sb = sa + 2u * 8192u;
          const uint64_t da = tc::make_smem_desc(sa, lbo_a, 128u), db = tc::make_smem_desc(sb, lbo_b, 128u);
#pragma unroll
          for (int pass = 0; pass < 3; ++pass) {
            const uint32_t ao = pass == 2 ? 8192u : 0u, bo = pass == 1 ? half_b : 0u;
#pragma unroll
            for (int k = 0; k < 2; ++k) {
              tc::mma_ss<tc::kFmtBF16>(tmem + 128u, da + ((ao + k * 2u * lbo_a) >> 4), db + ((bo + k * 2u * lbo_b) >> 4), idesc, acc);
              acc = 1;
            }
          }
          if (var >= 1) tc::mma_commit(bar + 2);  // a barrier nobody waits on (count 1: phases just roll over)
        }
        long long t1 = clock64();
        tc::mma_commit(bar);
        tc::mbar_wait(bar, phase);
        phase ^= 1;
        long long t2 = clock64();
        out[4 * rec + 0] = N;
        out[4 * rec + 1] = 7000 - 1000 * var + nm;
        out[4 * rec + 2] = t1 - t0;
        out[4 * rec + 3] = t2 - t0;
        ++rec;
      }
    }
    out[4 * rec] = -1;
  }
  __syncthreads();
  if (warp == 0) {
    // warp-convergent variant (code 8000 + n): every lane runs the loop, one elected lane issues
    int rec = 0;
    while (out[4 * rec] != -1) ++rec;
    __syncwarp();
    uint32_t phase = 0;
    {  // bar was used `rec` times by tid 0 above
      phase = (uint32_t)(rec & 1);
    }
    for (int ni = 0; ni < 5; ++ni) {
      const int Nt[5] = {16, 48, 64, 128, 256};
      const int N = Nt[ni];
      const uint32_t idesc = tc::make_idesc(tc::kFmtTF32, 128, N);
      const uint32_t lbo_b = (uint32_t)(N >> 3) * 128u;
      const uint32_t bbase = tc::smem_u32(smem);
      for (int nm = 24; nm <= 96; nm += 72) {
        long long t0 = clock64();
        for (int m = 0; m < nm; m += 8) {
#pragma unroll
          for (int k = 0; k < 8; ++k)
            tc::mma_ts_warp<tc::kFmtTF32>(tmem + 128u, tmem + 8u * k, tc::make_smem_desc(bbase + k * 2u * lbo_b, lbo_b, 128u),
                                          idesc, 1u);
        }
        long long t1 = clock64();
        tc::mma_commit_warp(bar);
        tc::mbar_wait(bar, phase);
        phase ^= 1;
        long long t2 = clock64();
        if ((tid & 31) == 0) {
          out[4 * rec + 0] = N;
          out[4 * rec + 1] = 8000 + nm;
          out[4 * rec + 2] = t1 - t0;
          out[4 * rec + 3] = t2 - t0;
        }
        ++rec;
      }
    }
    if ((tid & 31) == 0) out[4 * rec] = -1;
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}
}  // namespace flowcuda

extern "C" int flowcuda_selftest_mma_timing(long long* out_device, int capacity, void* stream) {
  FC_REQUIRE(out_device && capacity >= 4 * 64, "selftest_mma_timing: need room for 64 records");
  const size_t smem = 160 * 1024 + 64;
  FC_ENSURE_SMEM(flowcuda::tc_mma_timing_kernel, smem);
  flowcuda::tc_mma_timing_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(out_device);
  FC_LAUNCH_CHECK("tc_mma_timing_kernel");
  return FLOWCUDA_OK;
}

// this test library links its own copy of core.cu's error plumbing
extern "C" const char* flowcuda_selftest_last_error(void) { return flowcuda_last_error(); }
