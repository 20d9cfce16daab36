// dense_tc_chain.cuh — the hidden Linears of one conditioner as ONE kernel (coupling flows, 512-wide hidden layers).
//
// A conditioner MLP 784 -> 512 -> 512 -> 512 -> 19600 (src/maf_mnist.py:34-50) spends 4 % of its FLOPs in the three
// hidden Linears, but as three launches of dense_tc_gemm_kernel they took 39 of a flow layer's 112 us at batch 4096: each
// is a latency-bound chain (prologue, first operand stage, 16 K stages, epilogue, drain) of one 128 x 128 tile per CTA.
// Here a CLUSTER OF FOUR CTAs owns one 128-row m tile for the whole chain: CTA p computes columns [128 p, 128 p + 128)
// of every hidden Linear (tcgen05 128 x 128 x 16 MMAs, bf16x3 as everywhere), writes its quarter of the next Linear's
// activation image (the same [m tile][k stage][hi|lo] image the per-Linear kernels use, so the transform GEMM that
// follows is unchanged), and the four meet at a cluster barrier before the next Linear's operand stream starts.
// One launch, one prologue, no drain between the Linears; barriers, TMEM and the TMA ring are set up once.
//
// What the per-Linear kernels were really bound by is L2 -> SM bandwidth: 128 CTAs x 16 stages x 32 KB = 64 MB per Linear
// at the ~5.6 TB/s the L2 delivers = 11.7 us, against 3 us of MMAs.  The four CTAs of a cluster need the SAME activation
// stage: each fetches a quarter of it (4 KB) and MULTICASTS it to all four (cp.async.bulk ... .multicast::cluster), so
// the activation traffic drops from 32 to 8 MB per Linear.  A stage slot may be refilled only when all four issuers have
// consumed it: the MMA completion is multicast to the `empty` barrier of every CTA (count 4).
//
// Cross-proxy ordering: the image quarters are written with generic-proxy stores and read back by cp.async.bulk (async
// proxy) of ALL four CTAs — every writer issues fence.proxy.async.global (the unqualified form adds a CTA-wide membar: 8 % of the kernel) before it arrives at the cluster barrier
// (arrive.release / wait.acquire), and the producer thread issues it again after the wait.
#pragma once
#include "dense_tc2.cuh"

namespace flowcuda {

namespace tc {
// bulk copy global -> the same shared-memory offset of every CTA in `mask` (complete_tx on the same-offset barrier of each)
__device__ __forceinline__ void tma_bulk_g2s_mc(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar, uint16_t mask) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "h"(mask)
               : "memory");
}
// MMA completion -> one arrival on the same-offset barrier of every CTA in `mask`
__device__ __forceinline__ void mma_commit_mc(uint64_t* bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
               "h"(mask)
               : "memory");
}
__device__ __forceinline__ void mma_commit_mc_warp(uint64_t* bar, uint16_t mask) {  // whole warp executes, the elected lane commits
  asm volatile(
      "{\n\t.reg .pred q;\n\telect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n\t}" ::"r"(smem_u32(bar)),
      "h"(mask)
      : "memory");
}
}  // namespace tc

constexpr int kDcMaxLinears = FLOWCUDA_MAX_HIDDEN_LAYERS;
constexpr int kDcCluster = 4;                       // CTAs per m tile = 512 / 128 columns
constexpr int kDcStages = 6;
constexpr int kDcStageBytes = 2 * kDtABytes + kDtBBytes;  // A hi/lo (16 KB) + 128 rows of B hi/lo (16 KB)
constexpr int kDcEpiWarps = 8;
constexpr int kDcThreads = 32 * (kDcEpiWarps + 2);
constexpr size_t kDcSmem = (size_t)kDcStages * kDcStageBytes + kDtBN * 4 + (2 * kDcStages + 2) * 8 + 16;

struct DtChain {
  int nlin;                          // hidden Linears in the chain (every one 512 wide)
  int64_t N;
  const uint8_t* a_img0;             // input image of the first Linear, KS[0] stages
  int KS[kDcMaxLinears];             // K stages of Linear i's operands
  const uint8_t* b_img[kDcMaxLinears];   // weight image of Linear i in 128-ROW tiles: [4 n tiles][KS[i]][hi|lo][8 KB]
  const float* bias[kDcMaxLinears];
  uint8_t* out_img[kDcMaxLinears];   // activation image Linear i writes (out_KS[i] stages): the next Linear's / the transform GEMM's A operand
  int out_KS[kDcMaxLinears];
  float* store_h[kDcMaxLinears];     // training: relu(h_i) as fp32 [N, 512] (NULL: not stored)
};

static __global__ void __launch_bounds__(kDcThreads, 1) dense_tc_chain_kernel(DtChain a) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* stages = smem;
  float* sbias = reinterpret_cast<float*>(smem + kDcStages * kDcStageBytes);  // [256]: indexed by tile column
  uint64_t* full = reinterpret_cast<uint64_t*>(sbias + kDtBN);
  uint64_t* empty = full + kDcStages;
  uint64_t* acc_full = empty + kDcStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 2);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  constexpr int kLoadWarp = kDcEpiWarps, kMmaWarp = kDcEpiWarps + 1;
  const int part = (int)(blockIdx.x % kDcCluster), mt = (int)(blockIdx.x / kDcCluster);
  const int nt = part >> 1, half = part & 1, cshift = half * 128;

  if (warp == kMmaWarp) tc::tmem_alloc(tmem_slot, 128);
  if (tid == kLoadWarp * 32) {
    for (int s = 0; s < kDcStages; ++s) {
      tc::mbar_init(&full[s], 1);
      tc::mbar_init(&empty[s], kDcCluster);  // one arrival from the issuer of every CTA of the cluster
    }
    tc::mbar_init(&acc_full[0], 1);
    tc::fence_mbar_init();
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  tc::cluster_sync_all();  // every CTA's barriers exist before a peer's multicast or commit can reach them
  pdl_wait();
  pdl_launch_dependents();

  uint32_t it = 0;  // ring position, carried across the Linears (producer and issuer count alike)
  for (int l = 0; l < a.nlin; ++l) {
    const int KS = a.KS[l];
    if (warp == kLoadWarp) {
      if (lane == 0) {  // ===== TMA producer
        if (l > 0) asm volatile("fence.proxy.async.global;" ::: "memory");
        const uint8_t* ag = (l == 0 ? a.a_img0 : a.out_img[l - 1]) + (int64_t)mt * KS * 2 * kDtABytes;
        const uint8_t* bg = a.b_img[l] + (int64_t)part * KS * kDtBBytes;  // this CTA's 128-row tile: 16 KB (hi|lo) per stage
        uint32_t p = it;
        for (int ks = 0; ks < KS; ++ks, ++p) {
          const uint32_t st = p % kDcStages, ph = (p / kDcStages) & 1u;
          tc::mbar_wait(&empty[st], ph ^ 1u);
          uint8_t* dst = stages + st * kDcStageBytes;
          tc::mbar_arrive_expect_tx(&full[st], kDcStageBytes);
          // this CTA's quarter of the activation stage, to all four CTAs (theirs arrive the same way)
          tc::tma_bulk_g2s_mc(dst + part * (2 * kDtABytes / kDcCluster), ag + (int64_t)ks * 2 * kDtABytes + part * (2 * kDtABytes / kDcCluster),
                              2 * kDtABytes / kDcCluster, &full[st], (uint16_t)((1u << kDcCluster) - 1u));
          // ONE copy for the weight stage: with 256-row images a 128-row half is eight 2 KB runs, and nine bulk copies
          // per stage from one thread (~150 cycles each) were what a hidden Linear's 12 us really consisted of
          tc::tma_bulk_g2s(dst + 2 * kDtABytes, bg + (int64_t)ks * kDtBBytes, kDtBBytes, &full[st]);
        }
      }
      __syncwarp();
    } else if (warp == kMmaWarp) {
      {  // ===== MMA issuer: 3 passes x 2 K steps per stage, N = 128; whole warp converged, the elected lane issues (dense_tc.cuh)
        tc::tc_fence_after();
        constexpr uint32_t idesc = tc::make_idesc(tc::kFmtBF16, 128, 128);
        constexpr uint32_t lbo = 16u * 128u;
        uint32_t p = it, acc = 0;
        for (int ks = 0; ks < KS; ++ks, ++p) {
          const uint32_t st = p % kDcStages, ph = (p / kDcStages) & 1u;
          tc::mbar_wait(&full[st], ph);
          tc::tc_fence_after();
          const uint32_t sa = tc::smem_u32(stages + st * kDcStageBytes), sb = sa + 2 * kDtABytes;
          const uint64_t da = tc::make_smem_desc(sa, lbo, 128u), db = tc::make_smem_desc(sb, lbo, 128u);  // per stage; see dense_tc.cuh
#pragma unroll
          for (int pass = 0; pass < 3; ++pass) {
            const uint32_t ao = pass == 2 ? (uint32_t)kDtABytes : 0u, bo = pass == 1 ? (uint32_t)kDtABytes : 0u;
#pragma unroll
            for (int s = 0; s < kDtBK / 16; ++s) {
              tc::mma_ss_warp<tc::kFmtBF16>(tmem, da + ((ao + s * 2u * lbo) >> 4), db + ((bo + s * 2u * lbo) >> 4), idesc, acc);
              acc = 1;
            }
          }
          tc::mma_commit_mc_warp(&empty[st], (uint16_t)((1u << kDcCluster) - 1u));
        }
        tc::mma_commit_warp(&acc_full[0]);
      }
      __syncwarp();
    } else {
      // ===== epilogue: warp w reads TMEM lane quarter w % 4; the two warps of a quarter take 64 columns each
      for (int n = tid; n < 128; n += kDcEpiWarps * 32) sbias[cshift + n] = a.bias[l][nt * kDtBN + cshift + n];
      asm volatile("bar.sync 1, %0;" ::"n"(kDcEpiWarps * 32) : "memory");
      DtGemm g;
      g.out_img = a.out_img[l];
      g.out_KS = a.out_KS[l];
      g.N = a.N;
      g.n_out = 512;
      g.store_h = a.store_h[l];
      g.store_ld = 512;
      tc::mbar_wait(&acc_full[0], (uint32_t)l & 1u);
      tc::tc_fence_after();
      const uint32_t tb = tmem + ((uint32_t)((warp & 3) * 32) << 16);
      const int c_lo = cshift + (warp >> 2) * 64;
      dt_epilogue<kEpiHidden, false>(g, mt, nt, tb, sbias, nullptr, warp, lane, c_lo, c_lo + 64, cshift);
      tc::tc_fence_before();
      asm volatile("fence.proxy.async.global;" ::: "memory");  // the image quarter (global memory) is read back through the async proxy
    }
    it += (uint32_t)KS;
    if (l + 1 < a.nlin) tc::cluster_sync_all();  // all four quarters of the next operand image are in place
  }
  tc::tc_fence_before();
  tc::cluster_sync_all();  // no CTA leaves while a peer's multicast copy or commit can still land in it
  if (warp == kMmaWarp) {
    __syncwarp();
    tc::tmem_dealloc(tmem, 128);
  }
}

// When the chain kernel runs (FLOWCUDA_HIDDEN_CHAIN, read on every call so that tests can switch in-process:
// unset = by batch size, "1" = whenever the shapes allow, "0" = never).  Measured on B200 (profiles/r02/README.md):
// at the maf_mnist bench shape (batch 4096: 32 clusters, one wave) the step goes 0.998 -> 0.953 ms; at batch 16384 /
// 32768 the grid is several waves of non-persistent clusters and the per-Linear kernels, throughput-bound there, win
// (deep16_nvp 1.54 -> 1.84 ms, maf_mnist_b32768 6.93 -> 7.38 ms).  Default: one wave only.
// History worth keeping: with the MMAs issued from single-thread code (~95 cycles per narrow MMA, issue-bound) this
// kernel was no faster than the three launches it replaces; the multicast and the single-copy weight stage only paid
// once the issue loop ran warp-converged (tc::mma_ss_warp).
static int dt_chain_mode() {
  const char* e = getenv("FLOWCUDA_HIDDEN_CHAIN");
  if (e && e[0] == '1') return 1;
  if (e && e[0] == '0') return 0;
  return -1;
}
static bool dt_chain_possible() { return dt_chain_mode() != 0; }  // are the 128-row weight images packed at all
static bool dt_chain_wanted(int64_t MT) {
  const int m = dt_chain_mode();
  return m == 1 || (m < 0 && MT * kDcCluster <= sm_count());
}

static int dt_launch_chain(const DtChain& c, int64_t MT, cudaStream_t s) {
  FC_ENSURE_SMEM(dense_tc_chain_kernel, kDcSmem);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(MT * kDcCluster));
  cfg.blockDim = dim3(kDcThreads);
  cfg.dynamicSmemBytes = kDcSmem;
  cfg.stream = s;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = kDcCluster;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  FC_CUDA(cudaLaunchKernelEx(&cfg, dense_tc_chain_kernel, c));
  return FLOWCUDA_OK;
}

}  // namespace flowcuda
