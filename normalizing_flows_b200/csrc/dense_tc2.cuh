// dense_tc2.cuh — the bf16x3 GEMM of the wide-conditioner flows as CTA-PAIR tcgen05 MMAs (cta_group::2) fed by
// tensor-map TMA (cp.async.bulk.tensor -> UTMALDG) with multicast across the pairs of a 4-CTA cluster.
//
// Why: the 1-CTA kernel (dense_tc.cuh) pulls 48 KB per 128x256x32 stage per SM out of L2 — 62 B per SM-cycle at
// the full tensor rate, ~9.5 KB/cycle chip-wide, which is more than L2 delivers (measured ~8.6 TB/s), so its
// tensor pipe sat at 38-56 %.  Here
//   * a CTA pair shares one 256(M) x 256(N) accumulator tile: each CTA stages its own 128 rows of A and only HALF
//     of B (128 of the 256 weight rows); the pair's tensor cores read the other half from the peer's shared memory
//     (tcgen05.mma.cta_group::2, M = 256) — 32 KB per stage per SM instead of 48 KB;
//   * two pairs of a 4-CTA cluster work on the same 256 sample rows and adjacent n tiles, so the A tile of CTA
//     (pair p, rank r) is identical in both pairs: each CTA fetches half of it (the hi or the lo image) and
//     multicasts it to the CTA of the other pair (.multicast::cluster) — 24 KB per stage per SM.
// Operands stay in the library's pre-tiled "tile image" form (tc.cuh: 8 rows x 16 B core matrices, each 128
// contiguous bytes, K-major, SWIZZLE_NONE): the images are WRITTEN tiled by the producing kernels, so the
// tensor-core reads are conflict-free by construction and no 128-byte swizzle is needed (swizzling exists to make
// row-major TMA boxes conflict-free).  The tensor maps describe the images as arrays of 2 KB blocks, so one
// UTMALDG moves a whole 8-16 KB operand stage; with .cta_group::2 every copy of a pair completes on the
// LEADER's mbarrier, which is what lets one thread issue the pair's MMAs.
//
// Roles per CTA (one CTA per SM, persistent over the tile list): warp 0 = TMA producer, warp 1 = TMEM allocator
// and (leader CTA of a pair) MMA issuer, then two epilogue groups alternating over the two 256-column TMEM
// accumulators — the epilogues are dense_tc.cuh's (hidden / RQS / affine), unchanged: thread = sample row.
#pragma once
#include <stdlib.h>

#include <cuda.h>  // CUtensorMap (types only; the encoder is fetched through the runtime, no -lcuda)

#include "dense_tc.cuh"

namespace flowcuda {

namespace tc {

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_id_x() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t nclusters_x() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%nclusterid.x;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local` (a shared::cta address of this CTA) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
// arrive on an mbarrier of another CTA of the cluster.  Default semantics (release at CTA scope) on purpose: what
// these barriers order is TMEM traffic (tcgen05.fence + the barrier) and async-proxy copies, never generic-proxy
// memory of another CTA; the .release.cluster form costs MEMBAR.ALL.GPU per arrive and .acquire.cluster waits emit
// CCTL.IVALL (a full L1 invalidation) per success — measured 1.19 -> 1.46 ms on the maf_mnist step.
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// ---- TMEM allocation, pair-collective: the same warp of BOTH CTAs of the pair executes these ----------------
__device__ __forceinline__ void tmem_alloc2(uint32_t* result_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(result_smem)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

// ---- tensor-map TMA, pair form: the copy lands in THIS CTA's shared memory (and, with a mask, at the same offset
// in every CTA of the mask); its bytes complete on the mbarrier `mbar_cluster` (a shared::cluster address — the
// leader CTA's barrier of each destination pair) ------------------------------------------------------------------
__device__ __forceinline__ void tma2_load_2d(const CUtensorMap* map, uint32_t dst_smem, uint32_t mbar_cluster, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst_smem),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(mbar_cluster), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma2_load_2d_mc(const CUtensorMap* map, uint32_t dst_smem, uint32_t mbar_cluster, int c0, int c1,
                                                uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%4, %5}], [%2], "
      "%3;" ::"r"(dst_smem),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(mbar_cluster), "h"(mask), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma2_load_3d(const CUtensorMap* map, uint32_t dst_smem, uint32_t mbar_cluster, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
          dst_smem),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(mbar_cluster), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}

// ---- pair MMA: D[256 x N] (128 rows in each CTA's TMEM) (+)= A[256 x 16] · B[N x 16]^T; issued by ONE thread of the
// leader CTA, operands at the same shared-memory offsets in both CTAs (each holds its 128 rows of A, its N/2 rows of B)
__device__ __forceinline__ void mma2_ss_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// all MMAs issued so far arrive (once complete) on the mbarrier at this offset in every CTA of `mask`
__device__ __forceinline__ void mma2_commit_mc(uint64_t* bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
               "h"(mask)
               : "memory");
}
// the same two, executed by the WHOLE converged warp and issued by its elected lane (tc::mma_ss_warp: inside single-thread
// code every tcgen05.mma is wrapped in an ELECT / BRA.U.ANY loop and its uniform operands are rebuilt)
__device__ __forceinline__ void mma2_ss_warp(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\telect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma2_commit_mc_warp(uint64_t* bar, uint16_t mask) {
  asm volatile(
      "{\n\t.reg .pred q;\n\telect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n\t}" ::"r"(smem_u32(bar)),
      "h"(mask)
      : "memory");
}

}  // namespace tc

constexpr int kDt2AStage = 2 * kDtABytes;            // A of one CTA and K stage: [hi|lo][4 chunks][128 rows][16 B] = 16 KB
constexpr int kDt2BStage = 2 * kDtABytes;            // this CTA's half of B: same shape = 16 KB
constexpr int kDt2StageBytes = kDt2AStage + kDt2BStage;  // 32 KB

// CL = CTAs per cluster: 2 (one pair) or 4 (two pairs on adjacent n tiles sharing A by multicast)
template <int EPI, bool WIDE, int CL> struct Dt2Cfg {
  static constexpr int kEpiWarps = WIDE ? 8 : 4;
  static constexpr int kThreads = 64 + 2 * 32 * kEpiWarps;
  static constexpr int kStages = EPI == kEpiAffine ? 4 : 6;
  static constexpr int kScratch = EPI == kEpiAffine ? 2 * kDtTileScratchFloats : 0;
  static constexpr size_t kSmem = (size_t)kStages * kDt2StageBytes + kScratch * 4 + 2 * kDtBN * 4 + (2 * kStages + 4) * 8 + 16 + 1024;
};

// tile walk of a cluster: quad index q -> (pair-row mtp, n tile of THIS pair).  With CL = 4 the two pairs take n
// tiles 2*j and 2*j + 1; a pair whose n tile does not exist (odd NT) still runs the pipeline on a clamped tile
// (its A halves are needed by the other pair) and only skips the epilogue (`live` = false).
struct Dt2Tile {
  int mtp, nt;
  bool live;
  int ks;  // K stages this tile walks (masked Linears: fewer than the image holds)
};
__device__ __forceinline__ int dt2_ks_of(const DtGemm& a, int nt) {
  return (nt < kDtMaxKsTiles && a.ks_tile[nt] != 0) ? (int)a.ks_tile[nt] : a.KS;
}
template <int CL>
__device__ __forceinline__ Dt2Tile dt2_tile(const DtGemm& a, int q, int NT, int pair) {
  Dt2Tile t;
  if (CL == 2) {
    t.mtp = q / NT;
    t.nt = q - t.mtp * NT;
    t.live = true;
    t.ks = dt2_ks_of(a, t.nt);
  } else {
    const int NT2 = (NT + 1) >> 1;
    t.mtp = q / NT2;
    const int n0 = 2 * (q - t.mtp * NT2), n1 = min(n0 + 1, NT - 1);
    const int nt = n0 + pair;
    t.live = nt < NT;
    t.nt = t.live ? nt : NT - 1;
    // both pairs walk the same number of stages (they share the A stages): the larger of the two tiles' needs —
    // the extra stages of the other tile hold masked (zero) weights
    t.ks = max(dt2_ks_of(a, n0), dt2_ks_of(a, n1));
  }
  return t;
}

template <int EPI, bool INV, bool WIDE, int CL>
__global__ void __launch_bounds__(Dt2Cfg<EPI, WIDE, CL>::kThreads, 1)
dense_tc2_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, DtGemm a, int MTP, int NT,
                      int b_blk0) {
  using Cfg = Dt2Cfg<EPI, WIDE, CL>;
  constexpr int S = Cfg::kStages;
  constexpr int EW = Cfg::kEpiWarps;
  constexpr int kLoadWarp = 2 * EW, kMmaWarp = 2 * EW + 1;  // issuer and producer at the highest warp ids (dense_tc.cuh)
  extern __shared__ uint8_t smem_raw[];
  // the dynamic window is only 16-byte aligned by contract: TMA destinations want 128 B, keep 1 KB
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* stages = smem;                                                     // [S][A 16 KB | B 16 KB]
  float* Tbase = reinterpret_cast<float*>(smem + S * kDt2StageBytes);         // epilogue scratch (affine)
  float* sbias2 = Tbase + Cfg::kScratch;                                      // [2][256]
  uint64_t* full = reinterpret_cast<uint64_t*>(sbias2 + 2 * kDtBN);           // [S]  (used in the pair leader)
  uint64_t* empty = full + S;                                                 // [S]  (every CTA)
  uint64_t* acc_full = empty + S;                                             // [2]  (every CTA)
  uint64_t* acc_empty = acc_full + 2;                                         // [2]  (used in the pair leader)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t crank = tc::cluster_ctarank();   // 0 .. CL-1
  const uint32_t prank = crank & 1u;              // rank inside the pair: 0 = leader (issues the MMAs)
  const uint32_t pair = crank >> 1;               // pair inside the cluster
  const uint32_t leader = crank & ~1u;            // cluster rank of this pair's leader
  const int nq = CL == 2 ? MTP * NT : MTP * ((NT + 1) >> 1);
  const int q0 = (int)tc::cluster_id_x(), qstep = (int)tc::nclusters_x();

  if (tid == kLoadWarp * 32) {
    tc::tma_prefetch_desc(&tmA);
    tc::tma_prefetch_desc(&tmB);
    for (int s = 0; s < S; ++s) {
      tc::mbar_init(&full[s], 1);            // the leader's own arrive.expect_tx; every copy of the pair completes here
      tc::mbar_init(&empty[s], CL / 2);      // one tcgen05.commit per pair of the cluster (A halves are shared)
    }
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(&acc_full[s], 1);
      tc::mbar_init(&acc_empty[s], 4 * EW);  // both epilogue groups of BOTH CTAs of the pair drain every accumulator
    }
    tc::fence_mbar_init();
  }
  if (warp == kMmaWarp) tc::tmem_alloc2(tmem_slot, 2 * kDtBN);
  tc::tc_fence_before();
  __syncthreads();
  tc::cluster_sync_all();  // every CTA's barriers exist before any remote arrive / multicast can reach them
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  pdl_wait();  // from here on: operands written by the previous kernel of the chain (common.cuh, PDL)
  pdl_launch_dependents();

  if (warp == kLoadWarp) {
    if (lane == 0) {  // ===== TMA producer (every CTA)
      const uint32_t full_leader0 = tc::mapa_u32(tc::smem_u32(&full[0]), leader);
      uint32_t it = 0;
      for (int q = q0; q < nq; q += qstep) {
        const Dt2Tile tl = dt2_tile<CL>(a, q, NT, (int)pair);
        const int mt = 2 * tl.mtp + (int)prank;
        const int KS = tl.ks;
        for (int ks = 0; ks < KS; ++ks, ++it) {
          const uint32_t st = it % S, ph = (it / S) & 1u;
          tc::mbar_wait(&empty[st], ph ^ 1u);  // passes at once on a fresh barrier
          const uint32_t dst = tc::smem_u32(stages + st * kDt2StageBytes);
          const uint32_t bar = full_leader0 + st * 8u;
          if (prank == 0) tc::mbar_arrive_expect_tx(&full[st], 2 * kDt2StageBytes);
          const int ablk = (mt * a.KS + ks) * 8;  // 2 KB blocks: [hi: 4 chunks][lo: 4 chunks]
          if (CL == 2) {
            tc::tma2_load_2d(&tmA, dst, bar, 0, ablk);
          } else {
            // this CTA fetches the hi (pair 0) or lo (pair 1) half of the A stage and multicasts it to the CTA with the
            // same pair rank in the other pair: ranks {prank, prank + 2}
            const uint16_t mask = (uint16_t)((1u << prank) | (1u << (prank + 2)));
            tc::tma2_load_2d_mc(&tmA, dst + pair * kDtABytes, bar, 0, ablk + 4 * (int)pair, mask);
          }
          tc::tma2_load_3d(&tmB, dst + kDt2AStage, bar, 0, (int)prank, b_blk0 + (tl.nt * a.KS + ks) * 8);
        }
      }
    }
  } else if (warp == kMmaWarp) {
    if (prank == 0) {  // ===== MMA issuer (pair leader): 3 passes x 2 K steps per stage, M = 256 over the pair; the whole
                       // warp runs the loop converged, its elected lane issues (as in dense_tc.cuh)
      constexpr uint32_t idesc = tc::make_idesc(tc::kFmtBF16, 256, kDtBN);
      constexpr uint32_t lbo = 16u * 128u;  // both operands are 128-row images in each CTA
      constexpr uint16_t kAll = (uint16_t)((1u << CL) - 1u);
      const uint16_t kPair = (uint16_t)(3u << leader);
      uint32_t it = 0, tl_i = 0;
      for (int q = q0; q < nq; q += qstep, ++tl_i) {
        const Dt2Tile tl = dt2_tile<CL>(a, q, NT, (int)pair);
        const int KS = tl.ks;
        const uint32_t slot = tl_i & 1u, aph = (tl_i >> 1) & 1u;
        tc::mbar_wait(&acc_empty[slot], aph ^ 1u);  // both CTAs' epilogues have drained this accumulator
        tc::tc_fence_after();
        const uint32_t dtm = tmem + slot * (uint32_t)kDtBN;
        uint32_t acc = 0;
        for (int ks = 0; ks < KS; ++ks, ++it) {
          const uint32_t st = it % S, ph = (it / S) & 1u;
          tc::mbar_wait(&full[st], ph);
          tc::tc_fence_after();
          const uint32_t sa = tc::smem_u32(stages + st * kDt2StageBytes), sb = sa + kDt2AStage;
          // one descriptor pair per stage; the six MMAs add their operand offsets (>> 4) to the start-address field
          const uint64_t da = tc::make_smem_desc(sa, lbo, 128u), db = tc::make_smem_desc(sb, lbo, 128u);
#pragma unroll
          for (int pass = 0; pass < 3; ++pass) {
            const uint32_t ao = pass == 2 ? (uint32_t)kDtABytes : 0u, bo = pass == 1 ? (uint32_t)kDtABytes : 0u;
#pragma unroll
            for (int s = 0; s < kDtBK / 16; ++s) {
              tc::mma2_ss_warp(dtm, da + ((ao + s * 2u * lbo) >> 4), db + ((bo + s * 2u * lbo) >> 4), idesc, acc);
              acc = 1;
            }
          }
          tc::mma2_commit_mc_warp(&empty[st], kAll);  // the stage is free in every CTA of the cluster once these MMAs are done
        }
        tc::mma2_commit_mc_warp(&acc_full[slot], kPair);
      }
    }
  } else {
    // ===== epilogue warps (every CTA).  As in dense_tc.cuh both groups work on EVERY tile, each on its half of the
    // accumulator's columns (an eight-warp group splits its half again), so that the MMAs of tile i+1 run under the
    // epilogue of tile i: per-tile time max(MMA, epilogue/2).
    const uint32_t group = (uint32_t)warp / EW;
    const int sub = (warp >> 2) & 1;  // EW == 8 only
    const int et = tid % (EW * 32);
    auto group_sync = [&]() {
      if (group == 0) asm volatile("bar.sync 1, %0;" ::"n"(EW * 32) : "memory");
      else asm volatile("bar.sync 2, %0;" ::"n"(EW * 32) : "memory");
    };
    float* sbias = sbias2 + group * kDtBN;
    float* Tw = Tbase + ((int)group * EW + warp % EW) * 32 * 36;
    const int span = EW == 8 ? 64 : 128;
    const int c_lo = ((int)group * (EW == 8 ? 2 : 1) + (EW == 8 ? sub : 0)) * span, c_hi = c_lo + span;
    const uint32_t acc_empty_leader0 = tc::mapa_u32(tc::smem_u32(&acc_empty[0]), leader);
    uint32_t tl_i = 0;
    for (int q = q0; q < nq; q += qstep, ++tl_i) {
      const Dt2Tile tl = dt2_tile<CL>(a, q, NT, (int)pair);
      const int mt = 2 * tl.mtp + (int)prank, nt = tl.nt;
      const uint32_t slot = tl_i & 1u, aph = (tl_i >> 1) & 1u;
      if (tl.live) {
        for (int n = et; n < kDtBN; n += EW * 32) {  // bias of this tile's columns
          float b = 0.f;
          if (EPI == kEpiHidden) {
            const int col = nt * kDtBN + n;
            if (col < a.n_out) b = a.bias[col];
          } else {
            b = dt_tile_bias(a, nt, n);
            if (EPI == kEpiRqs8 && (n & 127) % 25 < 16) b *= kLog2e;  // bin logits reach the spline in log2 units
          }
          sbias[n] = b;
        }
      }
      group_sync();
      DtRqsX xin;
      int rd_lo = 0, rd_hi = 0;
      if (EPI == kEpiRqs8) dt_rqs_group_dims((int)tl_i, (int)group, rd_lo, rd_hi);
      if (EPI == kEpiRqs8 && tl.live) dt_rqs_load_x(a, mt, nt, rd_lo, rd_hi, warp, lane, xin);
      if (tl.live && EPI == kEpiAffine) {
        // pull this thread's row of the warp's state slice into L2 under the tile's MMAs
        const int64_t prow = (int64_t)mt * 128 + (warp & 3) * 32 + lane;
        const int dlo = nt * a.dpt + c_lo / a.ppd, nd = min(span / a.ppd, a.ndims - dlo);
        if (prow < a.N && nd > 0) {
          const float* p0 = a.v + prow * a.D + a.tp + a.dstride * dlo;
          const uintptr_t q1 = (uintptr_t)(p0 + a.dstride * (nd - 1));
          for (uintptr_t qq = (uintptr_t)p0 & ~(uintptr_t)127; qq <= q1; qq += 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(qq));
        }
      }
      tc::mbar_wait(&acc_full[slot], aph);
      tc::tc_fence_after();
      if (tl.live) {
        const uint32_t tb = tmem + slot * (uint32_t)kDtBN + ((uint32_t)((warp & 3) * 32) << 16);
        if (EPI == kEpiRqs8) dt_epilogue_rqs_tile<INV>(a, mt, nt, rd_lo, rd_hi, 2 * (int)group, 0, tb, sbias, warp, lane, xin);
        else dt_epilogue<EPI, INV>(a, mt, nt, tb, sbias, Tw, warp, lane, c_lo, c_hi, 0);
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive_cluster(acc_empty_leader0 + slot * 8u);
      group_sync();  // sbias is rewritten by the group's next tile
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::cluster_sync_all();  // no CTA leaves (or frees TMEM) while a peer can still signal its barriers or read its operands
  if (warp == kMmaWarp) {
    __syncwarp();
    tc::tmem_dealloc2(tmem, 2 * kDtBN);
  }
}

// ---- host side: tensor maps over the tile images ------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static inline PFN_encodeTiled dt2_encoder() {
  static PFN_encodeTiled fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  }
  return fn;
}

// image = `blocks` consecutive 2 KB blocks ([128 rows][16 B] of one K chunk of one of hi/lo); box = 8 blocks = one
// operand stage of a 128-row tile (16 KB); boxes of 4 blocks (one of hi/lo) for the multicast halves.
static inline int dt2_map_a(CUtensorMap* m, const void* img, uint64_t blocks, uint32_t box_blocks) {
  PFN_encodeTiled enc = dt2_encoder();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled is not available from this driver");
    return FLOWCUDA_ECUDA;
  }
  const cuuint64_t dims[2] = {256, blocks};
  const cuuint64_t strides[1] = {2048};
  const cuuint32_t box[2] = {256, box_blocks};
  const cuuint32_t es[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, const_cast<void*>(img), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(A image, %llu blocks) failed with %d", (unsigned long long)blocks, (int)r);
    return FLOWCUDA_ECUDA;
  }
  return FLOWCUDA_OK;
}
// weight image: [tile][k stage][hi|lo][4 chunks][256 rows][16 B]: per chunk two 2 KB halves (rows 0-127 / 128-255);
// dims {2 KB, half, chunk-blocks}; box = one half x 8 chunk-blocks = this CTA's 16 KB of a stage.
static inline int dt2_map_b(CUtensorMap* m, const void* img, uint64_t chunk_blocks) {
  PFN_encodeTiled enc = dt2_encoder();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled is not available from this driver");
    return FLOWCUDA_ECUDA;
  }
  const cuuint64_t dims[3] = {256, 2, chunk_blocks};
  const cuuint64_t strides[2] = {2048, 4096};
  const cuuint32_t box[3] = {256, 1, 8};
  const cuuint32_t es[3] = {1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, const_cast<void*>(img), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(B image, %llu chunk blocks) failed with %d", (unsigned long long)chunk_blocks, (int)r);
    return FLOWCUDA_ECUDA;
  }
  return FLOWCUDA_OK;
}

// Description of one launch: the A image holds `a_mt` 128-row tiles (EVEN: a pair always reads two), the weight image of
// this Linear holds `b_tiles` n tiles in total (all flow layers), of which this launch reads tiles [b_tile0, b_tile0 + NT).
struct Dt2Launch {
  int64_t a_mt;
  int64_t b_tiles;
  int64_t b_tile0;
  int prefer;  // kernel of this launch when FLOWCUDA_DENSE_TC is unset: 2 = CTA pairs, anything else = single CTAs (dt2_mode)
};

template <int EPI, bool INV, bool WIDE, int CL>
static int dt2_launch_w(const DtGemm& g, const Dt2Launch& L, int64_t MT, int NT, cudaStream_t s) {
  using Cfg = Dt2Cfg<EPI, WIDE, CL>;
  auto kern = dense_tc2_gemm_kernel<EPI, INV, WIDE, CL>;
  // one-time per template instance: the shared-memory opt-in and how many clusters the device really keeps
  // resident (GPC sizes can strand SMs for 4-CTA clusters; a persistent walk must not launch more than fit)
  FC_ENSURE_SMEM(kern, Cfg::kSmem);
  static int resident = 0;
  if (resident == 0) {
    cudaLaunchConfig_t qc = {};
    qc.gridDim = dim3((unsigned)(sm_count() / CL * CL));
    qc.blockDim = dim3(Cfg::kThreads);
    qc.dynamicSmemBytes = Cfg::kSmem;
    cudaLaunchAttribute qa[1];
    qa[0].id = cudaLaunchAttributeClusterDimension;
    qa[0].val.clusterDim.x = CL;
    qa[0].val.clusterDim.y = 1;
    qa[0].val.clusterDim.z = 1;
    qc.attrs = qa;
    qc.numAttrs = 1;
    int nc = 0;
    FC_CUDA(cudaOccupancyMaxActiveClusters(&nc, kern, &qc));
    if (nc < 1) {
      set_error("dense_tc2: no %d-CTA cluster of this kernel fits the device", CL);
      return FLOWCUDA_ECUDA;
    }
    resident = nc;
  }
  CUtensorMap tmA, tmB;
  int rc = dt2_map_a(&tmA, g.a_img, (uint64_t)L.a_mt * g.KS * 8, CL == 2 ? 8u : 4u);
  if (rc) return rc;
  rc = dt2_map_b(&tmB, g.b_img - (size_t)L.b_tile0 * g.KS * 2 * kDtBBytes, (uint64_t)L.b_tiles * g.KS * 8);
  if (rc) return rc;
  const int MTP = (int)((MT + 1) / 2);
  const int64_t nq = CL == 2 ? (int64_t)MTP * NT : (int64_t)MTP * ((NT + 1) / 2);
  const int clusters = (int)(nq < resident ? nq : resident);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(clusters * CL));
  cfg.blockDim = dim3(Cfg::kThreads);
  cfg.dynamicSmemBytes = Cfg::kSmem;
  cfg.stream = s;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = CL;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  FC_CUDA(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, g, MTP, NT, (int)(L.b_tile0 * g.KS * 8)));
  return FLOWCUDA_OK;
}

// Which kernel a launch takes.  FLOWCUDA_DENSE_TC = "pair" / "quad" / "1cta" forces the CTA-pair kernel, the 4-CTA
// multicast clusters or the single-CTA kernel (dense_tc.cuh) for every launch (tests/test_gpu_tc.py runs the parity cases
// through all of them); unset, the caller's preference (Dt2Launch::prefer) decides.  Measured on B200 at the end of
// round 2 (profiles/r02/gemm_modes_final.log), once the single-CTA kernel's issue loop and epilogue had been brought to
// the MMA rate: the RQS transform Linear [n,512]x[512,9800] reads 64 B per clock and SM as single CTAs — 9.5 KB per
// clock over the chip, above what the L2 slices deliver — and 43 B as pairs (each CTA stages half of B): pairs for that
// launch alone run the maf_mnist step in 0.914 instead of 0.940 ms (batch 32768: 6.47 vs 6.92 ms).  The affine output
// Linear of deep16_nvp is slower as pairs (1.67 vs 1.55 ms: a 784-column output is 4 n tiles, pairs halve the
// parallelism), the hidden Linears gain as pairs only where the cluster chain kernel does not run (batch 16384 /
// 32768: +2-3 %), and the MADE Linears (masked K ranges) were slower as pairs all round: those callers prefer 1.
static inline int dt2_mode(int prefer = 1) {
  // read on every call (a getenv is nothing next to a launch) so that tests can switch kernels in-process
  const char* e = getenv("FLOWCUDA_DENSE_TC");
  if (e && e[0] == 'p') return 2;
  if (e && e[0] == 'q') return 4;
  if (e && e[0] == '1') return 1;
  return prefer == 2 ? 2 : 1;
}

template <int EPI, bool INV>
static int dt2_launch_gemm(const DtGemm& g, const Dt2Launch& L, int64_t MT, int NT, cudaStream_t s) {
  const int mode = dt2_mode(L.prefer);
  if (mode == 1) return dt_launch_gemm<EPI, INV>(g, MT, NT, s);
  const bool quad = mode == 4 && NT >= 2;
  const int64_t MTP = (MT + 1) / 2;
  if (quad) {
    const bool wide = EPI != kEpiRqs8 && MTP * ((NT + 1) / 2) > sm_count() / 4;
    if (EPI != kEpiRqs8 && wide) return dt2_launch_w<EPI, INV, EPI != kEpiRqs8, 4>(g, L, MT, NT, s);
    return dt2_launch_w<EPI, INV, false, 4>(g, L, MT, NT, s);
  }
  const bool wide = EPI != kEpiRqs8 && MTP * NT > sm_count() / 2;
  if (EPI != kEpiRqs8 && wide) return dt2_launch_w<EPI, INV, EPI != kEpiRqs8, 2>(g, L, MT, NT, s);
  return dt2_launch_w<EPI, INV, false, 2>(g, L, MT, NT, s);
}

}  // namespace flowcuda
