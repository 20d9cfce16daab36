// common.cuh — shared helpers for libflowcuda (sm_100a only).
#pragma once
#include <mutex>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>

#include "../../include/flowcuda.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libflowcuda is written for sm_100a (B200) only"
#endif

namespace flowcuda {

// ---- error plumbing (thread-local message, C return codes) -------------------------
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);
int sm_count();

#define FC_REQUIRE(cond, ...)                 \
  do {                                        \
    if (!(cond)) {                            \
      flowcuda::set_error(__VA_ARGS__);       \
      return FLOWCUDA_EINVAL;                 \
    }                                         \
  } while (0)

#define FC_CUDA(call)                                              \
  do {                                                             \
    cudaError_t e__ = (call);                                      \
    if (e__ != cudaSuccess) return flowcuda::cuda_fail(e__, #call); \
  } while (0)

#define FC_LAUNCH_CHECK(name)                                       \
  do {                                                              \
    cudaError_t e__ = cudaGetLastError();                           \
    if (e__ != cudaSuccess) return flowcuda::cuda_fail(e__, name);  \
  } while (0)

// Opt a kernel in to `bytes` of dynamic shared memory ONCE per device (and again only if a later call needs more):
// cudaFuncSetAttribute is a driver round trip that does not belong on the per-launch path ("no global
// mutable state except one-time cudaFuncSetAttribute").  The cache is per call site (= per kernel instantiation)
// and per device; racing first calls both set the same value.
#define FC_ENSURE_SMEM(kern, bytes)                                                                              \
  do {                                                                                                           \
    static int fc_smem_set__[32] = {0};                                                                          \
    int fc_dev__ = 0;                                                                                            \
    FC_CUDA(cudaGetDevice(&fc_dev__));                                                                           \
    const int fc_need__ = (int)(bytes);                                                                          \
    if (fc_dev__ < 0 || fc_dev__ >= 32 || fc_need__ > fc_smem_set__[fc_dev__]) {                                 \
      FC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, fc_need__));               \
      if (fc_dev__ >= 0 && fc_dev__ < 32) fc_smem_set__[fc_dev__] = fc_need__;                                   \
    }                                                                                                            \
  } while (0)

// ---- programmatic dependent launch (PDL) ---------------------------------------------------------------------------
// The D=784 paths are chains of 40-100 short kernels, each depending on the one before.  Launched with
// cudaLaunchAttributeProgrammaticStreamSerialization, kernel k+1 is scheduled while kernel k is still running: its CTAs
// run their prologue (barrier init, TMEM allocation, descriptor prefetch) on SMs as they free up and then block in
// pdl_wait() until kernel k has COMPLETED and its writes are visible.  Rules kept by every kernel launched this way:
// nothing before pdl_wait() reads or writes global memory another kernel of the chain touches, and
// pdl_launch_dependents() comes after pdl_wait(), so at most one successor is ever resident early.
// FLOWCUDA_PDL=0 launches the same kernels with plain stream order (the waits are then no-ops).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool pdl_enabled();  // core.cu

template <typename... KArgs, typename... Args>
inline cudaError_t launch_chain(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr float kHalfLog2Pi = 0.9189385332046727f;  // 0.5*log(2*pi)

// ---- streaming 128-bit global access (inputs are read once, outputs written once) ---
__device__ __forceinline__ float4 ld_stream_f4(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ float2 ld_stream_f2(const float2* p) {
  float2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f32 {%0,%1}, [%2];" : "=f"(r.x), "=f"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ float ld_stream_f1(const float* p) {
  float r;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream_f4(float4* p, float4 v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w));
}
__device__ __forceinline__ void st_stream_f2(float2* p, float2 v) {
  asm volatile("st.global.L1::no_allocate.v2.f32 [%0], {%1,%2};" ::"l"(p), "f"(v.x), "f"(v.y));
}
__device__ __forceinline__ void st_stream_f1(float* p, float v) {
  asm volatile("st.global.L1::no_allocate.f32 [%0], %1;" ::"l"(p), "f"(v));
}

// ---- Blackwell packed fp32x2 arithmetic (FFMA2 / FMUL2 / FADD2 in SASS) ---------------
// sm_100 adds {fma,mul,add}.f32x2: one issue slot, two IEEE fp32 results.
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  float2 d;
  asm("{\n\t.reg .b64 ra, rb, rc, rd;\n\t"
      "mov.b64 ra, {%2,%3};\n\tmov.b64 rb, {%4,%5};\n\tmov.b64 rc, {%6,%7};\n\t"
      "fma.rn.f32x2 rd, ra, rb, rc;\n\t"
      "mov.b64 {%0,%1}, rd;\n\t}"
      : "=f"(d.x), "=f"(d.y)
      : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
  return d;
}
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
  float2 d;
  asm("{\n\t.reg .b64 ra, rb, rd;\n\t"
      "mov.b64 ra, {%2,%3};\n\tmov.b64 rb, {%4,%5};\n\t"
      "mul.rn.f32x2 rd, ra, rb;\n\t"
      "mov.b64 {%0,%1}, rd;\n\t}"
      : "=f"(d.x), "=f"(d.y)
      : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return d;
}
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
  float2 d;
  asm("{\n\t.reg .b64 ra, rb, rd;\n\t"
      "mov.b64 ra, {%2,%3};\n\tmov.b64 rb, {%4,%5};\n\t"
      "add.rn.f32x2 rd, ra, rb;\n\t"
      "mov.b64 {%0,%1}, rd;\n\t}"
      : "=f"(d.x), "=f"(d.y)
      : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return d;
}
__device__ __forceinline__ float2 splat2(float a) { return make_float2(a, a); }

// The same instructions on values that STAY packed: f2_t is one 64-bit register pair holding two fp32.
// Keeping operands in this form end to end lets ptxas allocate aligned pairs once; the float2 helpers
// above re-pack on every call, which costs one MOV per operand in hot loops.
typedef unsigned long long f2_t;
__device__ __forceinline__ f2_t f2_pack(float lo, float hi) {
  f2_t r;
  asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ f2_t f2_splat(float a) { return f2_pack(a, a); }
__device__ __forceinline__ void f2_unpack(f2_t v, float& lo, float& hi) {
  asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f2_t f2_fma(f2_t a, f2_t b, f2_t c) {
  f2_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ f2_t f2_mul(f2_t a, f2_t b) {
  f2_t d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ f2_t f2_add(f2_t a, f2_t b) {
  f2_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

// ---- scalar math with controlled accuracy -----------------------------------------
// exp via MUFU.EX2 (abs err ~2 ulp on the scaled argument; rel err ~1e-7·|x| — inside the
// 1e-5 parity budget for |x| <~ 30, which log-scales of a trained flow satisfy).
__device__ __forceinline__ float fast_exp(float x) { return exp2f(x * kLog2e); }
// bare MUFU.EX2 (no denormal-range fix-up code around it; results below 2^-126 flush to zero)
__device__ __forceinline__ float ex2_ftz(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float softplus_f(float x) {
  // log1p(exp(x)) evaluated as max(x,0) + log1p(exp(-|x|)) (no overflow, full fp32 accuracy)
  return fmaxf(x, 0.f) + log1pf(expf(-fabsf(x)));
}
__device__ __forceinline__ float leaky_relu_f(float x) { return fmaxf(x, 0.01f * x); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Persistent grid for a grid-stride kernel: exactly the CTAs that are resident at once (SMs x occupancy),
// so there is no partial last wave; fewer when the work does not fill them.
template <typename K>
inline int grid_resident(K kernel, int threads, size_t smem, int64_t work_items, int per_block) {
  int occ = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, threads, smem) != cudaSuccess || occ < 1) occ = 1;
  int64_t need = (work_items + per_block - 1) / per_block;
  int64_t cap = (int64_t)sm_count() * occ;
  if (need < 1) need = 1;
  return (int)(need < cap ? need : cap);
}

inline int grid_for(int64_t work_items, int per_block, int waves_cap = 8) {
  int64_t need = (work_items + per_block - 1) / per_block;
  int64_t cap = (int64_t)sm_count() * waves_cap;
  if (need < 1) need = 1;
  return (int)(need < cap ? need : cap);
}

// ---- the library's one internal stream ----------------------------------------------------------------------------
// Work that nothing downstream in the caller's stream depends on for a while (the weight-gradient jobs of the training
// step) runs on ONE non-blocking stream per device, forked from and joined back into the caller's stream with events —
// all created on first use and kept for the life of the process.  The events are reused from call to call, so host
// threads enqueue such calls one at a time per process (side_mutex()).  FLOWCUDA_TRAIN_OVERLAP=0 keeps everything on
// the caller's stream.  (Measured and NOT adopted: packing the weight images of the later flow layers of a density
// call on this stream beside the GEMM chain — maf_mnist 1.019 vs 1.014 ms, deep16_nvp 1.60 vs 1.54 ms: the HBM pass
// slows the chain it runs beside by as much as it saves.)
struct SideCtx {
  cudaStream_t q;
  cudaEvent_t ev_m, ev_last, ev_s[2];
};
int side_ctx_get(SideCtx** out);
std::recursive_mutex& side_mutex();
bool side_env_enabled(const char* name);

// one 256-bit store (sm_100: STG.256): eight 32-bit words to a 32-byte aligned address
__device__ __forceinline__ void st_global_v8(float* p, const uint32_t* w) {
  // streamed: nothing re-reads the line soon, and it must not push the GEMM's L2-resident operand images out
  asm volatile("st.global.L1::no_allocate.L2::evict_first.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]), "r"(w[4]),
               "r"(w[5]), "r"(w[6]), "r"(w[7])
               : "memory");
}

// Training forward through the density path's GEMMs (coupling_dense_tc.cu) — what coupling_dense_grad.cu asks the
// epilogues to keep for the reverse mode.  Per flow layer l one activation block of `per_layer_act` floats:
//   [0, N*D)            the layer's input state v
//   act_h[i], i < nh    relu(h_{i+1}) as fp32 [N, hidden_i]
//   act_h[nh]           the last Linear's RAW accumulators (no bias), 32 floats per transformed dim, fp32 [N, ldo]
struct DtTrainStores {
  float* act;
  size_t per_layer_act;
  size_t act_h[FLOWCUDA_MAX_HIDDEN_LAYERS + 1];
  int64_t ldo;
};
bool coupling_dense_tc_supported(const flowcuda_coupling_spec* sp);
size_t coupling_dense_tc_workspace_bytes(const flowcuda_coupling_spec* sp, int64_t n);
int64_t coupling_dense_tc_store_ldo(const flowcuda_coupling_spec* sp);
// data -> base pass over all layers with the stores above; out_z [N, D] and out_ld [N] (sum of the log-dets)
int coupling_dense_tc_train_forward(const flowcuda_coupling_spec* sp, const float* params, const float* x, int64_t n,
                                    const DtTrainStores& tr, float* out_z, float* out_ld, void* workspace,
                                    size_t workspace_bytes, cudaStream_t s);

}  // namespace flowcuda
