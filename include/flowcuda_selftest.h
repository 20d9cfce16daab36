/*
 * flowcuda_selftest.h — C ABI of libflowcuda_selftest.so: diagnostics of the hand-written tcgen05 primitives
 * (csrc/tc.cuh).  TEST LIBRARY: built next to libflowcuda.so, loaded only by tests/ and tools/; the product
 * library does not contain these kernels.  Conventions as in flowcuda.h (device pointers, `stream` = cudaStream_t,
 * 0 or a negative FLOWCUDA_E* code; the message comes from this library's own flowcuda_selftest_last_error()).
 */
#ifndef FLOWCUDA_SELFTEST_H_
#define FLOWCUDA_SELFTEST_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* flowcuda_selftest_last_error(void);

/* ------------------------------------------------------------------------------------
 * Diagnostics — one 128-row tile through every tensor-core primitive the flow kernels use
 * (tile-image staging, TMA bulk copy, tcgen05.mma kind::tf32 / kind::f16 in 1- and 3-pass
 * form, tcgen05.commit, tcgen05.ld at aligned and unaligned columns).
 *   d[128,n] = a[128,k] @ w[k,n];  mode 0 tf32x3, 1 bf16x3, 2 tf32, 3 bf16.
 * d_unaligned (optional, [128,n]) receives columns [0, 25*(n/25)) re-read in 25-column
 * groups at unaligned TMEM offsets.
 * ---------------------------------------------------------------------------------- */
size_t flowcuda_selftest_umma_workspace_bytes(int n, int k);
int flowcuda_selftest_umma(const float* a, const float* w, float* d, float* d_unaligned, int n, int k,
                           int mode, void* workspace, size_t workspace_bytes, void* stream);

/* Microbenchmark: device cycles of chains of small tcgen05.mma (M128, K8, tf32) for several N,
 * 1-3 independent accumulators, A from TMEM or shared memory.  out_device: int64 records
 * {N, nacc*1000 + n_mma (+100000: A from smem), issue cycles, cycles until the commit arrives},
 * terminated by -1. */
int flowcuda_selftest_mma_timing(long long* out_device, int capacity, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FLOWCUDA_SELFTEST_H_ */
