// densities.cu — the data step of the training loops, on the device.
//
// The reference draws every training batch on the host, through a Python generator around
// jax.random / tfd (src/densities/banana.py:31-41, src/densities/energies.py:47-127,
// src/densities/gaussian_blobs.py:34-59) and dequantises MNIST bytes with one more host
// uniform draw per step (src/maf_mnist.py:101-106).  Here a batch is a pure function of
// (seed, sample index): Philox4x32-10 keyed by the seed, counter = (index, stream), so a batch can
// be regenerated, sharded over ranks by index range, and needs no state.  JAX's threefry streams
// are not reproduced (JAX is not available) — the DISTRIBUTIONS are the reference's,
// the bit streams are this library's, restated by oracle/philox_oracle.py.
#include "common.cuh"

namespace flowcuda {

__host__ __device__ __forceinline__ void philox_round(uint32_t& c0, uint32_t& c1, uint32_t& c2, uint32_t& c3,
                                                      uint32_t k0, uint32_t k1) {
  const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
  const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
  const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
  const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
  c1 = (uint32_t)p1;
  c3 = (uint32_t)p0;
  c0 = n0;
  c2 = n2;
}

// Philox4x32-10 (Salmon et al. 2011): counter (c0..c3), key (k0,k1) -> four 32-bit words.
__host__ __device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                       uint32_t k0, uint32_t k1, uint32_t r[4]) {
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    philox_round(c0, c1, c2, c3, k0, k1);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  r[0] = c0; r[1] = c1; r[2] = c2; r[3] = c3;
}

// 24 random bits -> [0,1) exactly representable; the half-offset form lies in (0,1) (safe under log).
__device__ __forceinline__ float u01_closed_open(uint32_t r) { return (float)(r >> 8) * 5.9604644775390625e-8f; }
__device__ __forceinline__ float u01_open(uint32_t r) { return ((float)(r >> 8) + 0.5f) * 5.9604644775390625e-8f; }

// Box-Muller on two words: radius from r0, angle from r1.
__device__ __forceinline__ void normal_pair(uint32_t r0, uint32_t r1, float& n0, float& n1) {
  const float rad = sqrtf(-2.f * logf(u01_open(r0)));
  float sn, cs;
  sincospif(2.f * u01_closed_open(r1), &sn, &cs);
  n0 = rad * cs;
  n1 = rad * sn;
}

constexpr float kPi = 3.14159265358979323846f;

__device__ __forceinline__ float dens_w1(float z) { return sinpif(0.5f * z); }  // sin(2 pi z / 4)
__device__ __forceinline__ float dens_w2(float z) {
  const float t = (z - 1.f) / 0.6f;
  return 3.f * expf(-0.5f * t * t);
}
__device__ __forceinline__ float dens_w3(float z) { return 3.f / (1.f + expf(-(z - 1.f) / 0.3f)); }

struct DensityArgs {
  float* x;         // [n,2]
  int64_t n;
  uint64_t offset;  // index of the first sample in the stream
  uint32_t k0, k1;
  int kind, num_blobs;
  float r;
};

// One thread per sample; word use: (r0,r1) normal pair, r2 uniform, r3 category.
__global__ void __launch_bounds__(256) density_sample_kernel(DensityArgs a) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
    const uint64_t idx = a.offset + (uint64_t)i;
    uint32_t w[4];
    philox4x32_10((uint32_t)idx, (uint32_t)(idx >> 32), (uint32_t)a.kind, 0u, a.k0, a.k1, w);
    float n0, n1;
    normal_pair(w[0], w[1], n0, n1);
    const float uni = u01_closed_open(w[2]);
    const float x1u = (uni - 0.5f) * (3.f * kPi);  // U(-3pi/2, 3pi/2)
    const bool second = (w[3] >> 31) != 0;          // Categorical([.5,.5])
    float x1, x2;
    switch (a.kind) {
      case FLOWCUDA_DENSITY_BANANA:  // banana.py:35-40
        x2 = 1.7320508075688772f * n0;
        x1 = x2 * x2 / 3.f - 2.f + 0.8660254037844386f * n1;
        break;
      case FLOWCUDA_DENSITY_ENERGY1: {  // energies.py:47-66
        x2 = 2.f * n0;
        const float mu = 2.f - x2 * x2 / 32.f;
        x1 = (second ? -mu : mu) + 0.3f * n1;
        break;
      }
      case FLOWCUDA_DENSITY_ENERGY2:  // energies.py:69-82
        x1 = x1u;
        x2 = dens_w1(x1) + 0.5f * n0;
        break;
      case FLOWCUDA_DENSITY_ENERGY3: {  // energies.py:85-104
        x1 = x1u;
        const float m1 = dens_w1(x1);
        x2 = (second ? m1 - dens_w2(x1) : m1) + 0.5f * n0;
        break;
      }
      case FLOWCUDA_DENSITY_ENERGY4: {  // energies.py:107-127
        x1 = x1u;
        const float m1 = dens_w1(x1);
        x2 = (second ? m1 + dens_w3(x1) : m1) + 0.5f * n0;
        break;
      }
      default: {  // gaussian blobs, gaussian_blobs.py:34-59: uniform component on a circle of radius r
        const int c = (int)(((uint64_t)w[3] * (uint64_t)a.num_blobs) >> 32);
        float sn, cs;
        sincospif(2.f * (float)c / (float)a.num_blobs, &sn, &cs);
        x1 = a.r * cs + 0.3f * n0;
        x2 = a.r * sn + 0.3f * n1;
        break;
      }
    }
    reinterpret_cast<float2*>(a.x)[i] = make_float2(x1, x2);
  }
}

// (byte + U[0,1)) / 256 (maf_mnist.py:101-106); 16 elements per thread: four Philox calls against one
// 16-byte load and four 16-byte stores.  Element e uses word e%4 of counter e/4.
__global__ void __launch_bounds__(256) dequantize_kernel(const uint8_t* __restrict__ in, float* __restrict__ out,
                                                         int64_t count, uint64_t offset, uint32_t k0, uint32_t k1) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t groups = count >> 4;
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += stride) {
    const uint4 b = reinterpret_cast<const uint4*>(in)[g];
    const uint32_t bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint64_t ctr = (offset >> 2) + (uint64_t)(g * 4 + q);
      uint32_t w[4];
      philox4x32_10((uint32_t)ctr, (uint32_t)(ctr >> 32), 0xDE0u, 0u, k0, k1, w);
      float4 o;
      o.x = ((float)(bw[q] & 255u) + u01_closed_open(w[0])) * 0.00390625f;
      o.y = ((float)((bw[q] >> 8) & 255u) + u01_closed_open(w[1])) * 0.00390625f;
      o.z = ((float)((bw[q] >> 16) & 255u) + u01_closed_open(w[2])) * 0.00390625f;
      o.w = ((float)(bw[q] >> 24) + u01_closed_open(w[3])) * 0.00390625f;
      reinterpret_cast<float4*>(out)[g * 4 + q] = o;
    }
  }
  // ragged tail (count % 16), one element per thread of block 0
  if (blockIdx.x == 0) {
    for (int64_t e = (groups << 4) + threadIdx.x; e < count; e += blockDim.x) {
      const uint64_t el = offset + (uint64_t)e;
      const uint64_t ctr = el >> 2;
      uint32_t w[4];
      philox4x32_10((uint32_t)ctr, (uint32_t)(ctr >> 32), 0xDE0u, 0u, k0, k1, w);
      out[e] = ((float)in[e] + u01_closed_open(w[el & 3])) * 0.00390625f;
    }
  }
}

// standard normal / uniform fills for sample(seed): element e = word e%4 of counter e/4, normals pair up
// words (0,1) and (2,3).
__global__ void __launch_bounds__(256) base_noise_kernel(float* __restrict__ out, int64_t count, uint64_t offset,
                                                         uint32_t k0, uint32_t k1, int uniform) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t groups = (count + 3) >> 2;
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += stride) {
    const uint64_t ctr = (offset >> 2) + (uint64_t)g;
    uint32_t w[4];
    philox4x32_10((uint32_t)ctr, (uint32_t)(ctr >> 32), 0xBA5Eu, 0u, k0, k1, w);
    float v[4];
    if (uniform) {
#pragma unroll
      for (int q = 0; q < 4; ++q) v[q] = u01_closed_open(w[q]);
    } else {
      normal_pair(w[0], w[1], v[0], v[1]);
      normal_pair(w[2], w[3], v[2], v[3]);
    }
    if (g * 4 + 3 < count) {
      reinterpret_cast<float4*>(out)[g] = make_float4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if (g * 4 + q < count) out[g * 4 + q] = v[q];
    }
  }
}

}  // namespace flowcuda

using namespace flowcuda;

extern "C" {

int flowcuda_density_sample(int kind, uint64_t seed, uint64_t offset, int num_blobs, float r, float* x, int64_t n,
                            void* stream) {
  FC_REQUIRE(n >= 0 && (x || n == 0), "density_sample: NULL output or negative n");
  FC_REQUIRE(kind >= FLOWCUDA_DENSITY_BANANA && kind <= FLOWCUDA_DENSITY_BLOBS, "density_sample: unknown density kind");
  FC_REQUIRE(kind != FLOWCUDA_DENSITY_BLOBS || num_blobs >= 1, "density_sample: gaussian blobs need num_blobs >= 1");
  FC_REQUIRE(((uintptr_t)x & 7) == 0, "density_sample: output must be 8-byte aligned");
  if (n == 0) return FLOWCUDA_OK;
  DensityArgs a;
  a.x = x; a.n = n; a.offset = offset;
  a.k0 = (uint32_t)seed; a.k1 = (uint32_t)(seed >> 32);
  a.kind = kind; a.num_blobs = num_blobs; a.r = r;
  density_sample_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(a);
  FC_LAUNCH_CHECK("density_sample_kernel");
  return FLOWCUDA_OK;
}

int flowcuda_dequantize_u8(const uint8_t* bytes, uint64_t seed, uint64_t offset, float* out, int64_t count,
                           void* stream) {
  FC_REQUIRE(count >= 0 && ((bytes && out) || count == 0), "dequantize_u8: NULL pointer or negative count");
  FC_REQUIRE(((uintptr_t)bytes & 15) == 0 && ((uintptr_t)out & 15) == 0, "dequantize_u8: buffers must be 16-byte aligned");
  FC_REQUIRE((offset & 15) == 0, "dequantize_u8: offset must be a multiple of 16 elements");
  if (count == 0) return FLOWCUDA_OK;
  const int64_t groups = count >> 4;
  dequantize_kernel<<<grid_for(groups > 0 ? groups : 1, 256), 256, 0, (cudaStream_t)stream>>>(
      bytes, out, count, offset, (uint32_t)seed, (uint32_t)(seed >> 32));
  FC_LAUNCH_CHECK("dequantize_kernel");
  return FLOWCUDA_OK;
}

int flowcuda_base_noise(int uniform, uint64_t seed, uint64_t offset, float* out, int64_t count, void* stream) {
  FC_REQUIRE(count >= 0 && (out || count == 0), "base_noise: NULL output or negative count");
  FC_REQUIRE(((uintptr_t)out & 15) == 0 && (offset & 3) == 0, "base_noise: output must be 16-byte aligned, offset a multiple of 4");
  if (count == 0) return FLOWCUDA_OK;
  base_noise_kernel<<<grid_for((count + 3) >> 2, 256), 256, 0, (cudaStream_t)stream>>>(
      out, count, offset, (uint32_t)seed, (uint32_t)(seed >> 32), uniform);
  FC_LAUNCH_CHECK("base_noise_kernel");
  return FLOWCUDA_OK;
}

}  // extern "C"
