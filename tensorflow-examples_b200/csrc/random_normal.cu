// tf.random_normal on the device: the initial weights of distributed.py:85-86 drawn the way TensorFlow 1.x's CPU kernel
// draws them (Philox4x32-10 keyed by the op's (seed, seed2) attributes, four normals per Philox block by Box-Muller),
// so that a run can start from the reference's own initial weights instead of injected ones when the pair is known.
// Restated from the published algorithm [TF1: lib/random/philox_random.h, random_distributions.h]; the CPU oracle is
// oracle/tf1_random.py (pinned to the Random123 known-answer vectors; TensorFlow itself is not available to compare).
#include <math.h>

#include "common.cuh"

namespace rn {

__device__ __forceinline__ void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
    const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
    c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
}

__device__ __forceinline__ float uint32_to_float(uint32_t x) {
  return __uint_as_float(0x3F800000u | (x & 0x7FFFFFu)) - 1.0f;          // [0, 1)
}

__global__ void __launch_bounds__(256)
random_normal_kernel(float* __restrict__ out, long n, uint32_t key0, uint32_t key1, uint32_t ctr2, uint32_t ctr3,
                     float mean, float stddev) {
  const long groups = (n + 3) / 4;
  for (long g = (long)blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += (long)gridDim.x * blockDim.x) {
    uint32_t c[4] = {(uint32_t)((unsigned long long)g & 0xFFFFFFFFull), (uint32_t)((unsigned long long)g >> 32), ctr2, ctr3};
    philox4x32_10(c, key0, key1);
    float f[4];
#pragma unroll
    for (int i = 0; i < 4; i += 2) {
      const float u1 = fmaxf(uint32_to_float(c[i]), 1.0e-7f);
      const float v1 = 6.283185307179586f * uint32_to_float(c[i + 1]);
      const float rad = sqrtf(-2.0f * logf(u1));
      float sn, cs;
      sincosf(v1, &sn, &cs);
      f[i] = sn * rad * stddev + mean;
      f[i + 1] = cs * rad * stddev + mean;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (4 * g + i < n) out[4 * g + i] = f[i];
  }
}

}  // namespace rn

extern "C" int b2_random_normal(float* out_dev, long n, unsigned long long seed, unsigned long long seed2, float mean,
                                float stddev, b2_stream_t stream) {
  B2_CHECK_ARG(out_dev != nullptr && n >= 0, "b2_random_normal: bad output (%p, %ld elements)", (void*)out_dev, n);
  if (n == 0) return B2_OK;
  long blocks = ((n + 3) / 4 + 255) / 256;
  if (blocks > 4096) blocks = 4096;
  rn::random_normal_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(
      out_dev, n, (uint32_t)(seed & 0xFFFFFFFFull), (uint32_t)(seed >> 32), (uint32_t)(seed2 & 0xFFFFFFFFull),
      (uint32_t)(seed2 >> 32), mean, stddev);
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}
