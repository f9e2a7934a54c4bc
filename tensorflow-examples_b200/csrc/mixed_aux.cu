// CUDA-core kernels around the tensor-core GEMMs of the bf16 / fp32-accumulate path
// (BASELINE.json configs[4], 784-4096-4096-10): everything whose shape is not a dense contraction.
//   cast_bf16_kernel      fp32 -> bf16 (inputs x each step; weights inside the SGD apply)
//   pad_cast64_kernel     fp32 [rows,C<=16] -> bf16 [rows,64]: the 10-wide output layer as a tensor-core operand
//   colsum_f32_narrow     bias gradient of the output layer (fp32 dz), per-slice partials
//   colsum_bf16_kernel    bias gradient of a hidden layer: column sums of a bf16 [rows,N] matrix (two stages)
//   sgd_apply_mixed_kernel  var -= lr*grad on the fp32 masters + refresh of the bf16 weight copies
// All reductions use fixed slices and a fixed order (no float atomics).
#include <cuda_bf16.h>

#include "common.cuh"

namespace mx {

__device__ __forceinline__ float bf16r(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
__device__ __forceinline__ float bf_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf_hi(uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); }

__global__ void __launch_bounds__(256)
cast_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long n) {
  const long n4 = n >> 2;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
    const float4 v = reinterpret_cast<const float4*>(src)[i];
    const __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
    uint2 o;
    o.x = *reinterpret_cast<const uint32_t*>(&a);
    o.y = *reinterpret_cast<const uint32_t*>(&b);
    reinterpret_cast<uint2*>(dst)[i] = o;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) dst[(n4 << 2) + threadIdx.x] = __float2bfloat16_rn(src[(n4 << 2) + threadIdx.x]);
}

// ---------------------------------------------------------------------------------------------
// The 10-wide output layer (distributed.py:98) is run on the tensor cores too, padded to 64 columns:
// pad_cast64 makes bf16 [rows,64] copies of W_out [K,10] and of dz_out [rows,10]; its bias gradient
// (column sums of the fp32 dz) is reduced here in fixed slices.
// ---------------------------------------------------------------------------------------------
constexpr int NARROW = 64;

__global__ void __launch_bounds__(256)
pad_cast64_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long rows, int C) {
  // one thread per (row, 8-column group)
  const long total = rows * (NARROW / 8);
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const long r = i >> 3;
    const int g = (int)(i & 7) * 8;
    uint32_t w[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int c0 = g + 2 * u, c1 = c0 + 1;
      const float v0 = c0 < C ? src[r * C + c0] : 0.f, v1 = c1 < C ? src[r * C + c1] : 0.f;
      const __nv_bfloat162 h = __floats2bfloat162_rn(v0, v1);
      w[u] = *reinterpret_cast<const uint32_t*>(&h);
    }
    reinterpret_cast<uint4*>(dst)[i] = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

__global__ void __launch_bounds__(256)
colsum_f32_narrow_kernel(const float* __restrict__ dz, float* __restrict__ partB, long part_stride, int rows, int C,
                         int rows_per_slice) {
  __shared__ float sh[16][17];
  const int c = threadIdx.x & 15, sub = threadIdx.x >> 4;            // 16 columns x 16 row sub-slices
  const int r0 = blockIdx.x * rows_per_slice, r1 = min(rows, r0 + rows_per_slice);
  float s = 0.f;
  if (c < C)
    for (int r = r0 + sub; r < r1; r += 16) s += dz[(size_t)r * C + c];
  sh[sub][c] = s;
  __syncthreads();
  if (sub == 0 && c < C) {
    float t = 0.f;
#pragma unroll
    for (int u = 0; u < 16; ++u) t += sh[u][c];
    partB[(size_t)blockIdx.x * part_stride + c] = t;
  }
}

// ---------------------------------------------------------------------------------------------
// Column sums of a bf16 [rows, N] matrix.  Stage 1: grid = (N/512, slices): thread owns 2 columns, adds
// the rows of its slice in order.  Stage 2: adds the slices in order.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
colsum_bf16_kernel(const __nv_bfloat16* __restrict__ m, float* __restrict__ part, int rows, int N, int rows_per_slice) {
  const int c2 = blockIdx.x * 256 + threadIdx.x;      // column pair
  if (2 * c2 >= N) return;
  const int r0 = blockIdx.y * rows_per_slice, r1 = min(rows, r0 + rows_per_slice);
  float s0 = 0.f, s1 = 0.f;
  const uint32_t* p = reinterpret_cast<const uint32_t*>(m) + c2;
  const size_t ld = (size_t)N / 2;
  int r = r0;
  for (; r + 4 <= r1; r += 4) {
    const uint32_t v0 = p[(size_t)r * ld], v1 = p[(size_t)(r + 1) * ld], v2 = p[(size_t)(r + 2) * ld], v3 = p[(size_t)(r + 3) * ld];
    s0 += bf_lo(v0); s1 += bf_hi(v0);
    s0 += bf_lo(v1); s1 += bf_hi(v1);
    s0 += bf_lo(v2); s1 += bf_hi(v2);
    s0 += bf_lo(v3); s1 += bf_hi(v3);
  }
  for (; r < r1; ++r) { const uint32_t v = p[(size_t)r * ld]; s0 += bf_lo(v); s1 += bf_hi(v); }
  part[(size_t)blockIdx.y * N + 2 * c2] = s0;
  part[(size_t)blockIdx.y * N + 2 * c2 + 1] = s1;
}

// Stage 2: out[c] = sum over slices, in a fixed order.  Block = 32 columns x 8 interleaved slice groups.
__global__ void __launch_bounds__(256)
colsum_finish_kernel(const float* __restrict__ part, float* __restrict__ out, int N, int slices) {
  __shared__ float sh[8][33];
  const int cl = threadIdx.x & 31, zg = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cl;
  float s = 0.f;
  if (c < N) {
    int z = zg;
    for (; z + 24 < slices; z += 32) {
      const float v0 = part[(size_t)z * N + c], v1 = part[(size_t)(z + 8) * N + c], v2 = part[(size_t)(z + 16) * N + c],
                  v3 = part[(size_t)(z + 24) * N + c];
      s += v0; s += v1; s += v2; s += v3;
    }
    for (; z < slices; z += 8) s += part[(size_t)z * N + c];
  }
  sh[zg][cl] = s;
  __syncthreads();
  if (zg == 0 && c < N) {
    float t = sh[0][cl];
#pragma unroll
    for (int u = 1; u < 8; ++u) t += sh[u][cl];
    out[c] = t;
  }
}

// ---------------------------------------------------------------------------------------------
// ApplyGradientDescent on the fp32 masters (distributed.py:107-108) + the bf16 copies the next step's
// GEMMs read + the step's loss / correct trace + global_step.
// ---------------------------------------------------------------------------------------------
struct ApplyArgs {
  int n_seg;
  long off[B2_MAX_LAYERS], size[B2_MAX_LAYERS];   // W segments that have a bf16 copy
  __nv_bfloat16* wb[B2_MAX_LAYERS];
};

__global__ void __launch_bounds__(256)
sgd_apply_mixed_kernel(float* __restrict__ params, const float* __restrict__ grads, long P, float lr, ApplyArgs a,
                       float* __restrict__ loss_trace, int* __restrict__ correct_trace, float* __restrict__ global_step) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < P; i += (long)gridDim.x * blockDim.x) {
    const float v = params[i] - lr * grads[i];
    params[i] = v;
#pragma unroll
    for (int s = 0; s < B2_MAX_LAYERS; ++s)
      if (s < a.n_seg && i >= a.off[s] && i < a.off[s] + a.size[s]) a.wb[s][i - a.off[s]] = __float2bfloat16_rn(v);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    if (loss_trace) *loss_trace = grads[P];
    if (correct_trace) *correct_trace = (int)(grads[P + 1] + 0.5f);
    *global_step += 1.0f;
  }
}

}  // namespace mx

#define MX_CHECK()                                                                          \
  do {                                                                                      \
    cudaError_t _e = cudaGetLastError();                                                    \
    if (_e != cudaSuccess) {                                                                \
      b2_set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
      return B2_ERR_CUDA;                                                                   \
    }                                                                                       \
  } while (0)

int b2_launch_cast_bf16(const float* src, void* dst, long n, int sm_count, cudaStream_t s) {
  long blocks = (n / 4 + 255) / 256;
  if (blocks > 8L * sm_count) blocks = 8L * sm_count;
  if (blocks < 1) blocks = 1;
  mx::cast_bf16_kernel<<<(int)blocks, 256, 0, s>>>(src, reinterpret_cast<__nv_bfloat16*>(dst), n);
  MX_CHECK();
  return B2_OK;
}

int b2_launch_pad_cast64(const float* src, void* dst, long rows, int C, int sm_count, cudaStream_t s) {
  if (C > mx::NARROW) { b2_set_error("pad_cast64: C must be <= 64"); return B2_ERR_ARG; }
  long blocks = (rows * 8 + 255) / 256;
  if (blocks > 8L * sm_count) blocks = 8L * sm_count;
  if (blocks < 1) blocks = 1;
  mx::pad_cast64_kernel<<<(int)blocks, 256, 0, s>>>(src, reinterpret_cast<__nv_bfloat16*>(dst), rows, C);
  MX_CHECK();
  return B2_OK;
}

int b2_launch_colsum_f32_narrow(const float* dz, float* partB, long part_stride, int rows, int C, int slices,
                                cudaStream_t s) {
  if (C > 16) { b2_set_error("colsum_f32_narrow: C must be <= 16"); return B2_ERR_ARG; }
  const int rps = (rows + slices - 1) / slices;
  mx::colsum_f32_narrow_kernel<<<slices, 256, 0, s>>>(dz, partB, part_stride, rows, C, rps);
  MX_CHECK();
  return B2_OK;
}

int b2_colsum_slices(int rows, int N, int sm_count) {
  const int bx = (N / 2 + 255) / 256;
  int slices = (4 * sm_count + bx - 1) / bx;
  const int cap = (rows + 31) / 32;
  if (slices > cap) slices = cap;
  return slices < 1 ? 1 : slices;
}

int b2_launch_colsum_bf16(const void* m, float* part, float* out, int rows, int N, int slices, cudaStream_t s) {
  if (N % 2) { b2_set_error("colsum_bf16: N must be even"); return B2_ERR_ARG; }
  const int bx = (N / 2 + 255) / 256;
  const int rps = (rows + slices - 1) / slices;
  mx::colsum_bf16_kernel<<<dim3(bx, slices), 256, 0, s>>>(reinterpret_cast<const __nv_bfloat16*>(m), part, rows, N, rps);
  MX_CHECK();
  mx::colsum_finish_kernel<<<(N + 31) / 32, 256, 0, s>>>(part, out, N, slices);
  MX_CHECK();
  return B2_OK;
}

int b2_launch_colsum_finish(const float* part, float* out, int N, int slices, cudaStream_t s) {
  mx::colsum_finish_kernel<<<(N + 31) / 32, 256, 0, s>>>(part, out, N, slices);
  MX_CHECK();
  return B2_OK;
}

int b2_launch_sgd_apply_mixed(float* params, const float* grads, long P, float lr, int n_seg, const long* off,
                              const long* size, void* const* wb, float* loss_trace, int* correct_trace,
                              float* global_step, int sm_count, cudaStream_t s) {
  mx::ApplyArgs a;
  a.n_seg = n_seg;
  for (int i = 0; i < B2_MAX_LAYERS; ++i) {
    a.off[i] = i < n_seg ? off[i] : 0;
    a.size[i] = i < n_seg ? size[i] : 0;
    a.wb[i] = i < n_seg ? reinterpret_cast<__nv_bfloat16*>(wb[i]) : nullptr;
  }
  long blocks = (P + 255) / 256;
  if (blocks > 8L * sm_count) blocks = 8L * sm_count;
  mx::sgd_apply_mixed_kernel<<<(int)blocks, 256, 0, s>>>(params, grads, P, lr, a, loss_trace, correct_trace, global_step);
  MX_CHECK();
  return B2_OK;
}
