// CUDA-core kernels around the tensor-core GEMMs of the bf16 / fp32-accumulate path
// (BASELINE.json configs[4], 784-4096-4096-10): everything whose shape is not a dense contraction.
//   cast_bf16_kernel      fp32 -> bf16 (inputs x each step; weights inside the SGD apply)
//   skinny_fwd_kernel     logits = a[rows,K](bf16) . bf16(W[K,C]) + b     C <= 16 (distributed.py:98)
//   skinny_bwd_kernel     the three gradients of that narrow layer in one pass over a:
//                           dz_prev = (bf16(dz) . bf16(W)^T) * a(1-a) -> bf16      (MatMul grad b + SigmoidGrad)
//                           dW partials = a^T . bf16(dz), db partials = colsum(dz)  (MatMul grad a, BiasAdd grad)
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
// logits[r][c] = sum_k a[r][k] * bf16(W[k][c]) + b[c].   One warp per row pair; lanes split K in
// chunks of 8 (one 16-byte load of a per chunk); W^T (bf16-rounded, as floats) lives in shared memory.
// K % 256 == 0 is not required; K % 8 == 0 is.
// ---------------------------------------------------------------------------------------------
constexpr int SK_MAXC = 16;

__global__ void __launch_bounds__(512)
skinny_fwd_kernel(const __nv_bfloat16* __restrict__ a, const float* __restrict__ W, const float* __restrict__ b,
                  float* __restrict__ logits, int rows, int K, int C) {
  extern __shared__ float s_wt[];   // [C][K] : W transposed, bf16-rounded
  for (int i = threadIdx.x; i < K * C; i += blockDim.x) {
    const int k = i / C, c = i - k * C;
    s_wt[c * K + k] = bf16r(W[i]);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  for (long r = (long)blockIdx.x * nwarp + warp; r < rows; r += (long)gridDim.x * nwarp) {
    float acc[SK_MAXC];
#pragma unroll
    for (int c = 0; c < SK_MAXC; ++c) acc[c] = 0.f;
    const uint4* ar = reinterpret_cast<const uint4*>(a + r * K);
    for (int k8 = lane; k8 < K / 8; k8 += 32) {
      const uint4 q = ar[k8];
      const float x[8] = {bf_lo(q.x), bf_hi(q.x), bf_lo(q.y), bf_hi(q.y), bf_lo(q.z), bf_hi(q.z), bf_lo(q.w), bf_hi(q.w)};
#pragma unroll
      for (int c = 0; c < SK_MAXC; ++c) {
        if (c < C) {
          const float4 w0 = *reinterpret_cast<const float4*>(s_wt + c * K + k8 * 8);
          const float4 w1 = *reinterpret_cast<const float4*>(s_wt + c * K + k8 * 8 + 4);
          acc[c] = fmaf(x[0], w0.x, acc[c]); acc[c] = fmaf(x[1], w0.y, acc[c]);
          acc[c] = fmaf(x[2], w0.z, acc[c]); acc[c] = fmaf(x[3], w0.w, acc[c]);
          acc[c] = fmaf(x[4], w1.x, acc[c]); acc[c] = fmaf(x[5], w1.y, acc[c]);
          acc[c] = fmaf(x[6], w1.z, acc[c]); acc[c] = fmaf(x[7], w1.w, acc[c]);
        }
      }
    }
#pragma unroll
    for (int c = 0; c < SK_MAXC; ++c) {
      if (c < C) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], o);
      }
    }
    if (lane == 0) {
#pragma unroll
      for (int c = 0; c < SK_MAXC; ++c)
        if (c < C) logits[r * C + c] = acc[c] + b[c];
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Backward of the narrow layer.  grid = (row slices, K / (threads*4)); thread owns 4 inputs k.
// Rows of the slice stream through: per row  da[k] = sum_c dz[c]*W[k][c];  dz_prev = da*a*(1-a) -> bf16;
// gW[k][c] += a[k]*dz[c].  Partials are written in the layout the layered reduce kernel expects:
// partW[slice][k*C + c], partB[slice][c].
// ---------------------------------------------------------------------------------------------
constexpr int SB_THREADS = 256, SB_KPT = 4, SB_ROWCHUNK = 64;

__global__ void __launch_bounds__(SB_THREADS)
skinny_bwd_kernel(const __nv_bfloat16* __restrict__ a, const float* __restrict__ dz, const float* __restrict__ W,
                  __nv_bfloat16* __restrict__ dz_prev, float* __restrict__ partW, float* __restrict__ partB,
                  long part_stride, int rows, int K, int C, int rows_per_slice) {
  __shared__ float s_dz[SB_ROWCHUNK * SK_MAXC];
  const int slice = blockIdx.x;
  const int k0 = (blockIdx.y * SB_THREADS + threadIdx.x) * SB_KPT;
  const bool kact = k0 < K;
  const int r_begin = slice * rows_per_slice, r_end = min(rows, r_begin + rows_per_slice);
  float w[SB_KPT][SK_MAXC], g[SB_KPT][SK_MAXC];
#pragma unroll
  for (int j = 0; j < SB_KPT; ++j)
#pragma unroll
    for (int c = 0; c < SK_MAXC; ++c) {
      w[j][c] = (kact && c < C) ? bf16r(W[(size_t)(k0 + j) * C + c]) : 0.f;
      g[j][c] = 0.f;
    }
  float bsum = 0.f;   // thread c < C of blockIdx.y == 0 accumulates db[c]
  for (int rc = r_begin; rc < r_end; rc += SB_ROWCHUNK) {
    const int nr = min(SB_ROWCHUNK, r_end - rc);
    __syncthreads();
    for (int i = threadIdx.x; i < nr * C; i += SB_THREADS) {
      const int rr = i / C, c = i - rr * C;
      s_dz[rr * SK_MAXC + c] = dz[(size_t)(rc + rr) * C + c];
    }
    __syncthreads();
    if (blockIdx.y == 0 && threadIdx.x < C)
      for (int rr = 0; rr < nr; ++rr) bsum += s_dz[rr * SK_MAXC + threadIdx.x];   // fp32 dz, rows in order
    if (!kact) continue;
    for (int rr = 0; rr < nr; ++rr) {
      const size_t off = (size_t)(rc + rr) * K + k0;
      const uint2 q = *reinterpret_cast<const uint2*>(a + off);
      const float av[4] = {bf_lo(q.x), bf_hi(q.x), bf_lo(q.y), bf_hi(q.y)};
      float da[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int c = 0; c < SK_MAXC; ++c) {
        if (c < C) {
          const float d = bf16r(s_dz[rr * SK_MAXC + c]);       // GEMM operand precision
#pragma unroll
          for (int j = 0; j < SB_KPT; ++j) {
            da[j] = fmaf(d, w[j][c], da[j]);
            g[j][c] = fmaf(av[j], d, g[j][c]);
          }
        }
      }
      const __nv_bfloat162 o0 = __floats2bfloat162_rn(da[0] * av[0] * (1.f - av[0]), da[1] * av[1] * (1.f - av[1]));
      const __nv_bfloat162 o1 = __floats2bfloat162_rn(da[2] * av[2] * (1.f - av[2]), da[3] * av[3] * (1.f - av[3]));
      uint2 o;
      o.x = *reinterpret_cast<const uint32_t*>(&o0);
      o.y = *reinterpret_cast<const uint32_t*>(&o1);
      *reinterpret_cast<uint2*>(dz_prev + off) = o;
    }
  }
  if (kact) {
    float* pw = partW + (size_t)slice * part_stride;
#pragma unroll
    for (int j = 0; j < SB_KPT; ++j)
#pragma unroll
      for (int c = 0; c < SK_MAXC; ++c)
        if (c < C) pw[(size_t)(k0 + j) * C + c] = g[j][c];
  }
  if (blockIdx.y == 0 && threadIdx.x < C) partB[(size_t)slice * part_stride + threadIdx.x] = bsum;
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

__global__ void __launch_bounds__(256)
colsum_finish_kernel(const float* __restrict__ part, float* __restrict__ out, int N, int slices) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= N) return;
  float s = 0.f;
  for (int z = 0; z < slices; ++z) s += part[(size_t)z * N + c];
  out[c] = s;
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

int b2_launch_skinny_fwd(const void* a, const float* W, const float* b, float* logits, int rows, int K, int C,
                         int sm_count, cudaStream_t s) {
  if (C > mx::SK_MAXC || K % 8) { b2_set_error("skinny_fwd: needs C <= 16 and K %% 8 == 0 (C=%d K=%d)", C, K); return B2_ERR_ARG; }
  const int smem = K * C * (int)sizeof(float);
  if (smem > 200 * 1024) { b2_set_error("skinny_fwd: K*C too large for shared memory (K=%d C=%d)", K, C); return B2_ERR_ARG; }
  static int attr_smem = 0;
  if (smem > 48 * 1024 && smem > attr_smem) {
    B2_CUDA(cudaFuncSetAttribute(mx::skinny_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_smem = smem;
  }
  int blocks = (rows + 15) / 16;
  if (blocks > sm_count) blocks = sm_count;
  mx::skinny_fwd_kernel<<<blocks, 512, smem, s>>>(reinterpret_cast<const __nv_bfloat16*>(a), W, b, logits, rows, K, C);
  MX_CHECK();
  return B2_OK;
}

int b2_skinny_bwd_slices(int rows, int K, int sm_count) {
  const int ky = (K + mx::SB_THREADS * mx::SB_KPT - 1) / (mx::SB_THREADS * mx::SB_KPT);
  int slices = (2 * sm_count + ky - 1) / ky;
  const int cap = (rows + 63) / 64;
  if (slices > cap) slices = cap;
  return slices < 1 ? 1 : slices;
}

int b2_launch_skinny_bwd(const void* a, const float* dz, const float* W, void* dz_prev, float* partW, float* partB,
                         long part_stride, int rows, int K, int C, int slices, cudaStream_t s) {
  if (C > mx::SK_MAXC || K % 4) { b2_set_error("skinny_bwd: needs C <= 16 and K %% 4 == 0"); return B2_ERR_ARG; }
  const int ky = (K + mx::SB_THREADS * mx::SB_KPT - 1) / (mx::SB_THREADS * mx::SB_KPT);
  const int rps = (rows + slices - 1) / slices;
  mx::skinny_bwd_kernel<<<dim3(slices, ky), mx::SB_THREADS, 0, s>>>(
      reinterpret_cast<const __nv_bfloat16*>(a), dz, W, reinterpret_cast<__nv_bfloat16*>(dz_prev), partW, partB,
      part_stride, rows, K, C, rps);
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
  mx::colsum_finish_kernel<<<(N + 255) / 256, 256, 0, s>>>(part, out, N, slices);
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
