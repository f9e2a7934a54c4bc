// CUDA-core kernels around the tensor-core GEMMs of the bf16 / fp32-accumulate path
// (BASELINE.json configs[4], 784-4096-4096-10): everything whose shape is not a dense contraction.
//   cast_bf16_kernel      fp32 -> bf16 (inputs x each step; weights inside the SGD apply)
//   pad_cast64_kernel     fp32 [rows,C<=16] -> bf16 [rows,64]: the 10-wide output layer as a tensor-core operand
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
// pad_cast64 makes the bf16 [K,64] copy of W_out [K,10] (dz_out arrives padded from loss_head_rows_kernel).
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
    for (; z + 8 * 15 < slices; z += 8 * 16) {          // sixteen independent loads in flight, added in slice order
      float v[16];
#pragma unroll
      for (int u = 0; u < 16; ++u) v[u] = part[(size_t)(z + 8 * u) * N + c];
#pragma unroll
      for (int u = 0; u < 16; ++u) s += v[u];
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
  // four parameters per thread: every segment of the flat layout starts on a 16-byte boundary and P is a multiple of 4,
  // so a float4 never straddles two segments
  const long P4 = P >> 2;
  for (long q = (long)blockIdx.x * blockDim.x + threadIdx.x; q < P4; q += (long)gridDim.x * blockDim.x) {
    const long i = q << 2;
    float4 p = reinterpret_cast<const float4*>(params)[q];
    const float4 g = reinterpret_cast<const float4*>(grads)[q];
    p.x -= lr * g.x; p.y -= lr * g.y; p.z -= lr * g.z; p.w -= lr * g.w;
    reinterpret_cast<float4*>(params)[q] = p;
#pragma unroll
    for (int s = 0; s < B2_MAX_LAYERS; ++s) {
      if (s < a.n_seg && i >= a.off[s] && i < a.off[s] + a.size[s]) {
        const long j = i - a.off[s];
        __nv_bfloat16* w = a.wb[s] + j;
        if (j + 4 <= a.size[s] && (j & 3) == 0) {
          const __nv_bfloat162 lo = __floats2bfloat162_rn(p.x, p.y), hi = __floats2bfloat162_rn(p.z, p.w);
          *reinterpret_cast<uint2*>(w) = make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi));
        } else {
          const float v[4] = {p.x, p.y, p.z, p.w};
#pragma unroll
          for (int u = 0; u < 4; ++u)
            if (j + u < a.size[s]) w[u] = __float2bfloat16_rn(v[u]);
        }
      }
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    if (loss_trace) *loss_trace = grads[P];
    if (correct_trace) *correct_trace = (int)(grads[P + 1] + 0.5f);
    *global_step += 1.0f;
  }
}


// ---------------------------------------------------------------------------------------------
// Loss head of the bf16 path, one THREAD per row (<= 16 classes): softmax / loss / argmax / dz exactly as
// loss_head_kernel composes them (distributed.py:99-102,117-118), and in the same pass
//   * dz written as the zero-padded bf16 [rows, 64] operand the tensor-core products of the output layer read,
//   * the column sums of dz (BiasAdd gradient of the output layer) as one partial row per block.
// Replaces three launches (warp-per-row head, fp32 -> padded bf16 cast, 5-block column sum: 59 + 5 + 58 us at 65536 rows).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
loss_head_rows_kernel(const float* __restrict__ z, const float* __restrict__ y_, __nv_bfloat16* __restrict__ dzn,
                      float* __restrict__ loss_part, float* __restrict__ correct_part, float* __restrict__ db_part,
                      int rows, int C, int loss_kind, float inv_B) {
  __shared__ float sh[8][18];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float db[16];
#pragma unroll
  for (int c = 0; c < 16; ++c) db[c] = 0.f;
  float loss_acc = 0.f, corr_acc = 0.f;
  for (long row = (long)blockIdx.x * blockDim.x + threadIdx.x; row < rows; row += (long)gridDim.x * blockDim.x) {
    float zv[16], t[16], d[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      zv[c] = c < C ? z[row * C + c] : -INFINITY;
      t[c] = c < C ? y_[row * C + c] : 0.f;
      d[c] = 0.f;
    }
    if (loss_kind == B2_LOSS_SSE_SUM) {
      float l = 0.f;
#pragma unroll
      for (int c = 0; c < 16; ++c)
        if (c < C) { const float e = zv[c] - t[c]; l = fmaf(e, e, l); d[c] = 2.0f * e; }
      loss_acc += l;
    } else {
      float m = zv[0];
#pragma unroll
      for (int c = 1; c < 16; ++c) m = fmaxf(m, zv[c]);
      float e[16], ssum = 0.f;
#pragma unroll
      for (int c = 0; c < 16; ++c) { e[c] = c < C ? expf(zv[c] - m) : 0.f; ssum += e[c]; }
      const float inv = 1.0f / ssum;                         // softmax = exp(z - max) * (1 / sum)   [TF1]
      float l = 0.f, sy = 0.f, bv = -INFINITY, lv = -INFINITY;
      int bi = 0, li = 0;
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        if (c < C) {
          const float yv = e[c] * inv;
          l -= t[c] * logf(yv);                              // un-fused log(softmax), distributed.py:102
          sy += t[c];
          if (yv > bv) { bv = yv; bi = c; }                  // first index wins ties (argmax)
          if (t[c] > lv) { lv = t[c]; li = c; }
          e[c] = yv;
        }
      }
#pragma unroll
      for (int c = 0; c < 16; ++c)
        if (c < C) d[c] = (e[c] * sy - t[c]) * inv_B;
      loss_acc += l;
      corr_acc += (bi == li) ? 1.0f : 0.0f;
    }
    uint32_t w[8];
#pragma unroll
    for (int c = 0; c < 16; c += 2) {
      const __nv_bfloat162 h = __floats2bfloat162_rn(d[c], d[c + 1]);
      w[c >> 1] = *reinterpret_cast<const uint32_t*>(&h);
      db[c] += d[c];
      db[c + 1] += d[c + 1];
    }
    uint4* o = reinterpret_cast<uint4*>(dzn + row * NARROW);   // 64 bf16 = 128 B per row
    o[0] = make_uint4(w[0], w[1], w[2], w[3]);
    o[1] = make_uint4(w[4], w[5], w[6], w[7]);
#pragma unroll
    for (int q = 2; q < 8; ++q) o[q] = make_uint4(0u, 0u, 0u, 0u);
  }
  // block sums in a fixed order: shuffle tree over the lanes, then the eight warps in order
#pragma unroll
  for (int c = 0; c < 16; ++c) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) db[c] += __shfl_xor_sync(0xffffffffu, db[c], o);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    loss_acc += __shfl_xor_sync(0xffffffffu, loss_acc, o);
    corr_acc += __shfl_xor_sync(0xffffffffu, corr_acc, o);
  }
  if (lane == 0) {
#pragma unroll
    for (int c = 0; c < 16; ++c) sh[warp][c] = db[c];
    sh[warp][16] = loss_acc;
    sh[warp][17] = corr_acc;
  }
  __syncthreads();
  if (threadIdx.x < 18) {
    float v = 0.f;
#pragma unroll
    for (int wq = 0; wq < 8; ++wq) v += sh[wq][threadIdx.x];
    if (threadIdx.x < 16) db_part[(size_t)blockIdx.x * 16 + threadIdx.x] = v;
    else if (threadIdx.x == 16) loss_part[blockIdx.x] = (loss_kind == B2_LOSS_SSE_SUM) ? v : v * inv_B;
    else correct_part[blockIdx.x] = v;
  }
}


// ---------------------------------------------------------------------------------------------
// Input gradient of the 10-wide output layer on the bf16 path (TF1 MatMul gradient, transpose_b, then SigmoidGrad):
//   dz_prev[r][j] = bf16( (sum_c dz[r][c] * W[j][c]) * a[r][j] * (1 - a[r][j]) )        r < rows, j < K (4096)
// with K = 10 in the reduction this is an element-wise pass over two 512 MB matrices, not a GEMM: run through the
// tensor-core kernel (K padded to 64) it was epilogue-bound (398 us at 65536 x 4096); here every thread owns 8 columns
// (their 8 x C weights live in registers), streams rows, and also produces the column sums of the ROUNDED outputs (the
// bias gradient of the layer below) as one partial row per 64-row block.
// grid = (ceil(K / (8 CG)), ceil(rows / 64)), CG column groups x RL row lanes per CTA.  Two shapes: 64 x 4 (a 512-column
// strip: narrow layers) and 512 x 1 (rows of up to 4096 columns, walked one after the other: every CTA streams whole
// 8 KB rows instead of 1 KB pieces 8 KB apart, which is what HBM likes).
// ---------------------------------------------------------------------------------------------
constexpr int ND_ROWS = 64;

template <int CT, int CG, int RL>   // classes rounded up to an even number: the weights of a thread's 8 columns live in 8 x CT registers
__global__ void __launch_bounds__(CG * RL, CG * RL >= 512 ? 1 : 2)
narrow_dgrad_bf16_kernel(const __nv_bfloat16* __restrict__ dzn, const __nv_bfloat16* __restrict__ wpad,
                         const __nv_bfloat16* __restrict__ act, __nv_bfloat16* __restrict__ out,
                         float* __restrict__ colsum_part, int rows, int K) {
  constexpr int ND_COLS = 8 * CG;
  __shared__ float sh[RL > 1 ? RL : 1][RL > 1 ? ND_COLS + 8 : 1];
  const int cg = threadIdx.x % CG, rl = threadIdx.x / CG;
  const int j0 = blockIdx.x * ND_COLS + cg * 8;
  const bool col_ok = j0 < K;                                   // K is a multiple of 8
  float w[8][CT];
#pragma unroll
  for (int u = 0; u < 8; ++u) {
    uint4 q0 = make_uint4(0u, 0u, 0u, 0u), q1 = q0;
    if (col_ok) {
      const uint4* wp = reinterpret_cast<const uint4*>(wpad + (size_t)(j0 + u) * NARROW);
      q0 = wp[0]; q1 = wp[1];
    }
    const uint32_t ww[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
    for (int c = 0; c < CT / 2; ++c) { w[u][2 * c] = bf_lo(ww[c]); w[u][2 * c + 1] = bf_hi(ww[c]); }
  }
  float cs[8];
#pragma unroll
  for (int u = 0; u < 8; ++u) cs[u] = 0.f;
  const int r_begin = blockIdx.y * ND_ROWS, r_end = min(rows, r_begin + ND_ROWS);
#pragma unroll 2
  for (int r = r_begin + rl; r < r_end; r += RL) {
    const uint4* dp = reinterpret_cast<const uint4*>(dzn + (size_t)r * NARROW);
    const uint4 d0 = dp[0], d1 = dp[1];
    uint4 av = make_uint4(0u, 0u, 0u, 0u);
    if (col_ok) av = *reinterpret_cast<const uint4*>(act + (size_t)r * K + j0);
    const uint32_t dd[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
    float dz[CT];                                               // classes >= C are zero in dzn and in wpad
#pragma unroll
    for (int c = 0; c < CT / 2; ++c) { dz[2 * c] = bf_lo(dd[c]); dz[2 * c + 1] = bf_hi(dd[c]); }
    const uint32_t aa[4] = {av.x, av.y, av.z, av.w};
    uint32_t o[4];
#pragma unroll
    for (int u = 0; u < 8; u += 2) {
      float g0 = 0.f, g1 = 0.f;
#pragma unroll
      for (int c = 0; c < CT; ++c) { g0 = fmaf(dz[c], w[u][c], g0); g1 = fmaf(dz[c], w[u + 1][c], g1); }
      const float a0 = bf_lo(aa[u >> 1]), a1 = bf_hi(aa[u >> 1]);
      const __nv_bfloat162 h = __floats2bfloat162_rn(g0 * a0 * (1.0f - a0), g1 * a1 * (1.0f - a1));   // SigmoidGrad
      o[u >> 1] = *reinterpret_cast<const uint32_t*>(&h);
      cs[u] += bf_lo(o[u >> 1]);
      cs[u + 1] += bf_hi(o[u >> 1]);
    }
    if (col_ok) *reinterpret_cast<uint4*>(out + (size_t)r * K + j0) = make_uint4(o[0], o[1], o[2], o[3]);
  }
  if (colsum_part == nullptr) return;
  if constexpr (RL == 1) {                                        // one row lane: the thread's sums are the block's
    if (col_ok) {
      float4* o4 = reinterpret_cast<float4*>(colsum_part + (size_t)blockIdx.y * K + j0);
      o4[0] = make_float4(cs[0], cs[1], cs[2], cs[3]);
      o4[1] = make_float4(cs[4], cs[5], cs[6], cs[7]);
    }
  } else {
#pragma unroll
    for (int u = 0; u < 8; ++u) sh[rl][cg * 8 + u] = cs[u];
    __syncthreads();
    for (int j = threadIdx.x; j < ND_COLS; j += CG * RL) {
      const int col = blockIdx.x * ND_COLS + j;
      float t = sh[0][j];
#pragma unroll
      for (int q = 1; q < RL; ++q) t += sh[q][j];                 // row lanes in a fixed order
      if (col < K) colsum_part[(size_t)blockIdx.y * K + col] = t;
    }
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

int b2_launch_loss_head_rows(const float* z, const float* y, void* dzn, float* loss_part, float* correct_part, float* db_part,
                              int rows, int C, int loss, float inv_global_batch, int blocks, cudaStream_t s) {
  if (C > 16) { b2_set_error("loss_head_rows: C must be <= 16"); return B2_ERR_ARG; }
  mx::loss_head_rows_kernel<<<blocks, 256, 0, s>>>(z, y, reinterpret_cast<__nv_bfloat16*>(dzn), loss_part, correct_part, db_part,
                                                   rows, C, loss, inv_global_batch);
  MX_CHECK();
  return B2_OK;
}

// dz_prev = (dzn . wpad^T) * act * (1 - act) for the narrow output layer; colsum_part (may be NULL) gets one partial row of
// column sums per 64-row block: finish with b2_launch_colsum_finish(part, out, K, b2_narrow_dgrad_slices(rows))
int b2_narrow_dgrad_slices(int rows) { return (rows + mx::ND_ROWS - 1) / mx::ND_ROWS; }
int b2_launch_narrow_dgrad_bf16(const void* dzn, const void* wpad, const void* act, void* out, float* colsum_part, int rows,
                                int K, int C, cudaStream_t s) {
  if (K % 8 || C > 16) { b2_set_error("narrow_dgrad_bf16: K must be a multiple of 8 and C <= 16"); return B2_ERR_ARG; }
  const bool wide = K >= 2048;                                     // whole rows per CTA once a row is a few KB
  const int cols = wide ? 4096 : 512;
  const dim3 grid((K + cols - 1) / cols, (rows + mx::ND_ROWS - 1) / mx::ND_ROWS);
  const __nv_bfloat16* d = reinterpret_cast<const __nv_bfloat16*>(dzn);
  const __nv_bfloat16* w = reinterpret_cast<const __nv_bfloat16*>(wpad);
  const __nv_bfloat16* a = reinterpret_cast<const __nv_bfloat16*>(act);
  __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out);
#define ND_CASE(CT)                                                                                              \
  if (wide) mx::narrow_dgrad_bf16_kernel<CT, 512, 1><<<grid, 512, 0, s>>>(d, w, a, o, colsum_part, rows, K);     \
  else mx::narrow_dgrad_bf16_kernel<CT, 64, 4><<<grid, 256, 0, s>>>(d, w, a, o, colsum_part, rows, K);           \
  break;
  switch ((C + 1) / 2) {
    case 1: ND_CASE(2)
    case 2: ND_CASE(4)
    case 3: ND_CASE(6)
    case 4: ND_CASE(8)
    case 5: ND_CASE(10)
    case 6: ND_CASE(12)
    case 7: ND_CASE(14)
    default: ND_CASE(16)
  }
#undef ND_CASE
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
  long blocks = (P / 4 + 255) / 256;
  if (blocks > 8L * sm_count) blocks = 8L * sm_count;
  if (blocks < 1) blocks = 1;
  mx::sgd_apply_mixed_kernel<<<(int)blocks, 256, 0, s>>>(params, grads, P, lr, a, loss_trace, correct_trace, global_step);
  MX_CHECK();
  return B2_OK;
}
