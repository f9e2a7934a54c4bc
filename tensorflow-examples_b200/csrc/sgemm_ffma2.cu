// fp32 GEMM micro-kernel on the CUDA cores with packed fp32x2 FMAs (FFMA2): the two heavy products of the
// generic fp32 layer path (distributed.py's 784-100-10 model and every other non-fused shape):
//   forward  C[M,N]  = act(A[M,K] . W[K,N] + b)      (distributed.py:96-98)
//   wgrad    dW[K,N] = sum over a batch slice of A[r,K]^T dZ[r,N], db[N] = column sums of dZ   (TF1 MatMul / BiasAdd grads)
// Both are  C[m][n] = sum_kk Aop[kk][m] * Bop[kk][n]  over tiles staged in shared memory "reduction-major":
// a thread owns 8 rows x TN columns; the rows are FFMA2 pairs (a 64-bit register pair straight out of an
// LDS.128 of the A tile), the column value is FFMA2's scalar-broadcast operand, so there is no packing code.
// 128 x (16*TN) tile, BK = 16, 256 threads, register-prefetched global loads, two smem stages.
#include <math.h>

#include "common.cuh"

namespace sg {

typedef unsigned long long u64;
constexpr int BM = 128, BK = 16, THREADS = 256;
constexpr int BMP = BM + 4;   // padded row of the A tile (floats); keeps 16 B alignment

__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) {
  u64 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ u64 pack2(float lo, float hi) {
  u64 d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi));
  return d;
}
__device__ __forceinline__ void unpack2(u64 v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}

enum { EPI_FWD = 0, EPI_WGRAD = 1 };

struct Params {
  const float* A;      // EPI_FWD: [M, K] row-major (transposed on the way into smem); EPI_WGRAD: [rows, M] row-major
  const float* B;      // EPI_FWD: W [K, N];  EPI_WGRAD: dZ [rows, N]
  int M, N;            // output tile space
  int K;               // reduction length (EPI_FWD: in-features; EPI_WGRAD: batch rows)
  int chunk;           // EPI_WGRAD: rows per batch slice (blockIdx.z), multiple of BK
  const float* bias;   // EPI_FWD
  int act;
  float* C;            // EPI_FWD: [M, N]
  float* partW;        // EPI_WGRAD: partW[z*stride + m*N + n], partB[z*stride + n]
  float* partB;
  long part_stride;
};

template <int TN, int EPI>
__global__ void __launch_bounds__(THREADS, 2)
sgemm_kernel(Params p) {
  constexpr int BN = 16 * TN;
  constexpr int B_ELEMS = BK * BN / THREADS;          // elements of the B tile each thread stages (TN)
  __shared__ __align__(16) float As[2][BK][BMP];
  __shared__ __align__(16) float Bs[2][BK][BN];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  int k_begin = 0, k_end = p.K;
  if (EPI == EPI_WGRAD) {
    k_begin = blockIdx.z * p.chunk;
    k_end = min(p.K, k_begin + p.chunk);
  }

  u64 acc[4][TN];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0ull;
  float bsum[TN];
#pragma unroll
  for (int j = 0; j < TN; ++j) bsum[j] = 0.f;

  // ---- global -> register staging ----
  float ra[8], rb[B_ELEMS];
  auto load_tiles = [&](int k0) {
    if (EPI == EPI_FWD) {
      // A[M,K]: this thread takes row (tid & 127), 8 consecutive k starting at (tid >> 7) * 8
      const int m = m0 + (tid & 127), k = k0 + (tid >> 7) * 8;
#pragma unroll
      for (int j = 0; j < 8; ++j) ra[j] = 0.f;
      if (m < p.M) {
        const float* src = p.A + (size_t)m * p.K + k;
        if (((p.K & 3) == 0) && k + 8 <= k_end) {
          const float4 v0 = *reinterpret_cast<const float4*>(src), v1 = *reinterpret_cast<const float4*>(src + 4);
          ra[0] = v0.x; ra[1] = v0.y; ra[2] = v0.z; ra[3] = v0.w; ra[4] = v1.x; ra[5] = v1.y; ra[6] = v1.z; ra[7] = v1.w;
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (k + j < k_end) ra[j] = src[j];
        }
      }
    } else {
      // A[rows,M]: reduction row (tid >> 4), 8 consecutive m starting at (tid & 15) * 8
      const int r = k0 + (tid >> 4), m = m0 + (tid & 15) * 8;
#pragma unroll
      for (int j = 0; j < 8; ++j) ra[j] = 0.f;
      if (r < k_end) {
        const float* src = p.A + (size_t)r * p.M + m;
        if (((p.M & 3) == 0) && m + 8 <= p.M) {
          const float4 v0 = *reinterpret_cast<const float4*>(src), v1 = *reinterpret_cast<const float4*>(src + 4);
          ra[0] = v0.x; ra[1] = v0.y; ra[2] = v0.z; ra[3] = v0.w; ra[4] = v1.x; ra[5] = v1.y; ra[6] = v1.z; ra[7] = v1.w;
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (m + j < p.M) ra[j] = src[j];
        }
      }
    }
#pragma unroll
    for (int i = 0; i < B_ELEMS; ++i) {
      const int e = tid + THREADS * i, kk = e / BN, n = e - kk * BN;
      const int r = k0 + kk;
      rb[i] = (r < k_end && n0 + n < p.N) ? p.B[(size_t)r * p.N + n0 + n] : 0.f;
    }
  };
  auto store_tiles = [&](int buf) {
    if (EPI == EPI_FWD) {
      const int row = tid & 127, kq = (tid >> 7) * 8;
#pragma unroll
      for (int j = 0; j < 8; ++j) As[buf][kq + j][row] = ra[j];
    } else {
      const int kk = tid >> 4, mq = (tid & 15) * 8;
      *reinterpret_cast<float4*>(&As[buf][kk][mq]) = make_float4(ra[0], ra[1], ra[2], ra[3]);
      *reinterpret_cast<float4*>(&As[buf][kk][mq + 4]) = make_float4(ra[4], ra[5], ra[6], ra[7]);
    }
#pragma unroll
    for (int i = 0; i < B_ELEMS; ++i) {
      const int e = tid + THREADS * i, kk = e / BN, n = e - kk * BN;
      Bs[buf][kk][n] = rb[i];
    }
  };

  int buf = 0;
  if (k_begin < k_end) {
    load_tiles(k_begin);
    store_tiles(0);
  }
  __syncthreads();
  for (int k0 = k_begin; k0 < k_end; k0 += BK) {
    const bool more = k0 + BK < k_end;
    if (more) load_tiles(k0 + BK);            // global loads of the next tile fly while this one is multiplied
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      const ulonglong2 a01 = *reinterpret_cast<const ulonglong2*>(&As[buf][kk][ty * 8]);
      const ulonglong2 a23 = *reinterpret_cast<const ulonglong2*>(&As[buf][kk][ty * 8 + 4]);
#pragma unroll
      for (int j = 0; j < TN; ++j) {
        const float b = Bs[buf][kk][tx + 16 * j];
        const u64 bb = pack2(b, b);
        acc[0][j] = ffma2(a01.x, bb, acc[0][j]);
        acc[1][j] = ffma2(a01.y, bb, acc[1][j]);
        acc[2][j] = ffma2(a23.x, bb, acc[2][j]);
        acc[3][j] = ffma2(a23.y, bb, acc[3][j]);
        if (EPI == EPI_WGRAD && ty == 0) bsum[j] += b;
      }
    }
    if (more) store_tiles(buf ^ 1);
    __syncthreads();
    buf ^= 1;
  }

  // ---- epilogue ----
#pragma unroll
  for (int i = 0; i < 4; ++i) {
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int n = n0 + tx + 16 * j;
      if (n >= p.N) continue;
      float lo, hi;
      unpack2(acc[i][j], lo, hi);
      const int m = m0 + ty * 8 + 2 * i;
      if (EPI == EPI_FWD) {
        const float b = p.bias[n];
        if (m < p.M) {
          const float z = lo + b;
          p.C[(size_t)m * p.N + n] = p.act == B2_ACT_SIGMOID ? 1.0f / (1.0f + expf(-z)) : z;
        }
        if (m + 1 < p.M) {
          const float z = hi + b;
          p.C[(size_t)(m + 1) * p.N + n] = p.act == B2_ACT_SIGMOID ? 1.0f / (1.0f + expf(-z)) : z;
        }
      } else {
        float* pw = p.partW + (size_t)blockIdx.z * p.part_stride;
        if (m < p.M) pw[(size_t)m * p.N + n] = lo;
        if (m + 1 < p.M) pw[(size_t)(m + 1) * p.N + n] = hi;
      }
    }
  }
  if (EPI == EPI_WGRAD && blockIdx.y == 0 && ty == 0) {
    float* pb = p.partB + (size_t)blockIdx.z * p.part_stride;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int n = n0 + tx + 16 * j;
      if (n < p.N) pb[n] = bsum[j];
    }
  }
}

template <int EPI>
static int launch(const Params& p, int nz, cudaStream_t s) {
  // widest column tile that does not waste more than it must: N <= 64 -> 64, <= 112 -> 112, else 128
  const dim3 block(THREADS);
  if (p.N <= 64) {
    sgemm_kernel<4, EPI><<<dim3((p.N + 63) / 64, (p.M + BM - 1) / BM, nz), block, 0, s>>>(p);
  } else if (p.N <= 112 || (p.N % 128 != 0 && p.N % 112 == 0)) {
    sgemm_kernel<7, EPI><<<dim3((p.N + 111) / 112, (p.M + BM - 1) / BM, nz), block, 0, s>>>(p);
  } else {
    sgemm_kernel<8, EPI><<<dim3((p.N + 127) / 128, (p.M + BM - 1) / BM, nz), block, 0, s>>>(p);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    b2_set_error("sgemm launch failed: %s", cudaGetErrorString(e));
    return B2_ERR_CUDA;
  }
  return B2_OK;
}

}  // namespace sg

int b2_launch_sgemm_fwd(const float* A, const float* W, const float* bias, float* C, int M, int N, int K, int act,
                        cudaStream_t s) {
  sg::Params p;
  memset(&p, 0, sizeof(p));
  p.A = A; p.B = W; p.M = M; p.N = N; p.K = K; p.bias = bias; p.act = act; p.C = C;
  return sg::launch<sg::EPI_FWD>(p, 1, s);
}

// dW partials: output space is [Kin, N]; reduction over the batch rows, cut into `splits` slices
int b2_launch_sgemm_wgrad(const float* A, const float* dZ, float* partW, float* partB, long part_stride, int rows,
                          int Kin, int N, int splits, cudaStream_t s) {
  sg::Params p;
  memset(&p, 0, sizeof(p));
  int chunk = (rows + splits - 1) / splits;
  chunk = ((chunk + sg::BK - 1) / sg::BK) * sg::BK;
  p.A = A; p.B = dZ; p.M = Kin; p.N = N; p.K = rows; p.chunk = chunk;
  p.partW = partW; p.partB = partB; p.part_stride = part_stride;
  return sg::launch<sg::EPI_WGRAD>(p, splits, s);
}

// ---------------------------------------------------------------------------------------------
// Forward for SMALL batches (M <= 256 rows, e.g. the reference's shipped batch of 100): the tiled kernel above
// would put the whole K loop on one or two CTAs.  Here a CTA owns 8 output columns (its slice of W sits in shared
// memory), its warps share the rows, and the lanes of a warp split K -- x is read coalesced, each row ends with a
// warp-shuffle reduction.  grid = ceil(N/8); latency ~ one pass over K/32 elements per row.
// ---------------------------------------------------------------------------------------------
namespace sg {

constexpr int SM_COLS = 8, SM_THREADS = 256;

__global__ void __launch_bounds__(SM_THREADS)
small_m_fwd_kernel(const float* __restrict__ A, const float* __restrict__ W, const float* __restrict__ bias,
                   float* __restrict__ C, int M, int N, int K, int act) {
  extern __shared__ __align__(16) float s_w[];   // [K][8]
  const int n0 = blockIdx.x * SM_COLS;
  for (int i0 = 0; i0 < K * SM_COLS; i0 += 8 * SM_THREADS) {
    float v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {          // eight independent loads in flight per thread
      const int i = i0 + u * SM_THREADS + threadIdx.x;
      const int k = i >> 3, c = i & 7;
      v[u] = (i < K * SM_COLS && n0 + c < N) ? W[(size_t)k * N + n0 + c] : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = i0 + u * SM_THREADS + threadIdx.x;
      if (i < K * SM_COLS) s_w[i] = v[u];
    }
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // blockIdx.y cuts the rows into gridDim.y interleaved groups so that a small batch still fills many SMs
  for (int m = blockIdx.y * (SM_THREADS / 32) + warp; m < M; m += gridDim.y * (SM_THREADS / 32)) {
    const float* xr = A + (size_t)m * K;
    float acc[SM_COLS];
#pragma unroll
    for (int c = 0; c < SM_COLS; ++c) acc[c] = 0.f;
    for (int kc = 0; kc < K; kc += 1024) {
      // all of this lane's x values of the chunk are requested before any is used: one memory latency per row
      float xv[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const int k = kc + lane + 32 * i;
        xv[i] = k < K ? xr[k] : 0.f;
      }
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const int k = kc + lane + 32 * i;
        if (k < K) {
          const float x = xv[i];
          const float4 w0 = *reinterpret_cast<const float4*>(s_w + k * SM_COLS);
          const float4 w1 = *reinterpret_cast<const float4*>(s_w + k * SM_COLS + 4);
          acc[0] = fmaf(x, w0.x, acc[0]); acc[1] = fmaf(x, w0.y, acc[1]);
          acc[2] = fmaf(x, w0.z, acc[2]); acc[3] = fmaf(x, w0.w, acc[3]);
          acc[4] = fmaf(x, w1.x, acc[4]); acc[5] = fmaf(x, w1.y, acc[5]);
          acc[6] = fmaf(x, w1.z, acc[6]); acc[7] = fmaf(x, w1.w, acc[7]);
        }
      }
    }
#pragma unroll
    for (int c = 0; c < SM_COLS; ++c) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], o);
    }
    if (lane < SM_COLS && n0 + lane < N) {
      float z = acc[0];
#pragma unroll
      for (int c = 1; c < SM_COLS; ++c) z = (lane == c) ? acc[c] : z;
      z += bias[n0 + lane];
      C[(size_t)m * N + n0 + lane] = act == B2_ACT_SIGMOID ? 1.0f / (1.0f + expf(-z)) : z;
    }
  }
}

}  // namespace sg

bool b2_small_m_fwd_applies(int M, int K) { return M <= 256 && (size_t)K * sg::SM_COLS * sizeof(float) <= 160 * 1024; }

int b2_launch_small_m_fwd(const float* A, const float* W, const float* bias, float* C, int M, int N, int K, int act,
                          cudaStream_t s) {
  const int smem = K * sg::SM_COLS * (int)sizeof(float);
  static int attr_smem = 48 * 1024;   // opt-in only ever grows; per process (one device per process in this library's use)
  if (smem > attr_smem) {
    B2_CUDA(cudaFuncSetAttribute(sg::small_m_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_smem = smem;
  }
  const int col_blocks = (N + sg::SM_COLS - 1) / sg::SM_COLS;
  int row_groups = (M + 15) / 16;                       // about two rows per warp
  if (row_groups * col_blocks > 592) row_groups = (592 + col_blocks - 1) / col_blocks;
  if (row_groups < 1) row_groups = 1;
  sg::small_m_fwd_kernel<<<dim3(col_blocks, row_groups), sg::SM_THREADS, smem, s>>>(A, W, bias, C, M, N, K, act);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    b2_set_error("small_m_fwd launch failed: %s", cudaGetErrorString(e));
    return B2_ERR_CUDA;
  }
  return B2_OK;
}
