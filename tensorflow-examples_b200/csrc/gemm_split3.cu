// fp32-accurate GEMM on the bf16 tensor cores ("split-3"): the wide layer of the reference's own model
// (distributed.py:96-97, 784 -> 100 at fp32) is compute-bound on the CUDA cores, so its two heavy products
//   forward  a[M,N]  = act(x[M,K] . W[K,N] + b)          (distributed.py:96-97)
//   wgrad    dW[K,N] = x[M,K]^T . dz[M,N]                (TF1 MatMul gradient, transpose_a)
// run on tcgen05 instead WITHOUT giving up fp32 results: every fp32 operand v is written as the exact sum of three
// bfloat16 planes v = v1 + v2 + v3 (8 + 8 + 8 mantissa bits, split3_kernel below), and a product a.b is accumulated as
// the six partial products whose weight is >= 2^-16 of the leading one,
//   a1b1 + (a1b2 + a2b1) + (a1b3 + a2b2 + a3b1),
// each exact in the fp32 accumulator (8 x 8 bit mantissas); the dropped terms are <= 2^-24 relative, i.e. below fp32
// rounding.  Six bf16 MMAs cost what three TF32 MMAs would, and the planes take 6 B per element instead of 8.
//
// Kernel: persistent, one CTA per SM, 128 x 128 output tiles, BLOCK_K = 32.  A shared-memory stage holds the three
// planes of the A tile and the three planes of the B tile (48 KB; four stages), so every plane is fetched once per
// k-block and all six products read it from shared memory.  (BLOCK_K = 64 with two 96 KB stages measured 2.5x
// slower: one stage in flight cannot cover the load latency.)  K-major tiles have 64-byte rows -> 64B swizzle;
// MN-major tiles keep the 128B swizzle (64 MN elements per k row).
//   warp 0      TMA producer (3-D tensor maps: column, row, plane)
//   warp 1      TMEM allocation + tcgen05.mma issue (24 MMAs per stage), two accumulator buffers
//   warps 2-5   epilogue (one per TMEM lane quarter)
#include <cuda.h>
#include <cuda_bf16.h>
#include <math.h>
#include <stdlib.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace s3 {
using namespace tcx;

constexpr int BLOCK_M = 128, BLOCK_N = 128, BLOCK_K = 32, UMMA_K = 16;
constexpr int STAGES = 4, PLANES = 3;
constexpr int THREADS = 192;
constexpr int TILE_BYTES = 128 * BLOCK_K * 2;                 // one plane of an operand tile: 8 KB
constexpr int STAGE_BYTES = 2 * PLANES * TILE_BYTES;          // 48 KB
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
constexpr int TMEM_COLS = 2 * BLOCK_N;

enum { EPI_BIAS_ACT_F32 = 0, EPI_STORE_F32 = 1 };

struct Params {
  int M, N, K;
  int a_mn, b_mn;          // 1 = MN-major operand (global [K, M] / [K, N] row-major planes), 0 = K-major ([M, K] / [N, K])
  int epi, act;
  const float* bias;       // EPI_BIAS_ACT_F32: [N]
  float* out;              // fp32 [M, ldc] row-major
  int ldc;
  int ksplits;             // EPI_STORE_F32: K cut into slices, slice z writes out + z*split_stride (fixed-order partials)
  long split_stride;
  int debug;
};

// Shared-memory matrix descriptor (bit layout as in gemm_tcgen05.cu).
//   MN-major tile: 128B swizzle (layout 2); 64 MN elements per k row, 8-k-row atoms 1024 B apart (SBO), the next 64 MN
//                  elements BLOCK_K rows * 128 B further (LBO)
//   K-major tile:  rows of BLOCK_K = 32 bf16 = 64 B -> 64B swizzle (layout 4); 8-row groups 512 B apart (SBO)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, bool mn_major) {
  const uint64_t lbo = mn_major ? (uint64_t)((BLOCK_K * 128) >> 4) : 1ull;
  const uint64_t sbo = mn_major ? (1024 >> 4) : (512 >> 4);
  const uint64_t layout = mn_major ? 2ull : 4ull;
  return (uint64_t)((saddr & 0x3FFFF) >> 4) | (lbo << 16) | (sbo << 32) | (1ull << 46) | (layout << 61);
}

__global__ void __launch_bounds__(THREADS, 1)
gemm_split3_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, Params p) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* full = bars;                 // [STAGES]  TMA -> MMA
  uint64_t* empty = bars + STAGES;       // [STAGES]  MMA -> TMA
  uint64_t* tfull = bars + 2 * STAGES;   // [2]       MMA -> epilogue
  uint64_t* tempty = tfull + 2;          // [2]       epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_m = (p.M + BLOCK_M - 1) / BLOCK_M, tiles_n = (p.N + BLOCK_N - 1) / BLOCK_N;
  const int mn_tiles = tiles_m * tiles_n;
  const int ntiles = mn_tiles * p.ksplits;
  const int num_kb_total = (p.K + BLOCK_K - 1) / BLOCK_K;
  const int kb_per = (num_kb_total + p.ksplits - 1) / p.ksplits;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull[a], 1); mbar_init(&tempty[a], 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int it = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int mn = tile % mn_tiles, z = tile / mn_tiles;
        const int m0 = (mn / tiles_n) * BLOCK_M, n0 = (mn % tiles_n) * BLOCK_N;
        const int kb_begin = z * kb_per, kb_end = min(num_kb_total, kb_begin + kb_per);
        for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
          const int s = it % STAGES;
          mbar_wait(&empty[s], ((it / STAGES) & 1) ^ 1);
          if ((p.debug & 1) && it >= STAGES) { mbar_arrive(&full[s]); continue; }
          mbar_expect_tx(&full[s], STAGE_BYTES);
          const int k0 = kb * BLOCK_K;
          unsigned char* st = smem + s * STAGE_BYTES;
#pragma unroll
          for (int pl = 0; pl < PLANES; ++pl) {
            unsigned char* a_dst = st + pl * TILE_BYTES;
            unsigned char* b_dst = st + (PLANES + pl) * TILE_BYTES;
            if (p.a_mn) {
              tma_load_3d(a_dst, &tmA, m0, k0, pl, &full[s]);
              tma_load_3d(a_dst + BLOCK_K * 128, &tmA, m0 + 64, k0, pl, &full[s]);
            } else {
              tma_load_3d(a_dst, &tmA, k0, m0, pl, &full[s]);
            }
            if (p.b_mn) {
              tma_load_3d(b_dst, &tmB, n0, k0, pl, &full[s]);
              tma_load_3d(b_dst + BLOCK_K * 128, &tmB, n0 + 64, k0, pl, &full[s]);
            } else {
              tma_load_3d(b_dst, &tmB, k0, n0, pl, &full[s]);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // instruction descriptor: D = F32, A = B = BF16, operand majors, N >> 3, M >> 4
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)p.a_mn << 15) | ((uint32_t)p.b_mn << 16) |
                             ((uint32_t)(BLOCK_N >> 3) << 17) | ((uint32_t)(BLOCK_M >> 4) << 24);
      const uint32_t a_step = p.a_mn ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
      const uint32_t b_step = p.b_mn ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
      int it = 0, t = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++t) {
        const int acc = t & 1;
        mbar_wait(&tempty[acc], ((t >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BLOCK_N;
        const int z = tile / mn_tiles;
        const int kb_begin = z * kb_per, kb_end = min(num_kb_total, kb_begin + kb_per);
        for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
          const int s = it % STAGES;
          mbar_wait(&full[s], (it / STAGES) & 1);
          tc_fence_after();
          const uint32_t st = smem_u32(smem + s * STAGE_BYTES);
          // the six kept partial products, smallest weight first: (a3,b1) (a2,b2) (a1,b3) (a2,b1) (a1,b2) (a1,b1)
          constexpr int PA[6] = {2, 1, 0, 1, 0, 0};
          constexpr int PB[6] = {0, 1, 2, 0, 1, 0};
#pragma unroll
          for (int q = 0; q < 6; ++q) {
            const uint64_t da = make_desc(st + PA[q] * TILE_BYTES, p.a_mn);
            const uint64_t db = make_desc(st + (PLANES + PB[q]) * TILE_BYTES, p.b_mn);
#pragma unroll
            for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
              tc_mma(d_tmem, da + (uint64_t)(k * a_step), db + (uint64_t)(k * b_step), idesc,
                     (kb != kb_begin) || (q != 0) || (k != 0));
          }
          tc_commit(&empty[s]);
        }
        tc_commit(&tfull[acc]);
      }
    }
  } else {
    const int quarter = warp & 3;                           // TMEM lanes 32*quarter .. +31 belong to this warp
    int t = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++t) {
      const int acc = t & 1;
      const int mn = tile % mn_tiles, z = tile / mn_tiles;
      const int m0 = (mn / tiles_n) * BLOCK_M, n0 = (mn % tiles_n) * BLOCK_N;
      mbar_wait(&tfull[acc], (t >> 1) & 1);
      tc_fence_after();
      const int row = m0 + quarter * 32 + lane;
      const bool row_ok = row < p.M;
      const bool empty_split = z * kb_per >= num_kb_total;
      const bool vec4 = (p.ldc & 3) == 0 && (p.split_stride & 3) == 0;
#pragma unroll 1
      for (int c0 = 0; c0 < BLOCK_N; c0 += 32) {
        const int n = n0 + c0;
        if (n >= p.N) break;                                // warp-uniform
        uint32_t r[32];
        tc_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BLOCK_N + c0), r);
        if (empty_split) {
#pragma unroll
          for (int j = 0; j < 32; ++j) r[j] = 0u;
        }
        if (!row_ok || (p.debug & 2)) continue;
        float* o = p.out + (size_t)z * p.split_stride + (size_t)row * p.ldc + n;
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          float v[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            v[u] = __uint_as_float(r[j + u]);
            if (p.epi == EPI_BIAS_ACT_F32) {
              const float zv = v[u] + (n + j + u < p.N ? p.bias[n + j + u] : 0.f);
              v[u] = p.act == B2_ACT_SIGMOID ? 1.0f / (1.0f + expf(-zv)) : zv;       // distributed.py:97
            }
          }
          if (vec4 && n + j + 4 <= p.N) {
            *reinterpret_cast<float4*>(o + j) = make_float4(v[0], v[1], v[2], v[3]);
          } else {
#pragma unroll
            for (int u = 0; u < 4; ++u)
              if (n + j + u < p.N) o[j + u] = v[u];
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)TMEM_COLS)
                 : "memory");
  }
}

// v -> three bf16 planes with v == p1 + p2 + p3 exactly (for |v| well inside the bf16 range).  in: fp32 [rows, cols]
// (row pitch ld_in); out: planes [3][rows, ld_out] (plane pitch plane_stride elements), columns >= cols zero filled.
// One thread per group of 8 columns.
__global__ void __launch_bounds__(256)
split3_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, long rows, int cols, int ld_in, int ld_out,
              long plane_stride) {
  const int groups = ld_out >> 3;
  const long total = rows * groups;
  const bool vec_in = (ld_in & 3) == 0 && (reinterpret_cast<uintptr_t>(in) & 15) == 0;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const long r = i / groups;
    const int c = (int)(i - r * groups) * 8;
    float v[8];
    const float* src = in + r * ld_in + c;
    if (vec_in && c + 8 <= cols) {
      const float4 a = *reinterpret_cast<const float4*>(src), b = *reinterpret_cast<const float4*>(src + 4);
      v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = c + j < cols ? src[j] : 0.f;
    }
    uint32_t w[3][4];
#pragma unroll
    for (int j = 0; j < 8; j += 2) {
      float r0 = v[j], r1 = v[j + 1];
#pragma unroll
      for (int pl = 0; pl < 3; ++pl) {
        const __nv_bfloat16 h0 = __float2bfloat16_rn(r0), h1 = __float2bfloat16_rn(r1);
        r0 -= __bfloat162float(h0);     // exact: the difference has at most 16 significant bits left, then 8
        r1 -= __bfloat162float(h1);
        w[pl][j >> 1] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
      }
    }
#pragma unroll
    for (int pl = 0; pl < 3; ++pl)
      *reinterpret_cast<uint4*>(out + pl * plane_stride + r * ld_out + c) = make_uint4(w[pl][0], w[pl][1], w[pl][2], w[pl][3]);
  }
}

// fp32 column sums of dz over `slices` row slices: partB[z*stride + n] (the BiasAdd gradient partials of a wide layer)
__global__ void __launch_bounds__(256)
colsum_f32_kernel(const float* __restrict__ dz, float* __restrict__ partB, long part_stride, int rows, int N, int rps) {
  __shared__ float red[8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n = blockIdx.x * 32 + lane;
  const int r0 = blockIdx.y * rps, r1 = min(rows, r0 + rps);
  float acc = 0.f;
  if (n < N) {
    int r = r0 + warp;
    for (; r + 56 < r1; r += 64) {            // eight independent loads in flight per lane
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = dz[(size_t)(r + 8 * u) * N + n];
      acc += ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]));
    }
    for (; r < r1; r += 8) acc += dz[(size_t)r * N + n];
  }
  red[warp][lane] = acc;
  __syncthreads();
  if (warp == 0 && n < N) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += red[w][lane];
    partB[(size_t)blockIdx.y * part_stride + n] = t;
  }
}

// 3-D map over the planes of a row-major bf16 matrix [rows, ld] (only `cols` columns are real): {col, row, plane}
// K-major operand: box = {BLOCK_K columns (64 B), box_rows}, 64B swizzle.  MN-major: box = {64 columns (128 B), BLOCK_K rows},
// 128B swizzle.
static int make_map3(CUtensorMap* m, const void* base, long rows, long cols, long ld, long plane_stride, int box_rows,
                     bool mn_major) {
  pfn_encode enc = b2_tensor_map_encoder();
  if (!enc) return B2_ERR_CUDA;
  const cuuint64_t gdim[3] = {(cuuint64_t)cols, (cuuint64_t)rows, 3};
  const cuuint64_t gstr[2] = {(cuuint64_t)ld * 2, (cuuint64_t)plane_stride * 2};
  const cuuint32_t box[3] = {mn_major ? 64u : (cuuint32_t)BLOCK_K, (cuuint32_t)box_rows, 1u};
  const cuuint32_t estr[3] = {1u, 1u, 1u};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, mn_major ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    b2_set_error("cuTensorMapEncodeTiled failed (%d) for the planes of a [%ld, %ld] matrix (pitch %ld)", (int)r, rows, cols, ld);
    return B2_ERR_CUDA;
  }
  return B2_OK;
}

}  // namespace s3

int b2_launch_split3(const float* in, void* planes, long rows, int cols, int ld_in, int ld_out, long plane_stride,
                     int sm_count, cudaStream_t s) {
  if (ld_out % 8 || ld_out < cols || plane_stride % 8 || (reinterpret_cast<uintptr_t>(planes) & 15)) {
    b2_set_error("split3: plane pitch must be a multiple of 8 elements and the planes 16-byte aligned");
    return B2_ERR_ARG;
  }
  long blocks = (rows * (ld_out / 8) + 255) / 256;
  if (blocks > 16L * sm_count) blocks = 16L * sm_count;
  if (blocks < 1) blocks = 1;
  s3::split3_kernel<<<(int)blocks, 256, 0, s>>>(in, reinterpret_cast<__nv_bfloat16*>(planes), rows, cols, ld_in, ld_out,
                                               plane_stride);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { b2_set_error("split3 launch failed: %s", cudaGetErrorString(e)); return B2_ERR_CUDA; }
  return B2_OK;
}

int b2_launch_colsum_f32(const float* dz, float* partB, long part_stride, int rows, int N, int slices, cudaStream_t s) {
  const int rps = (rows + slices - 1) / slices;
  s3::colsum_f32_kernel<<<dim3((N + 31) / 32, slices), 256, 0, s>>>(dz, partB, part_stride, rows, N, rps);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { b2_set_error("colsum_f32 launch failed: %s", cudaGetErrorString(e)); return B2_ERR_CUDA; }
  return B2_OK;
}

// A planes: K-major -> [3][M, lda] (K real columns); MN-major -> [3][K, lda] (M real columns).  B planes likewise with N.
// out: fp32 [M, ldc].  epi 0: act(acc + bias); epi 1: raw partial sums, K cut into `ksplits` slices.
int b2_launch_gemm_split3(const void* A, int a_mn, long lda, long a_plane, const void* B, int b_mn, long ldb, long b_plane,
                          int M, int N, int K, int epi, int act, const float* bias, float* out, int ldc, int ksplits,
                          long split_stride, int sm_count, cudaStream_t s) {
  using namespace s3;
  if (((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B)) & 15) || lda % 8 || ldb % 8 || a_plane % 8 ||
      b_plane % 8) {
    b2_set_error("gemm_split3: planes must be 16-byte aligned with pitches that are multiples of 8 elements");
    return B2_ERR_ARG;
  }
  CUtensorMap ta, tb;
  int rc = a_mn ? make_map3(&ta, A, K, M, lda, a_plane, BLOCK_K, true) : make_map3(&ta, A, M, K, lda, a_plane, BLOCK_M, false);
  if (rc) return rc;
  rc = b_mn ? make_map3(&tb, B, K, N, ldb, b_plane, BLOCK_K, true) : make_map3(&tb, B, N, K, ldb, b_plane, BLOCK_N, false);
  if (rc) return rc;
  Params p;
  p.M = M; p.N = N; p.K = K; p.a_mn = a_mn; p.b_mn = b_mn; p.epi = epi; p.act = act; p.bias = bias; p.out = out; p.ldc = ldc;
  const int num_kb = (K + BLOCK_K - 1) / BLOCK_K;
  p.ksplits = (epi != EPI_STORE_F32 || ksplits < 1) ? 1 : (ksplits > num_kb ? num_kb : ksplits);
  p.split_stride = split_stride;
  { const char* e = getenv("B2_S3_DEBUG"); p.debug = e ? atoi(e) : 0; }
  static unsigned long long attr_done = 0ull;
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  if (dev >= 64 || !((attr_done >> dev) & 1ull)) {
    B2_CUDA(cudaFuncSetAttribute(gemm_split3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    if (dev < 64) attr_done |= 1ull << dev;
  }
  const int ntiles = ((M + BLOCK_M - 1) / BLOCK_M) * ((N + BLOCK_N - 1) / BLOCK_N) * p.ksplits;
  const int grid = ntiles < sm_count ? ntiles : sm_count;
  gemm_split3_kernel<<<grid, THREADS, SMEM_BYTES, s>>>(ta, tb, p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { b2_set_error("gemm_split3 launch failed: %s", cudaGetErrorString(e)); return B2_ERR_CUDA; }
  return B2_OK;
}
