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
// Kernel: persistent, one CTA per SM.  The products this path exists for have one NARROW output dimension (the 100
// hidden units) and one long one (batch rows in the forward pass, 784 input features in the weight gradient), so the
// narrow side goes on the 128 TMEM lanes (UMMA "A", called P here) and the long side on up to 256 accumulator columns
// (UMMA "B", called Q): 128 x BLOCK_Q x 16 MMAs with BLOCK_Q ~ 208-256 fetch 12 KB of operands per 128+ tensor-pipe
// clocks, where 128 x 128 tiles (the first version, 2.2x slower) were bound by shared-memory operand fetch.  The
// accumulator is the TRANSPOSE of the output tile; the epilogue stores it transposed, which also makes the stores
// coalesced (a warp's 32 lanes are 32 consecutive output columns).
// BLOCK_K = 32; a stage holds the three planes of the P tile (24 KB) and of the Q tile (<= 48 KB); three stages.
// K-major tiles have 64-byte rows -> 64B swizzle; MN-major tiles keep the 128B swizzle (64 MN elements per k row).
//   warp 0      TMA producer (3-D tensor maps: column, row, plane)
//   warp 1      TMEM allocation + tcgen05.mma issue (12 MMAs per stage), two accumulator buffers
//   warps 2-5   epilogue (one per TMEM lane quarter)
#include <cuda.h>
#include <cuda_bf16.h>
#include <math.h>
#include <stdlib.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace s3 {
using namespace tcx;

constexpr int BLOCK_P = 128, MAX_BLOCK_Q = 256, BLOCK_K = 32, UMMA_K = 16;
constexpr int STAGES = 3, PLANES = 3;
constexpr int THREADS = 192;
constexpr int P_TILE_BYTES = BLOCK_P * BLOCK_K * 2;           // one plane of the P tile: 8 KB
constexpr int Q_TILE_BYTES = MAX_BLOCK_Q * BLOCK_K * 2;       // one plane of the Q tile: 16 KB
constexpr int STAGE_BYTES = PLANES * (P_TILE_BYTES + Q_TILE_BYTES);   // 72 KB
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
constexpr int TMEM_COLS = 2 * MAX_BLOCK_Q;

enum { EPI_BIAS_ACT_F32 = 0, EPI_STORE_F32 = 1 };

struct Params {
  int NP, NQ, K;           // narrow extent (lanes), long extent (columns), reduction length
  int block_q;             // columns per tile: multiple of 16, <= 256
  int p_mn, q_mn;          // 1 = MN-major operand (global [K, NP] / [K, NQ] row-major planes), 0 = K-major ([NP, K] / [NQ, K])
  int epi, act;
  const float* bias;       // EPI_BIAS_ACT_F32: [NP]
  float* out;              // fp32 [NQ, ldc] row-major: out[q * ldc + p]
  int ldc;
  int ksplits;             // EPI_STORE_F32: K cut into slices, slice z writes out + z*split_stride (fixed-order partials)
  long split_stride;
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
gemm_split3_kernel(const __grid_constant__ CUtensorMap tmP, const __grid_constant__ CUtensorMap tmQ, Params p) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* full = bars;                 // [STAGES]  TMA -> MMA
  uint64_t* empty = bars + STAGES;       // [STAGES]  MMA -> TMA
  uint64_t* tfull = bars + 2 * STAGES;   // [2]       MMA -> epilogue
  uint64_t* tempty = tfull + 2;          // [2]       epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int BQ = p.block_q;
  const int tiles_p = (p.NP + BLOCK_P - 1) / BLOCK_P, tiles_q = (p.NQ + BQ - 1) / BQ;
  const int pq_tiles = tiles_p * tiles_q;
  const int ntiles = pq_tiles * p.ksplits;
  const int num_kb_total = (p.K + BLOCK_K - 1) / BLOCK_K;
  const int kb_per = (num_kb_total + p.ksplits - 1) / p.ksplits;
  const int q_atoms = (BQ + 63) / 64;                              // MN-major Q tile: 64-column boxes
  const uint32_t stage_tx = PLANES * (P_TILE_BYTES + (p.q_mn ? q_atoms * BLOCK_K * 128 : BQ * BLOCK_K * 2));

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
        const int pq = tile % pq_tiles, z = tile / pq_tiles;
        const int p0 = (pq % tiles_p) * BLOCK_P, q0 = (pq / tiles_p) * BQ;
        const int kb_begin = z * kb_per, kb_end = min(num_kb_total, kb_begin + kb_per);
        for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
          const int s = it % STAGES;
          mbar_wait(&empty[s], ((it / STAGES) & 1) ^ 1);
          mbar_expect_tx(&full[s], stage_tx);
          const int k0 = kb * BLOCK_K;
          unsigned char* st = smem + s * STAGE_BYTES;
#pragma unroll
          for (int pl = 0; pl < PLANES; ++pl) {
            unsigned char* p_dst = st + pl * P_TILE_BYTES;
            unsigned char* q_dst = st + PLANES * P_TILE_BYTES + pl * Q_TILE_BYTES;
            if (p.p_mn) {
              tma_load_3d(p_dst, &tmP, p0, k0, pl, &full[s]);
              tma_load_3d(p_dst + BLOCK_K * 128, &tmP, p0 + 64, k0, pl, &full[s]);
            } else {
              tma_load_3d(p_dst, &tmP, k0, p0, pl, &full[s]);
            }
            if (p.q_mn) {
              for (int a = 0; a < q_atoms; ++a) tma_load_3d(q_dst + a * (BLOCK_K * 128), &tmQ, q0 + 64 * a, k0, pl, &full[s]);
            } else {
              tma_load_3d(q_dst, &tmQ, k0, q0, pl, &full[s]);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // instruction descriptor: D = F32, A = B = BF16, operand majors, N >> 3, M >> 4
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)p.p_mn << 15) | ((uint32_t)p.q_mn << 16) |
                             ((uint32_t)(BQ >> 3) << 17) | ((uint32_t)(BLOCK_P >> 4) << 24);
      const uint32_t p_step = p.p_mn ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
      const uint32_t q_step = p.q_mn ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
      int it = 0, t = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++t) {
        const int acc = t & 1;
        mbar_wait(&tempty[acc], ((t >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * MAX_BLOCK_Q;
        const int z = tile / pq_tiles;
        const int kb_begin = z * kb_per, kb_end = min(num_kb_total, kb_begin + kb_per);
        for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
          const int s = it % STAGES;
          mbar_wait(&full[s], (it / STAGES) & 1);
          tc_fence_after();
          const uint32_t st = smem_u32(smem + s * STAGE_BYTES);
          // the six kept partial products, smallest weight first: (p3,q1) (p2,q2) (p1,q3) (p2,q1) (p1,q2) (p1,q1)
          constexpr int PP[6] = {2, 1, 0, 1, 0, 0};
          constexpr int PQ[6] = {0, 1, 2, 0, 1, 0};
#pragma unroll
          for (int q = 0; q < 6; ++q) {
            const uint64_t dp = make_desc(st + PP[q] * P_TILE_BYTES, p.p_mn);
            const uint64_t dq = make_desc(st + PLANES * P_TILE_BYTES + PQ[q] * Q_TILE_BYTES, p.q_mn);
#pragma unroll
            for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
              tc_mma(d_tmem, dp + (uint64_t)(k * p_step), dq + (uint64_t)(k * q_step), idesc,
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
      const int pq = tile % pq_tiles, z = tile / pq_tiles;
      const int p0 = (pq % tiles_p) * BLOCK_P, q0 = (pq / tiles_p) * BQ;
      mbar_wait(&tfull[acc], (t >> 1) & 1);
      tc_fence_after();
      const int pi = p0 + quarter * 32 + lane;              // this thread's output column (narrow index)
      const bool p_ok = pi < p.NP;
      const bool empty_split = z * kb_per >= num_kb_total;
      const float bias = (p.epi == EPI_BIAS_ACT_F32 && p_ok) ? p.bias[pi] : 0.f;
      float* obase = p.out + (size_t)z * p.split_stride + pi;
#pragma unroll 1
      for (int c0 = 0; c0 < BQ; c0 += 32) {
        if (q0 + c0 >= p.NQ) break;                         // warp-uniform
        uint32_t r[32];
        tc_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * MAX_BLOCK_Q + c0), r);
        if (!p_ok) continue;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int qi = q0 + c0 + j;
          if (c0 + j < BQ && qi < p.NQ) {
            float v = empty_split ? 0.f : __uint_as_float(r[j]);
            if (p.epi == EPI_BIAS_ACT_F32) {
              v += bias;
              if (p.act == B2_ACT_SIGMOID) v = 1.0f / (1.0f + expf(-v));         // distributed.py:97
            }
            obase[(size_t)qi * p.ldc] = v;                  // lanes = consecutive columns of one output row: coalesced
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

// Tile width along the long dimension and number of K slices: fewest waves of the persistent grid, widest tile on ties.
// max_splits <= 1: no K split (the forward pass).
void b2_gemm_split3_plan(int M, int N, int K, int max_splits, int sm_count, int* block_q, int* splits) {
  using namespace s3;
  const int num_kb = (K + BLOCK_K - 1) / BLOCK_K;
  const long tp = (N + BLOCK_P - 1) / BLOCK_P;
  long best_cost = -1;
  for (int bq = 256; bq >= 128; bq -= 16) {
    const long tq = (M + bq - 1) / bq;
    long S = 1;
    if (max_splits > 1) {
      S = sm_count / (tp * tq);
      if (S > max_splits) S = max_splits;
      if (S > num_kb) S = num_kb;
      if (S < 1) S = 1;
    }
    const long waves = (tp * tq * S + sm_count - 1) / sm_count;
    const long kb_per = (num_kb + S - 1) / S;
    const long cost = waves * kb_per * (bq + 64);
    if (best_cost < 0 || cost < best_cost) { best_cost = cost; *block_q = bq; *splits = (int)S; }
  }
}

// out[M, ldc] (fp32) = A[M, K] . B[K, N], both operands as three bf16 planes.
// A planes: K-major -> [3][M, lda] (K real columns); MN-major -> [3][K, lda] (M real columns).  B planes likewise with N.
// N is the narrow side (TMEM lanes), M the long one (accumulator columns, block_q per tile from b2_gemm_split3_plan).
// epi 0: act(acc + bias[n]); epi 1: raw partial sums, K cut into `ksplits` slices, slice z at out + z*split_stride.
int b2_launch_gemm_split3(const void* A, int a_mn, long lda, long a_plane, const void* B, int b_mn, long ldb, long b_plane,
                          int M, int N, int K, int epi, int act, const float* bias, float* out, int ldc, int block_q,
                          int ksplits, long split_stride, int sm_count, cudaStream_t s) {
  using namespace s3;
  if (((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B)) & 15) || lda % 8 || ldb % 8 || a_plane % 8 ||
      b_plane % 8) {
    b2_set_error("gemm_split3: planes must be 16-byte aligned with pitches that are multiples of 8 elements");
    return B2_ERR_ARG;
  }
  if (block_q < 16 || block_q > MAX_BLOCK_Q || block_q % 16) {
    b2_set_error("gemm_split3: tile width %d must be a multiple of 16 in [16, 256]", block_q);
    return B2_ERR_ARG;
  }
  // P = the B operand (narrow side), Q = the A operand (long side)
  CUtensorMap tp, tq;
  int rc = b_mn ? make_map3(&tp, B, K, N, ldb, b_plane, BLOCK_K, true) : make_map3(&tp, B, N, K, ldb, b_plane, BLOCK_P, false);
  if (rc) return rc;
  rc = a_mn ? make_map3(&tq, A, K, M, lda, a_plane, BLOCK_K, true) : make_map3(&tq, A, M, K, lda, a_plane, block_q, false);
  if (rc) return rc;
  Params p;
  p.NP = N; p.NQ = M; p.K = K; p.block_q = block_q; p.p_mn = b_mn; p.q_mn = a_mn; p.epi = epi; p.act = act; p.bias = bias;
  p.out = out; p.ldc = ldc;
  const int num_kb = (K + BLOCK_K - 1) / BLOCK_K;
  p.ksplits = (epi != EPI_STORE_F32 || ksplits < 1) ? 1 : (ksplits > num_kb ? num_kb : ksplits);
  p.split_stride = split_stride;
  static unsigned long long attr_done = 0ull;
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  if (dev >= 64 || !((attr_done >> dev) & 1ull)) {
    B2_CUDA(cudaFuncSetAttribute(gemm_split3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    if (dev < 64) attr_done |= 1ull << dev;
  }
  const long ntiles = (long)((N + BLOCK_P - 1) / BLOCK_P) * ((M + block_q - 1) / block_q) * p.ksplits;
  const int grid = ntiles < sm_count ? (int)ntiles : sm_count;
  gemm_split3_kernel<<<grid, THREADS, SMEM_BYTES, s>>>(tp, tq, p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { b2_set_error("gemm_split3 launch failed: %s", cudaGetErrorString(e)); return B2_ERR_CUDA; }
  return B2_OK;
}
