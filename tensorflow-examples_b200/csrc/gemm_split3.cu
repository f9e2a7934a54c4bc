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
//
// The long operand Q is always the layer INPUT x (fp32, 205 MB per step at batch 65536).  It is never written out as
// planes: TMA brings fp32 tiles into a staging ring and four converter warps split them into the three bf16 planes
// directly in the swizzled shared-memory layout the MMA reads (HBM: 4 B per element per GEMM instead of 4 + 6 + 6 + 6
// with a separate split pass).  The narrow operand P (weights, dz) is small and arrives as planes written by split3_kernel.
// BLOCK_K = 32.  K-major tiles have 64-byte rows -> 64B swizzle; MN-major tiles keep the 128B swizzle.
//   warp 0      TMA producer: fp32 Q tiles (staging ring, 2 x 28 KB) and P planes (3 x 24 KB)
//   warp 1      TMEM allocation + tcgen05.mma issue (12 MMAs per k-block), two accumulator buffers
//   warps 2-5   epilogue (one per TMEM lane quarter)
//   warps 6-13  converters: fp32 tile -> three bf16 planes (2 x 48 KB)
#include <cuda.h>
#include <cuda_bf16.h>
#include <math.h>
#include <stdlib.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace s3 {
using namespace tcx;

constexpr int BLOCK_P = 128, MAX_BLOCK_Q = 224, BLOCK_K = 32, UMMA_K = 16;
constexpr int ACC_COLS = 256;                                 // TMEM columns between the two accumulator buffers
constexpr int PLANES = 3;
constexpr int F_STAGES = 2;                                   // fp32 staging ring (Q tiles as they come from memory)
constexpr int Q_STAGES = 2;                                   // converted Q planes
constexpr int P_STAGES = 3;                                   // P planes (their TMA needs two k-blocks of lead time)
constexpr int CONV_WARPS = 8, CONV_THREADS = 32 * CONV_WARPS;
constexpr int THREADS = 6 * 32 + CONV_THREADS;
constexpr int P_TILE_BYTES = BLOCK_P * BLOCK_K * 2;           // one plane of the P tile: 8 KB
constexpr int Q_TILE_BYTES = ((MAX_BLOCK_Q + 63) / 64) * BLOCK_K * 128;   // one plane of the Q tile: 16 KB (four 64-column atoms)
constexpr int P_STAGE_BYTES = PLANES * P_TILE_BYTES;          // 24 KB
constexpr int Q_STAGE_BYTES = PLANES * Q_TILE_BYTES;          // 48 KB
constexpr int F_STAGE_BYTES = MAX_BLOCK_Q * BLOCK_K * 4;      // 28 KB
constexpr int SMEM_BYTES = P_STAGES * P_STAGE_BYTES + Q_STAGES * Q_STAGE_BYTES + F_STAGES * F_STAGE_BYTES + 1024 + 256;
constexpr int TMEM_COLS = 2 * ACC_COLS;

enum { EPI_BIAS_ACT_F32 = 0, EPI_STORE_F32 = 1 };

struct Params {
  int NP, NQ, K;           // narrow extent (lanes), long extent (columns), reduction length
  int block_q;             // columns per tile: multiple of 16, <= 256
  int p_mn, q_mn;          // 1 = MN-major operand (global [K, NP] planes / [K, NQ] fp32), 0 = K-major ([NP, K] / [NQ, K])
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

// two fp32 values -> their bf16 roundings packed in one word, and the exact remainders
__device__ __forceinline__ uint32_t split_pair(float& a, float& b) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  const uint32_t w = *reinterpret_cast<const uint32_t*>(&h);
  a -= __uint_as_float(w << 16);
  b -= __uint_as_float(w & 0xFFFF0000u);
  return w;
}

__global__ void __launch_bounds__(THREADS, 1)
gemm_split3_kernel(const __grid_constant__ CUtensorMap tmP, const __grid_constant__ CUtensorMap tmQ, Params p) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  unsigned char* sm_p = smem;                                         // [P_STAGES][P_STAGE_BYTES]
  unsigned char* sm_q = sm_p + P_STAGES * P_STAGE_BYTES;              // [Q_STAGES][Q_STAGE_BYTES]
  unsigned char* sm_f32 = sm_q + Q_STAGES * Q_STAGE_BYTES;            // [F_STAGES][F_STAGE_BYTES]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm_f32 + F_STAGES * F_STAGE_BYTES);
  uint64_t* f_full = bars;                       // [F_STAGES]  TMA -> converters
  uint64_t* f_empty = f_full + F_STAGES;         // [F_STAGES]  converters -> TMA
  uint64_t* q_full = f_empty + F_STAGES;         // [Q_STAGES]  converters -> MMA
  uint64_t* q_empty = q_full + Q_STAGES;         // [Q_STAGES]  MMA retired -> converters
  uint64_t* p_full = q_empty + Q_STAGES;         // [P_STAGES]  TMA -> MMA
  uint64_t* p_empty = p_full + P_STAGES;         // [P_STAGES]  MMA retired -> TMA
  uint64_t* tfull = p_empty + P_STAGES;          // [2]         MMA -> epilogue
  uint64_t* tempty = tfull + 2;                  // [2]         epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int BQ = p.block_q;
  const int tiles_p = (p.NP + BLOCK_P - 1) / BLOCK_P, tiles_q = (p.NQ + BQ - 1) / BQ;
  const int pq_tiles = tiles_p * tiles_q;
  const int ntiles = pq_tiles * p.ksplits;
  const int num_kb_total = (p.K + BLOCK_K - 1) / BLOCK_K;
  const int kb_per = (num_kb_total + p.ksplits - 1) / p.ksplits;

  if (threadIdx.x == 0) {
    for (int s = 0; s < F_STAGES; ++s) { mbar_init(&f_full[s], 1); mbar_init(&f_empty[s], CONV_WARPS); }
    for (int s = 0; s < Q_STAGES; ++s) { mbar_init(&q_full[s], CONV_WARPS); mbar_init(&q_empty[s], 1); }
    for (int s = 0; s < P_STAGES; ++s) { mbar_init(&p_full[s], 1); mbar_init(&p_empty[s], 1); }
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
    // ------------------------------------ fp32 Q tiles ------------------------------------
    if (lane == 0) {
      int it = 0;
      // a second cursor runs PF k-blocks ahead of the loads over the same (tile, k-block) sequence and asks L2 for the
      // fp32 tiles: the staging ring has only two stages, which covers an L2 hit but not an HBM round trip
      constexpr int PF = 4;
      int pf_tile = blockIdx.x, pf_kb = 0, pf_end = 0, pf_ahead = 0;
      if (pf_tile < ntiles) {
        const int z = pf_tile / pq_tiles;
        pf_kb = z * kb_per;
        pf_end = min(num_kb_total, pf_kb + kb_per);
      }
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int pq = tile % pq_tiles, z = tile / pq_tiles;
        const int p0 = (pq % tiles_p) * BLOCK_P, q0 = (pq / tiles_p) * BQ;
        const int kb_begin = z * kb_per, kb_end = min(num_kb_total, kb_begin + kb_per);
        for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
          while (pf_ahead <= PF && pf_tile < ntiles) {
            if (pf_kb < pf_end) {
              const int pq0 = ((pf_tile % pq_tiles) / tiles_p) * BQ;
              if (p.q_mn) tma_prefetch_2d(&tmQ, pq0, pf_kb * BLOCK_K);
              else tma_prefetch_2d(&tmQ, pf_kb * BLOCK_K, pq0);
              ++pf_kb;
              ++pf_ahead;
            } else {
              pf_tile += gridDim.x;
              if (pf_tile < ntiles) {
                const int zz = pf_tile / pq_tiles;
                pf_kb = zz * kb_per;
                pf_end = min(num_kb_total, pf_kb + kb_per);
              }
            }
          }
          --pf_ahead;
          const int f = it % F_STAGES, ps = it % P_STAGES;
          const int k0 = kb * BLOCK_K;
          // the fp32 tile first: its staging slot frees up (conversion of k-block i-2 done) before the P slot does
          mbar_wait(&f_empty[f], ((it / F_STAGES) & 1) ^ 1);
          mbar_expect_tx(&f_full[f], (uint32_t)(BQ * BLOCK_K * 4));
          if (p.q_mn) tma_load_2d(sm_f32 + f * F_STAGE_BYTES, &tmQ, q0, k0, &f_full[f]);    // [32 k rows][BQ] fp32
          else tma_load_2d(sm_f32 + f * F_STAGE_BYTES, &tmQ, k0, q0, &f_full[f]);           // [BQ rows][32 k] fp32
          // P planes (from L2 mostly): three stages, so the request leaves two k-blocks before the MMAs need it
          mbar_wait(&p_empty[ps], ((it / P_STAGES) & 1) ^ 1);
          mbar_expect_tx(&p_full[ps], P_STAGE_BYTES);
#pragma unroll
          for (int pl = 0; pl < PLANES; ++pl) {
            unsigned char* p_dst = sm_p + ps * P_STAGE_BYTES + pl * P_TILE_BYTES;
            if (p.p_mn) {
              tma_load_3d(p_dst, &tmP, p0, k0, pl, &p_full[ps]);
              tma_load_3d(p_dst + BLOCK_K * 128, &tmP, p0 + 64, k0, pl, &p_full[ps]);
            } else {
              tma_load_3d(p_dst, &tmP, k0, p0, pl, &p_full[ps]);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------ MMA issuer ------------------------------------
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
        const uint32_t d_tmem = tmem_base + acc * ACC_COLS;
        const int z = tile / pq_tiles;
        const int kb_begin = z * kb_per, kb_end = min(num_kb_total, kb_begin + kb_per);
        for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
          const int qs = it % Q_STAGES, ps = it % P_STAGES;
          mbar_wait(&p_full[ps], (it / P_STAGES) & 1);
          mbar_wait(&q_full[qs], (it / Q_STAGES) & 1);
          tc_fence_after();
          const uint32_t sp = smem_u32(sm_p + ps * P_STAGE_BYTES), sq = smem_u32(sm_q + qs * Q_STAGE_BYTES);
          // the six kept partial products, smallest weight first: (p3,q1) (p2,q2) (p1,q3) (p2,q1) (p1,q2) (p1,q1)
          constexpr int PP[6] = {2, 1, 0, 1, 0, 0};
          constexpr int PQ[6] = {0, 1, 2, 0, 1, 0};
#pragma unroll
          for (int q = 0; q < 6; ++q) {
            const uint64_t dp = make_desc(sp + PP[q] * P_TILE_BYTES, p.p_mn);
            const uint64_t dq = make_desc(sq + PQ[q] * Q_TILE_BYTES, p.q_mn);
#pragma unroll
            for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
              tc_mma(d_tmem, dp + (uint64_t)(k * p_step), dq + (uint64_t)(k * q_step), idesc,
                     (kb != kb_begin) || (q != 0) || (k != 0));
          }
          tc_commit(&p_empty[ps]);
          tc_commit(&q_empty[qs]);
        }
        tc_commit(&tfull[acc]);
      }
    }
  } else if (warp < 6) {
    // ------------------------------------ epilogue ------------------------------------
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
        tc_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * ACC_COLS + c0), r);
        if (!p_ok) continue;
        // all 32 values first, branch-free (independent chains for the exp / reciprocal latencies), then the stores
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          v[j] = empty_split ? 0.f : __uint_as_float(r[j]);
          if (p.epi == EPI_BIAS_ACT_F32) v[j] += bias;
        }
        if (p.epi == EPI_BIAS_ACT_F32 && p.act == B2_ACT_SIGMOID) {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            // 1 / (1 + e^-z), distributed.py:97.  The reciprocal is MUFU.RCP plus one Newton step (<= 1 ulp) instead
            // of an IEEE division: the division's special-case branch per element serialised the 32 chains (the
            // epilogue took 5x longer than the tile's MMAs).  -z is clamped at 80 so that 1 + e^-z stays finite.
            const float x = 1.0f + expf(fminf(-v[j], 80.0f));
            float rc;
            asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(x));
            v[j] = fmaf(rc, fmaf(-x, rc, 1.0f), rc);
          }
        }
        const int nvalid = min(min(32, BQ - c0), p.NQ - (q0 + c0));
        float* o = obase + (size_t)(q0 + c0) * p.ldc;
        if (nvalid == 32) {
#pragma unroll
          for (int j = 0; j < 32; ++j) o[(size_t)j * p.ldc] = v[j];   // lanes = consecutive columns of one output row: coalesced
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < nvalid) o[(size_t)j * p.ldc] = v[j];
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);
    }
  } else {
    // ------------------------------------ converters ------------------------------------
    // The staged tile is dense row-major fp32 [R][CW]: K-major Q -> R = BQ rows of CW = 32 k; MN-major Q -> R = 32 k rows
    // of CW = BQ columns.  An item is one float4; it becomes 8 bytes in each of the three planes.
    const int ct = threadIdx.x - 6 * 32;
    const int CW4 = (p.q_mn ? BQ : BLOCK_K) >> 2;           // float4 per staged row
    const float inv_cw4 = 1.0f / (float)CW4;                // row = floor((i + 0.5) / CW4), exact for i < 2^22 / CW4
    const int items = BQ * BLOCK_K / 4;
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const int z = tile / pq_tiles;
      const int kb_begin = z * kb_per, kb_end = min(num_kb_total, kb_begin + kb_per);
      for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
        const int qs = it % Q_STAGES, f = it % F_STAGES;
        mbar_wait(&q_empty[qs], ((it / Q_STAGES) & 1) ^ 1);           // the MMAs that read this plane stage have retired
        mbar_wait(&f_full[f], (it / F_STAGES) & 1);
        const float4* src = reinterpret_cast<const float4*>(sm_f32 + f * F_STAGE_BYTES);
        unsigned char* q1 = sm_q + qs * Q_STAGE_BYTES;
#pragma unroll 4
        for (int i = ct; i < items; i += CONV_THREADS) {
          float4 v = src[i];
          const int row = (int)(((float)i + 0.5f) * inv_cw4), c4 = i - row * CW4;
          uint32_t off;
          if (p.q_mn) {
            // MN-major, 128B swizzle: atom of 64 columns, k row `row`, 16-byte chunk (col/8) ^ (row & 7)
            const int col = c4 << 2, fl = col & 63;
            off = (uint32_t)((col >> 6) * (BLOCK_K * 128) + row * 128 + ((((fl >> 3) ^ (row & 7)) << 4) | ((fl & 7) << 1)));
          } else {
            // K-major, 64B swizzle: row of 32 k = 64 B, 16-byte chunk (k/8) ^ ((row >> 1) & 3)
            off = (uint32_t)(row * 64 + ((((c4 >> 1) ^ ((row >> 1) & 3)) << 4) | ((c4 & 1) << 3)));
          }
          const uint32_t a0 = split_pair(v.x, v.y), a1 = split_pair(v.z, v.w);
          const uint32_t b0 = split_pair(v.x, v.y), b1 = split_pair(v.z, v.w);
          const __nv_bfloat162 c01 = __floats2bfloat162_rn(v.x, v.y), c23 = __floats2bfloat162_rn(v.z, v.w);
          *reinterpret_cast<uint2*>(q1 + off) = make_uint2(a0, a1);
          *reinterpret_cast<uint2*>(q1 + Q_TILE_BYTES + off) = make_uint2(b0, b1);
          *reinterpret_cast<uint2*>(q1 + 2 * Q_TILE_BYTES + off) =
              make_uint2(*reinterpret_cast<const uint32_t*>(&c01), *reinterpret_cast<const uint32_t*>(&c23));
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // my plane writes -> visible to the tensor core
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&q_full[qs]);
          mbar_arrive(&f_empty[f]);
        }
      }
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
  if (!b2_tensor_map_encoder()) return B2_ERR_CUDA;
  const cuuint64_t gdim[3] = {(cuuint64_t)cols, (cuuint64_t)rows, 3};
  const cuuint64_t gstr[2] = {(cuuint64_t)ld * 2, (cuuint64_t)plane_stride * 2};
  const cuuint32_t box[3] = {mn_major ? 64u : (cuuint32_t)BLOCK_K, (cuuint32_t)box_rows, 1u};
  const int r = b2_tensor_map_cached(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, base, gdim, gstr, box,
                                     mn_major ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B);
  if (r != (int)CUDA_SUCCESS) {
    b2_set_error("cuTensorMapEncodeTiled failed (%d) for the planes of a [%ld, %ld] matrix (pitch %ld)", (int)r, rows, cols, ld);
    return B2_ERR_CUDA;
  }
  return B2_OK;
}

// 2-D map over a row-major fp32 matrix [rows, cols] (row pitch ld floats), dense box {box_cols, box_rows}, no swizzle
static int make_map_f32(CUtensorMap* m, const float* base, long rows, long cols, long ld, int box_cols, int box_rows) {
  if (!b2_tensor_map_encoder()) return B2_ERR_CUDA;
  const cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t gstr[1] = {(cuuint64_t)ld * 4};
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  const int r = b2_tensor_map_cached(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, base, gdim, gstr, box, CU_TENSOR_MAP_SWIZZLE_NONE);
  if (r != (int)CUDA_SUCCESS) {
    b2_set_error("cuTensorMapEncodeTiled failed (%d) for a [%ld, %ld] fp32 matrix (pitch %ld)", (int)r, rows, cols, ld);
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
  for (int bq = MAX_BLOCK_Q; bq >= 32; bq -= 16) {      // narrow tiles only win when the long side is short (few tiles)
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

// out[M, ldc] (fp32) = A . B with fp32 accuracy.  A is the layer input in fp32, converted inside the kernel:
//   a_mn = 0: A is [M, K] row-major (pitch lda floats)      a_mn = 1: A is [K, M] row-major (the product is A^T . B)
// B arrives as three bf16 planes: K-major -> [3][N, ldb] (K real columns); MN-major -> [3][K, ldb] (N real columns).
// N is the narrow side (TMEM lanes), M the long one (accumulator columns, block_q per tile from b2_gemm_split3_plan).
// epi 0: act(acc + bias[n]); epi 1: raw partial sums, K cut into `ksplits` slices, slice z at out + z*split_stride.
int b2_launch_gemm_split3(const float* A, int a_mn, long lda, const void* B, int b_mn, long ldb, long b_plane,
                          int M, int N, int K, int epi, int act, const float* bias, float* out, int ldc, int block_q,
                          int ksplits, long split_stride, int sm_count, cudaStream_t s) {
  using namespace s3;
  if (((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B)) & 15) || lda % 4 || ldb % 8 || b_plane % 8) {
    b2_set_error("gemm_split3: operands must be 16-byte aligned with row pitches that are multiples of 16 bytes");
    return B2_ERR_ARG;
  }
  if (block_q < 16 || block_q > MAX_BLOCK_Q || block_q % 16) {
    b2_set_error("gemm_split3: tile width %d must be a multiple of 16 in [16, %d]", block_q, MAX_BLOCK_Q);
    return B2_ERR_ARG;
  }
  // P = the B operand (narrow side, planes), Q = the A operand (long side, fp32)
  CUtensorMap tp, tq;
  int rc = b_mn ? make_map3(&tp, B, K, N, ldb, b_plane, BLOCK_K, true) : make_map3(&tp, B, N, K, ldb, b_plane, BLOCK_P, false);
  if (rc) return rc;
  rc = a_mn ? make_map_f32(&tq, A, K, M, lda, block_q, BLOCK_K) : make_map_f32(&tq, A, M, K, lda, BLOCK_K, block_q);
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
