// bf16 x bf16 -> fp32 GEMM on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM),
// operands staged by TMA into 128B-swizzled shared memory.  This is the layer GEMM of the scaled-up
// MLP (BASELINE.json configs[4], 784-4096-4096-10 at batch 65536), where the contractions are real
// dense ones; the small models stay on the CUDA-core kernels.
//
// One kernel covers the three products of a dense layer (distributed.py:96,98 and their TF1 gradients):
//   forward   C[M,N]  = A[M,K] . W[K,N]          A K-major,  B = W  MN-major   + bias, sigmoid  -> bf16
//   dgrad     dA[M,K] = dZ[M,N] . W[K,N]^T       A K-major,  B = W  K-major    * a(1-a)         -> bf16
//   wgrad     dW[K,N] = A[M,K]^T . dZ[M,N]       A MN-major, B = dZ MN-major                    -> fp32
// so no transposed copy of any operand is ever made: the UMMA descriptors describe the transposition.
//
// Persistent grid (one CTA per SM), 128 x 256 output tiles, BLOCK_K = 64, 4-stage TMA ring.
//   warp 0      TMA producer (one elected lane)
//   warp 1      allocates TMEM (2 x 256 columns: two accumulator buffers) and issues tcgen05.mma
//   warps 2-9   epilogue: tcgen05.ld accumulators -> registers -> fused epilogue -> global, while the
//               next tile's MMAs already fill the other accumulator buffer
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdlib.h>

#include "common.cuh"
#include <mutex>
#include <string.h>

#include "tc_common.cuh"

namespace tg {
using namespace tcx;

constexpr int BLOCK_M = 128, BLOCK_K = 64, UMMA_K = 16;
constexpr int STAGES = 4;
constexpr int THREADS = 320;   // TMA warp, MMA warp, 8 epilogue warps (two per TMEM lane quarter, splitting the columns)
constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;   // 16 KB

template <int BLOCK_N>
struct Cfg {
  static constexpr int B_STAGE_BYTES = BLOCK_N * BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
  static constexpr int TMEM_COLS = 2 * BLOCK_N;          // 512 for N = 256, 256 for N = 128 (power of two)
};

// Shared-memory matrix descriptor, 128B swizzle (cute::UMMA::SmemDescriptor bit layout):
//   [0,14) start>>4   [16,30) LBO>>4   [32,46) SBO>>4   [46,48) version=1   [61,64) layout=2 (SWIZZLE_128B)
// K-major tile  (rows of 64 bf16 = 128 B): 8-row groups 1024 B apart -> SBO = 1024, LBO unused (1)
// MN-major tile (64 MN elements = 128 B per k row): 8-k-row atoms 1024 B apart -> SBO = 1024,
//                                                   next 64 MN elements LBO = 64 k rows * 128 B = 8192 B apart
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, bool mn_major) {
  const uint64_t lbo = mn_major ? (uint64_t)((BLOCK_K * 128) >> 4) : 1ull;
  const uint64_t sbo = 1024 >> 4;
  return (uint64_t)((saddr & 0x3FFFF) >> 4) | (lbo << 16) | (sbo << 32) | (1ull << 46) | (2ull << 61);
}

enum { EPI_BIAS_SIGMOID_BF16 = 0, EPI_DSIGMOID_BF16 = 1, EPI_STORE_F32 = 2, EPI_BIAS_F32 = 3 };

struct Params {
  int M, N, K;
  int a_mn, b_mn;          // 1 = MN-major operand (global [K, M] / [K, N] row-major), 0 = K-major ([M, K] / [N, K])
  int epi;
  const float* bias;       // EPI_BIAS_*: [n_store]
  const __nv_bfloat16* act;  // EPI_DSIGMOID: [M, ldc] activations a; out = acc * a * (1 - a)
  void* out;               // bf16 or fp32 [M, ldc] row-major; only columns < n_store are written
  int ldc, n_store;
  int ksplits;             // > 1: the K range is cut into this many slices, slice z writes out + z*split_stride
  long split_stride;       //      (fixed-order partials; a reduction kernel adds them)
  float* colsum_part;      // EPI_DSIGMOID only, may be NULL: [ceil(M/32)][ldc] column sums of the bf16 OUTPUT over each
                           //      32-row block (the layer's bias gradient, finished by colsum_finish)
};

__device__ __forceinline__ float fast_sigmoid(float z) {
  // sigmoid(z) = 0.5 * tanh(z/2) + 0.5 ; one MUFU op.  The result is rounded to bf16 (8 bits) right after.
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.5f * z));
  return fmaf(0.5f, t, 0.5f);
}

// CL = CTAs per cluster.  The CL CTAs of a cluster work on CL vertically adjacent output tiles (same columns), so
// they need the same B tile: each loads 1/CL of it and TMA-multicasts that share into every CTA's shared memory.
// That cuts the L2 -> SM traffic per k-block from 48 KB to 32 KB (CL = 2) / 24 KB (CL = 4) per CTA, which is what
// bounds the single-CTA kernel (128x256x64 MMAs consume operands faster than L2 can deliver 48 KB per k-block).
template <int BLOCK_N, int CL>
__global__ void __launch_bounds__(THREADS, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, Params p) {
  using C = Cfg<BLOCK_N>;
  static_assert(CL == 1 || (BLOCK_N / 64) % CL == 0, "each CTA multicasts a whole number of 64-column atoms");
  const int cta_rank = CL == 1 ? 0 : (int)cluster_ctarank();
  const int cluster_id = blockIdx.x / CL, num_clusters = gridDim.x / CL;
  constexpr uint16_t kMask = (uint16_t)((1u << CL) - 1u);
  extern __shared__ unsigned char smem_raw[];
  // 1024-byte alignment: required by the 128B swizzle atoms (descriptor base_offset = 0)
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  unsigned char* smA = smem;
  unsigned char* smB = smem + STAGES * A_STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * C::STAGE_BYTES);
  uint64_t* full = bars;                 // [STAGES]  TMA -> MMA
  uint64_t* empty = bars + STAGES;       // [STAGES]  MMA -> TMA
  uint64_t* tfull = bars + 2 * STAGES;   // [2]       MMA -> epilogue
  uint64_t* tempty = tfull + 2;          // [2]       epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_m = (p.M + BLOCK_M - 1) / BLOCK_M, tiles_n = (p.N + BLOCK_N - 1) / BLOCK_N;
  const int tiles_ms = (tiles_m + CL - 1) / CL;            // groups of CL vertically adjacent tiles
  const int mn_tiles = tiles_ms * tiles_n;
  const int ntiles = mn_tiles * p.ksplits;
  const int num_kb_total = (p.K + BLOCK_K - 1) / BLOCK_K;
  const int kb_per = (num_kb_total + p.ksplits - 1) / p.ksplits;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], CL); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull[a], 1); mbar_init(&tempty[a], 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)C::TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();                            // every CTA's barriers exist before anyone multicasts
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------ TMA producer ------------------------------------
    if (lane == 0) {
      int it = 0;
      for (int tile = cluster_id; tile < ntiles; tile += num_clusters) {
        const int mn = tile % mn_tiles, z = tile / mn_tiles;
        const int m0 = ((mn / tiles_n) * CL + cta_rank) * BLOCK_M, n0 = (mn % tiles_n) * BLOCK_N;
        const int kb_begin = z * kb_per, kb_end = min(num_kb_total, kb_begin + kb_per);
        for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
          const int s = it % STAGES;
          mbar_wait(&empty[s], ((it / STAGES) & 1) ^ 1);
          mbar_expect_tx(&full[s], C::STAGE_BYTES);
          const int k0 = kb * BLOCK_K;
          unsigned char* a_dst = smA + s * A_STAGE_BYTES;
          unsigned char* b_dst = smB + s * C::B_STAGE_BYTES;
          if (p.a_mn) {
#pragma unroll
            for (int a = 0; a < BLOCK_M / 64; ++a) tma_load_2d(a_dst + a * (BLOCK_K * 128), &tmA, m0 + 64 * a, k0, &full[s]);
          } else {
            tma_load_2d(a_dst, &tmA, k0, m0, &full[s]);
          }
          if (CL == 1) {
            if (p.b_mn) {
#pragma unroll
              for (int b = 0; b < BLOCK_N / 64; ++b) tma_load_2d(b_dst + b * (BLOCK_K * 128), &tmB, n0 + 64 * b, k0, &full[s]);
            } else {
              tma_load_2d(b_dst, &tmB, k0, n0, &full[s]);
            }
          } else if (p.b_mn) {
            // my share of the B tile: 64-column atoms cta_rank*APC .. , multicast to the whole cluster
            constexpr int APC = (BLOCK_N / 64) / CL;
#pragma unroll
            for (int b = 0; b < APC; ++b) {
              const int atom = cta_rank * APC + b;
              tma_load_2d_mc(b_dst + atom * (BLOCK_K * 128), &tmB, n0 + 64 * atom, k0, &full[s], kMask);
            }
          } else {
            // K-major B: rows cta_rank*(BLOCK_N/CL) .. of the tile (the tensor map's box is BLOCK_N/CL rows high)
            constexpr int RPC = BLOCK_N / CL;
            tma_load_2d_mc(b_dst + cta_rank * RPC * 128, &tmB, k0, n0 + cta_rank * RPC, &full[s], kMask);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------ MMA issuer ------------------------------------
    if (lane == 0) {
      // instruction descriptor (cute::UMMA::InstrDescriptor): D = F32, A = B = BF16, majors, N >> 3, M >> 4
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)p.a_mn << 15) | ((uint32_t)p.b_mn << 16) |
                             ((uint32_t)(BLOCK_N >> 3) << 17) | ((uint32_t)(BLOCK_M >> 4) << 24);
      const uint32_t a_step = p.a_mn ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;   // descriptor units of 16 B
      const uint32_t b_step = p.b_mn ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
      int it = 0, t = 0;
      for (int tile = cluster_id; tile < ntiles; tile += num_clusters, ++t) {
        const int acc = t & 1;
        mbar_wait(&tempty[acc], ((t >> 1) & 1) ^ 1);       // epilogue has drained this accumulator buffer
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BLOCK_N;
        const int z = tile / mn_tiles;
        const int kb_begin = z * kb_per, kb_end = min(num_kb_total, kb_begin + kb_per);
        for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
          const int s = it % STAGES;
          mbar_wait(&full[s], (it / STAGES) & 1);
          tc_fence_after();
          const uint64_t da = make_desc(smem_u32(smA + s * A_STAGE_BYTES), p.a_mn);
          const uint64_t db = make_desc(smem_u32(smB + s * C::B_STAGE_BYTES), p.b_mn);
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
            tc_mma(d_tmem, da + (uint64_t)(k * a_step), db + (uint64_t)(k * b_step), idesc, (kb != kb_begin) || (k != 0));
          if (CL == 1) tc_commit(&empty[s]);                // frees the smem stage when these MMAs retire
          else tc_commit_mc(&empty[s], kMask);              // ... in every CTA: their multicasts write into my B stage too
        }
        tc_commit(&tfull[acc]);                             // accumulator complete
      }
    }
  } else {
    // ------------------------------------ epilogue ------------------------------------
    const int quarter = warp & 3;                           // TMEM lanes 32*quarter .. +31 belong to this warp
    const int chalf = (warp - 2) >> 2;                      // which half of the tile's columns this warp drains
    int t = 0;
    for (int tile = cluster_id; tile < ntiles; tile += num_clusters, ++t) {
      const int acc = t & 1;
      const int mn = tile % mn_tiles, z = tile / mn_tiles;
      const int m0 = ((mn / tiles_n) * CL + cta_rank) * BLOCK_M, n0 = (mn % tiles_n) * BLOCK_N;
      mbar_wait(&tfull[acc], (t >> 1) & 1);
      tc_fence_after();
      const int row = m0 + quarter * 32 + lane;
      const bool row_ok = row < p.M;
      const bool empty_split = z * kb_per >= num_kb_total;   // a K slice past the end contributes zeros
#pragma unroll 1
      for (int c0 = chalf * (BLOCK_N / 2); c0 < (chalf + 1) * (BLOCK_N / 2); c0 += 32) {
        const int n = n0 + c0;
        const size_t off = (size_t)z * p.split_stride + (size_t)row * p.ldc + n;
        const bool vec_ok = (n + 32 <= p.n_store) && ((p.ldc & 7) == 0);
        // the activations this chunk is multiplied with: issue the loads before waiting on tensor memory
        uint4 aq[4];
        if (p.epi == EPI_DSIGMOID_BF16 && vec_ok) {
          const uint4* ap = reinterpret_cast<const uint4*>(p.act + off);
#pragma unroll
          for (int j = 0; j < 4; ++j) aq[j] = row_ok ? ap[j] : make_uint4(0u, 0u, 0u, 0u);
        }
        uint32_t r[32];
        tc_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BLOCK_N + c0), r);
        if (empty_split) {
#pragma unroll
          for (int j = 0; j < 32; ++j) r[j] = 0u;
        }
        if (n >= p.n_store) continue;                       // warp-uniform
        if (p.epi == EPI_STORE_F32 || p.epi == EPI_BIAS_F32) {
          if (!row_ok) continue;
          float* o = reinterpret_cast<float*>(p.out) + off;
          if (vec_ok) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              float4 v = make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]),
                                     __uint_as_float(r[j + 3]));
              if (p.epi == EPI_BIAS_F32) {
                const float4 b = *reinterpret_cast<const float4*>(p.bias + n + j);
                v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w;
              }
              *reinterpret_cast<float4*>(o + j) = v;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)      // fully unrolled: r[] must stay in registers
              if (n + j < p.n_store) o[j] = __uint_as_float(r[j]) + (p.epi == EPI_BIAS_F32 ? p.bias[n + j] : 0.f);
          }
        } else if (!vec_ok) {
          // ragged last column chunk: element-wise (no fused column sums here)
          if (!row_ok) continue;
          __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + off;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            if (n + j < p.n_store) {
              float v = __uint_as_float(r[j]);
              if (p.epi == EPI_BIAS_SIGMOID_BF16) {
                v = fast_sigmoid(v + p.bias[n + j]);
              } else {
                const float a = __bfloat162float(p.act[off + j]);
                v = v * a * (1.0f - a);
              }
              o[j] = __float2bfloat16_rn(v);
            }
          }
        } else {
          float v[32];
          if (p.epi == EPI_BIAS_SIGMOID_BF16) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 b = *reinterpret_cast<const float4*>(p.bias + n + j);
              const float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
              for (int u = 0; u < 4; ++u) {
                v[j + u] = fast_sigmoid(__uint_as_float(r[j + u]) + bb[u]);   // distributed.py:97
              }
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              const uint4 q = aq[j / 8];
              const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
              for (int u = 0; u < 4; ++u) {
                const float a0 = __uint_as_float(w[u] << 16), a1 = __uint_as_float(w[u] & 0xFFFF0000u);
                v[j + 2 * u] = __uint_as_float(r[j + 2 * u]) * a0 * (1.0f - a0);         // SigmoidGrad
                v[j + 2 * u + 1] = __uint_as_float(r[j + 2 * u + 1]) * a1 * (1.0f - a1);
              }
            }
          }
          uint32_t w[16];
#pragma unroll
          for (int u = 0; u < 16; ++u) {
            const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * u], v[2 * u + 1]);
            w[u] = *reinterpret_cast<const uint32_t*>(&h);
          }
          if (row_ok) {
            __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + off;
#pragma unroll
            for (int j = 0; j < 4; ++j)
              *reinterpret_cast<uint4*>(o + 8 * j) = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
          }
          if (p.epi == EPI_DSIGMOID_BF16 && p.colsum_part != nullptr) {
            // BiasAdd gradient fused in: sum the ROUNDED outputs over this warp's 32 rows with a reduce-scatter
            // butterfly (lanes whose bit is set keep the upper half), so lane L ends with column n+L.
            float cs[32];
#pragma unroll
            for (int u = 0; u < 16; ++u) {
              cs[2 * u] = row_ok ? __uint_as_float(w[u] << 16) : 0.f;
              cs[2 * u + 1] = row_ok ? __uint_as_float(w[u] & 0xFFFF0000u) : 0.f;
            }
#define B2_BFLY(HW)                                                                 \
            {                                                                          \
              const bool up = (lane & HW) != 0;                                        \
              _Pragma("unroll") for (int i = 0; i < HW; ++i) {                          \
                const float send = up ? cs[i] : cs[i + HW];                            \
                const float keep = up ? cs[i + HW] : cs[i];                            \
                cs[i] = keep + __shfl_xor_sync(0xffffffffu, send, HW);                 \
              }                                                                        \
            }
            B2_BFLY(16) B2_BFLY(8) B2_BFLY(4) B2_BFLY(2) B2_BFLY(1)
#undef B2_BFLY
            p.colsum_part[(size_t)((m0 + quarter * 32) >> 5) * p.ldc + n + lane] = cs[0];
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
  if (CL > 1) cluster_sync_all();                            // nobody leaves while a peer can still write into it
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)C::TMEM_COLS)
                 : "memory");
  }
}

// ---- host side: tensor maps -----------------------------------------------------------------
static pfn_encode g_encode = nullptr;

static int get_encode() {
  if (!g_encode) g_encode = b2_tensor_map_encoder();
  return g_encode ? B2_OK : B2_ERR_CUDA;
}

// 2-D bf16 row-major [rows, cols]; box = {64 (inner, 128 B), box_rows}; 128B swizzle; out-of-bounds reads give 0
static int make_map(CUtensorMap* m, const void* base, long rows, long cols, int box_rows) {
  const cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t gstr[1] = {(cuuint64_t)cols * 2};
  const cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
  const int r = tcx::b2_tensor_map_cached(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, gdim, gstr, box, CU_TENSOR_MAP_SWIZZLE_128B);
  if (r != (int)CUDA_SUCCESS) {
    b2_set_error("cuTensorMapEncodeTiled failed (%d) for a [%ld, %ld] bf16 matrix", (int)r, rows, cols);
    return B2_ERR_CUDA;
  }
  return B2_OK;
}

}  // namespace tg

namespace tcx {
pfn_encode b2_tensor_map_encoder() {
  static pfn_encode fn_cached = nullptr;
  if (fn_cached) return fn_cached;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
    b2_set_error("cuTensorMapEncodeTiled is not available from this driver");
    return nullptr;
  }
  fn_cached = (pfn_encode)fn;
  return fn_cached;
}

namespace {
struct map_key {
  int type, rank, swizzle, pad;
  const void* base;
  unsigned long long gdim[3], gstr[2];
  unsigned box[3], pad2;
};
struct map_slot {
  bool used;
  map_key key;
  CUtensorMap map;
};
constexpr int MAP_SLOTS = 128;          // direct-mapped; a collision just re-encodes
map_slot g_maps[MAP_SLOTS];
std::mutex g_maps_mu;
}  // namespace

int b2_tensor_map_cached(CUtensorMap* out, CUtensorMapDataType type, unsigned rank, const void* base, const cuuint64_t* gdim,
                         const cuuint64_t* gstr, const cuuint32_t* box, CUtensorMapSwizzle swizzle) {
  pfn_encode enc = b2_tensor_map_encoder();
  if (!enc) return (int)CUDA_ERROR_NOT_SUPPORTED;
  map_key k;
  memset(&k, 0, sizeof(k));
  k.type = (int)type; k.rank = (int)rank; k.swizzle = (int)swizzle; k.base = base;
  for (unsigned i = 0; i < rank; ++i) { k.gdim[i] = gdim[i]; k.box[i] = box[i]; }
  for (unsigned i = 0; i + 1 < rank; ++i) k.gstr[i] = gstr[i];
  unsigned long long h = 1469598103934665603ull;      // FNV-1a over the key
  const unsigned char* kb = reinterpret_cast<const unsigned char*>(&k);
  for (size_t i = 0; i < sizeof(k); ++i) h = (h ^ kb[i]) * 1099511628211ull;
  map_slot& slot = g_maps[h % MAP_SLOTS];
  std::lock_guard<std::mutex> lock(g_maps_mu);
  if (slot.used && memcmp(&slot.key, &k, sizeof(k)) == 0) {
    *out = slot.map;
    return (int)CUDA_SUCCESS;
  }
  const cuuint32_t estr[3] = {1u, 1u, 1u};
  CUresult r = enc(out, type, rank, const_cast<void*>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r == CUDA_SUCCESS) {
    slot.used = true;
    slot.key = k;
    slot.map = *out;
  }
  return (int)r;
}
}  // namespace tcx

template <int BN, int CL>
static int launch_bn(const CUtensorMap& ta, const CUtensorMap& tb, const tg::Params& p, int sm_count, cudaStream_t s) {
  using namespace tg;
  // the shared-memory opt-in is per device: remember which devices of this process have it
  static unsigned long long attr_done = 0ull;
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  if (dev >= 64 || !((attr_done >> dev) & 1ull)) {
    B2_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<BN, CL>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::SMEM_BYTES));
    if (dev < 64) attr_done |= 1ull << dev;
  }
  const int tiles_ms = ((p.M + BLOCK_M - 1) / BLOCK_M + CL - 1) / CL;
  const int groups = tiles_ms * ((p.N + BN - 1) / BN) * p.ksplits;
  int clusters = sm_count / CL;
  if (clusters > groups) clusters = groups;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(clusters * CL);
  cfg.blockDim = dim3(THREADS);
  cfg.dynamicSmemBytes = Cfg<BN>::SMEM_BYTES;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_bf16_kernel<BN, CL>, ta, tb, p);
  if (e != cudaSuccess) {
    b2_set_error("gemm_bf16 launch failed: %s", cudaGetErrorString(e));
    return B2_ERR_CUDA;
  }
  return B2_OK;
}

// A: K-major -> global [M, K]; MN-major -> global [K, M].   B: K-major -> [N, K]; MN-major -> [K, N].
// out: [M, ldc], columns < n_store written.  ksplits > 1 cuts K into slices; slice z writes out + z*split_stride.
int b2_launch_gemm_bf16_ex(const void* A, int a_mn, const void* B, int b_mn, int M, int N, int K, int epi,
                           const float* bias, const void* act, void* out, int ldc, int n_store, int ksplits,
                           long split_stride, float* colsum_part, int sm_count, cudaStream_t s, int local_operands) {
  using namespace tg;
  if ((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B)) & 15) {
    b2_set_error("gemm_bf16: operands must be 16-byte aligned");
    return B2_ERR_ARG;
  }
  if ((a_mn ? M : K) % 8 || (b_mn ? N : K) % 8) {
    b2_set_error("gemm_bf16: operand row pitches must be multiples of 8 elements (M=%d N=%d K=%d)", M, N, K);
    return B2_ERR_ARG;
  }
  int rc = get_encode();
  if (rc) return rc;
  const int BN = (N <= 64) ? 64 : 256;
  // cluster size: 2 CTAs share each B tile by multicast when there are at least two rows of tiles to pair up
  const char* env = getenv("B2_GEMM_CLUSTER");
  int CL = (BN == 256 && M > BLOCK_M) ? 2 : 1;
  if (env && *env && BN == 256) CL = atoi(env) == 4 ? 4 : (atoi(env) == 1 ? 1 : 2);
  if (!local_operands) CL = 1;   // operands of unknown origin: no multicast loads (see common.cuh)
  CUtensorMap ta, tb;
  rc = a_mn ? make_map(&ta, A, K, M, BLOCK_K) : make_map(&ta, A, M, K, BLOCK_M);
  if (rc) return rc;
  rc = b_mn ? make_map(&tb, B, K, N, BLOCK_K) : make_map(&tb, B, N, K, BN / CL);
  if (rc) return rc;
  Params p;
  p.M = M; p.N = N; p.K = K; p.a_mn = a_mn; p.b_mn = b_mn; p.epi = epi;
  p.bias = bias; p.act = reinterpret_cast<const __nv_bfloat16*>(act); p.out = out;
  p.ldc = ldc; p.n_store = n_store;
  const int num_kb = (K + BLOCK_K - 1) / BLOCK_K;
  p.ksplits = ksplits < 1 ? 1 : (ksplits > num_kb ? num_kb : ksplits);
  p.split_stride = split_stride;
  p.colsum_part = colsum_part;
  if (BN == 64) return launch_bn<64, 1>(ta, tb, p, sm_count, s);
  if (CL == 4) return launch_bn<256, 4>(ta, tb, p, sm_count, s);
  if (CL == 2) return launch_bn<256, 2>(ta, tb, p, sm_count, s);
  return launch_bn<256, 1>(ta, tb, p, sm_count, s);
}

int b2_launch_gemm_bf16(const void* A, int a_mn, const void* B, int b_mn, int M, int N, int K, int epi,
                        const float* bias, const void* act, void* out, int sm_count, cudaStream_t s, int local_operands) {
  if (N % 8) {
    b2_set_error("gemm_bf16: N must be a multiple of 8 (N=%d)", N);
    return B2_ERR_ARG;
  }
  return b2_launch_gemm_bf16_ex(A, a_mn, B, b_mn, M, N, K, epi, bias, act, out, N, N, 1, 0, nullptr, sm_count, s,
                                local_operands);
}
