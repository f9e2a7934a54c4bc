// Shared pieces of the fused 784 -> 10 training-step kernels (fused_linear.cu: one step per launch;
// fused_linear_ms.cu: many steps per launch): tile geometry, the shared-memory layout, mbarrier / bulk-copy /
// packed-FMA wrappers.
#pragma once
#include <math.h>

#include "common.cuh"

namespace fl {

constexpr int K = 784, NC = 10;
constexpr int TILE_ROWS = 16, STAGE_ROWS = 8, STAGES_PER_TILE = 2, NSTAGES = 7;
constexpr int XSTRIDE = 784;  // floats per staged row: rows lie back to back (3136 B = 16 mod 32 words: the forward warps'
                              // 64-bit loads of four rows x eight input pairs then touch every bank exactly twice)
constexpr int F_WARPS = 7, G_WARPS = 7;
constexpr int P_WARP = 14, H_WARP = 15;
constexpr int THREADS = 16 * 32;
constexpr int KSLICES = 7, KW = K / KSLICES;    // forward: 7 slices of 112 inputs (28 quads, 7 per lane sub-slice)
constexpr int KGROUPS = K / 8;                  // backward: 98 groups of 8 inputs (quads kg and kg+98)
constexpr int ROW_BYTES = K * 4;                // 3136
constexpr int PART_ROW = 11;                    // floats per row of a partial-logit set (odd: conflict-free)
constexpr int PART_SET = TILE_ROWS * PART_ROW;  // 176
constexpr int PART_BUF = KSLICES * PART_SET;    // 1232
constexpr int DZ_ROW = 16;                      // dz of class c at (c/5)*8 + c%5
constexpr int DZ_BUF = TILE_ROWS * DZ_ROW;
constexpr int NBUF = 3;                        // partial-logit / dz buffers between the pipeline stages
constexpr long SCRATCH_STRIDE = 7856;           // 7840 dW + 10 db + loss + correct, padded to 16 B
constexpr int LOSS_IDX = 7850, CORRECT_IDX = 7851;
// reduction over the per-CTA partials: 7852 outputs, each the sum of four quarter-sums of the grid (CTA order inside a
// quarter, quarters combined 0,1,2,3) -- the same association in every kernel that reduces them
constexpr int RED_OUT = 32, RED_Q = 4, RED_N = 7852;
constexpr int RED_CHUNK = 40;   // partials a reducing thread keeps in flight (a quarter of a 148-CTA grid is 37)

struct __align__(16) Smem {
  float W[K * NC];                             // 31360 B   [784][10], as in global memory
  float x[NSTAGES * STAGE_ROWS * XSTRIDE];     // 175616 B
  float part[NBUF * PART_BUF];                 // 14784 B   part[buf][slice][row][class]
  float ys[NSTAGES * STAGE_ROWS * NC];         // 2240 B
  float dz[NBUF * DZ_BUF];                     // 3072 B
  float red_db[32 * 5];
  float red_loss[16];
  float red_corr[16];
  unsigned long long w_full;                   // producer -> F: weight tile landed
  unsigned long long full[NSTAGES];            // producer -> F, H, G   (tx bytes)
  unsigned long long empty[NSTAGES];           // G -> producer
  unsigned long long part_full[NBUF];             // F -> H
  unsigned long long part_empty[NBUF];            // H -> F
  unsigned long long dz_full[NBUF];               // H -> G
  unsigned long long dz_empty[NBUF];              // G -> H
};
static_assert(sizeof(Smem) <= 232448, "shared memory budget");

typedef unsigned long long u64;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(unsigned long long* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t ok = 0;
  // try_wait suspends the warp in hardware for a bounded time; the loop around it is kept to three
  // instructions so that waiting stages do not take issue slots from the working ones.
  for (uint32_t spins = 0;; ++spins) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(addr), "r"(parity)
        : "memory");
    if (ok) break;
    __nanosleep(128);                   // back off: the issue slots belong to the stages that have work
    if (spins > (1u << 22)) __trap();  // watchdog: a protocol bug must not hang the GPU
  }
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes,
                                         unsigned long long* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// Programmatic dependent launch: the kernel may start while its predecessor on the stream is still running
// (launch latency and prologue overlap); pdl_wait() blocks until the predecessor's writes are visible.  Both
// are no-ops when the kernel was launched without the attribute.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// packed fp32x2 FMA (SASS FFMA2): d.lo = a.lo*b.lo + c.lo, d.hi = a.hi*b.hi + c.hi
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


// ---- F: forward of one 16-row tile for one warp (slice = 112 inputs of the 784) -----------------------------------------
// lane = (input sub-slice ksub 0..7) x (row quad rq 0..3): 4 rows (rq, rq+4, rq+8, rq+12) x 10 classes per thread, input
// pairs 8j + ksub of the slice.  Packed fp32x2 FMAs over class pairs: the W pair comes straight from the [784][10] tile,
// the x value is the scalar-broadcast operand.  Four rows per thread (r01: two) halve the loads of W, which were 41 % of the
// kernel's shared-memory wavefronts (ncu r02f: 980 of 2394 per tile) -- and shared memory is what bounds the kernel.
// The eight partial sums per (row, class) are combined with a reduce-scatter over the sub-slice bits of the lane id
// (20 + 10 + 5 shuffles): every lane ends up with five classes of one row and writes them to the partial-logit set.
__device__ __forceinline__ void forward_tile(Smem& sm, int slice, int lane, int na, int nb, int buf, int i_t) {
  const int rq = lane & 3, ksub = lane >> 2;
  const float* xa = sm.x + ((na % NSTAGES) * STAGE_ROWS + rq) * XSTRIDE + slice * KW + 2 * ksub;   // tile rows rq, rq+4
  const float* xb = sm.x + ((nb % NSTAGES) * STAGE_ROWS + rq) * XSTRIDE + slice * KW + 2 * ksub;   // tile rows rq+8, rq+12
  const float* wp = sm.W + (slice * KW + 2 * ksub) * NC;
  u64 acc[4][5];                                  // [row][class pair]
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int p = 0; p < 5; ++p) acc[r][p] = 0ull;
#pragma unroll
  for (int j = 0; j < KW / 16; ++j) {             // 7 input pairs per lane
    float2 xr[4];
    xr[0] = *reinterpret_cast<const float2*>(xa + 16 * j);
    xr[1] = *reinterpret_cast<const float2*>(xa + 4 * XSTRIDE + 16 * j);
    xr[2] = *reinterpret_cast<const float2*>(xb + 16 * j);
    xr[3] = *reinterpret_cast<const float2*>(xb + 4 * XSTRIDE + 16 * j);
    const ulonglong2* w0 = reinterpret_cast<const ulonglong2*>(wp + j * 16 * NC);   // 2 inputs x 10 classes = 5 x 16 B
    const ulonglong2 w01 = w0[0], w23 = w0[1], w45 = w0[2], w67 = w0[3], w89 = w0[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      // first input of the pair: class pairs (0,1)(2,3)(4,5)(6,7)(8,9) = w01.x w01.y w23.x w23.y w45.x
      const u64 x0 = pack2(xr[r].x, xr[r].x);
      acc[r][0] = ffma2(x0, w01.x, acc[r][0]);
      acc[r][1] = ffma2(x0, w01.y, acc[r][1]);
      acc[r][2] = ffma2(x0, w23.x, acc[r][2]);
      acc[r][3] = ffma2(x0, w23.y, acc[r][3]);
      acc[r][4] = ffma2(x0, w45.x, acc[r][4]);
      // second input: w45.y w67.x w67.y w89.x w89.y
      const u64 x1 = pack2(xr[r].y, xr[r].y);
      acc[r][0] = ffma2(x1, w45.y, acc[r][0]);
      acc[r][1] = ffma2(x1, w67.x, acc[r][1]);
      acc[r][2] = ffma2(x1, w67.y, acc[r][2]);
      acc[r][3] = ffma2(x1, w89.x, acc[r][3]);
      acc[r][4] = ffma2(x1, w89.y, acc[r][4]);
    }
  }
  // reduce-scatter over the eight input sub-slices (lane bits 2, 3, 4): 40 -> 20 -> 10 -> 5 values per lane
  float v[40];
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int p = 0; p < 5; ++p) unpack2(acc[r][p], v[r * 10 + 2 * p], v[r * 10 + 2 * p + 1]);
  const bool b0 = ksub & 1, b1 = ksub & 2, b2 = ksub & 4;
  float v20[20], v10[10], v5[5];
#pragma unroll
  for (int i = 0; i < 20; ++i) {
    const float got = __shfl_xor_sync(0xffffffffu, b0 ? v[i] : v[20 + i], 4);
    v20[i] = (b0 ? v[20 + i] : v[i]) + got;
  }
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    const float got = __shfl_xor_sync(0xffffffffu, b1 ? v20[i] : v20[10 + i], 8);
    v10[i] = (b1 ? v20[10 + i] : v20[i]) + got;
  }
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const float got = __shfl_xor_sync(0xffffffffu, b2 ? v10[i] : v10[5 + i], 16);
    v5[i] = (b2 ? v10[5 + i] : v10[i]) + got;
  }
  // this lane now holds classes 5*b2 .. 5*b2+4 of tile row rq + 4*(2*b0 + b1)
  mbar_wait(&sm.part_empty[buf], ((uint32_t)(i_t / NBUF) & 1u) ^ 1u);   // H is done with this buffer
  float* pa = sm.part + buf * PART_BUF + slice * PART_SET + (rq + 4 * ((b0 ? 2 : 0) + (b1 ? 1 : 0))) * PART_ROW + (b2 ? 5 : 0);
#pragma unroll
  for (int i = 0; i < 5; ++i) pa[i] = v5[i];
  __syncwarp();
  if (lane == 0) mbar_arrive(&sm.part_full[buf]);
}

}  // namespace fl
