// Shared pieces of the fused 784 -> 10 training-step kernels (fused_linear.cu: one step per launch;
// fused_linear_ms.cu: many steps per launch): tile geometry, the shared-memory layout, mbarrier / bulk-copy /
// packed-FMA wrappers.
#pragma once
#include <math.h>

#include "common.cuh"

namespace fl {

constexpr int K = 784, NC = 10;
constexpr int TILE_ROWS = 16, STAGE_ROWS = 8, STAGES_PER_TILE = 2, NSTAGES = 7;
constexpr int XSTRIDE = 788;  // floats per staged row: 3152 B, keeps 16 B alignment, conflict-free LDS.128
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
  float x[NSTAGES * STAGE_ROWS * XSTRIDE];     // 176512 B
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
// the same with an L2 eviction policy (createpolicy): the x rows are read once per step and should not push the
// per-CTA partials and the weights, which the reduction kernel re-reads a few microseconds later, out of L2
__device__ __forceinline__ unsigned long long l2_evict_first_policy() {
  unsigned long long pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void bulk_g2s_hint(void* dst, const void* src, uint32_t bytes, unsigned long long* bar,
                                              unsigned long long policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
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

}  // namespace fl
