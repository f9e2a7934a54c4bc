// Small batches of the 784-wide models, all steps in ONE launch, on ONE thread-block cluster.
//
// The reference's shipped configuration is exactly this: distributed.py trains 784-100-10 with 100 rows per step
// for 27 500 steps (distributed.py:46-48,141-150).  At that size a step is a few microseconds of arithmetic; run as
// one launch sequence per step it costs 45-50 us, nearly all of it launch and grid-level synchronisation latency.
// Here the loop runs inside one kernel on a cluster of 14 CTAs (14 SMs), which synchronise with the hardware cluster
// barrier and exchange their per-step data through a small L2-resident scratch buffer (why not through distributed shared
// memory: see below):
//
//   CTA j owns inputs [56j, 56j+56) of the first layer for good: its slice of W1 lives in its shared memory for the whole
//   call and is updated there (nothing of W1 moves between steps), and it streams only its 56 columns of x.
//   per step:
//     1. partial pre-activations  zpart_j = x[:, slice j] . W1[slice j, :]   [B, N1]   (FFMA2), stored to an L2-resident
//        scratch buffer, one slot per CTA
//        -- cluster barrier --
//     2. CTA j finishes rows [RO*j, RO*j+RO): z = b1 + sum_c zpart_c (CTA order), sigmoid, output layer, softmax / loss /
//        accuracy (16 lanes per row), dz3, dz2 = (dz3 W2^T) a(1-a); its rows of dz2 go to the scratch buffer; partial
//        sums of the small gradients (W2, b2, b1, loss, correct) stay in shared memory
//        -- cluster barrier --
//     3. every CTA copies the whole dz2 [B, N1] into its shared memory (cp.async from L2); meanwhile it adds the partial
//        small gradients of all CTAs (from L2, CTA order) and applies them to its replica of W2, b2, b1
//     4. dW1[slice j] = x[:, slice j]^T . dz2, W1[slice j] -= lr dW1            (FFMA2; local, no exchange)
//   With one layer (784 -> 10, batch <= 256) step 2 is the loss head itself and step 3's small gradients are just the bias.
// The two bulk exchanges go through L2 and not through distributed shared memory: a CTA has to take in ~40 KB in each, and
// DSMEM delivers ~20 B/clk per SM where L2 delivers several times that (measured: 3.5 us -> 0.8 us for the dz2 gather).
// The cluster is what makes the barriers cheap (hardware cluster barrier, ~0.3 us) and the CTAs co-scheduled.
// Several workers (distributed.py:11-13: 1 ps + 2 workers): one cluster per GPU; at the end of a step the rank's gradients
// (every CTA's slice of dW1, the small ones once) are summed over the ranks through the NVLink peer cells of common.cuh
// (exchange_block: push all, collect peer by peer in rank order) and only then applied -- still one launch for all steps.
// Every sum runs in a fixed order (CTA order, row order, ascending k, rank order), so a run is bit-reproducible.
#include <cooperative_groups.h>
#include <math.h>
#include <stdlib.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace cs {

typedef unsigned long long u64;
constexpr int C = 14;              // CTAs per cluster
constexpr int K0 = 784;            // inputs
constexpr int KS = K0 / C;         // 56 inputs per CTA
constexpr int XS = 60;             // floats per staged x row (240 B: 16-byte aligned, rows spread over the banks)
constexpr int THREADS = 512;
constexpr int NC = 10;             // classes

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
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async4(void* dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

struct args {
  const float* x_pool;
  const float* y_pool;
  long span, first_row, global_batch;
  int rows, n_steps;
  int act, loss;
  float lr, inv_B;
  float* params;
  long w1_off, b1_off, w2_off, b2_off;   // flat parameter layout (w2/b2 unused with one layer)
  float* grads;                          // [P + 2]: loss and correct count of the last step at P, P+1
  long P;
  float* loss_trace;
  int* correct_trace;
  float* global_step;
  long rank_off;                         // rank * rows: this rank's slice of every global batch
  int exchange;                          // != 0: several ranks, fx describes the peers (step1 / parity are those of step 0)
  b2_fx_args fx;
  float* scratch;                        // [slots][rows][N1] partial pre-activations, then [rows][N1] dz (L2-resident exchange buffers)
  long long* prof;                       // diagnostics (B2_CS_PROF=1): SM clocks CTA 0 / thread 0 spent per phase, summed over steps
};
constexpr int NPROF = 12;
#define CS_TICK(slot)                                     \
  do {                                                    \
    if (a.prof && j == 0 && t == 0) {                     \
      const long long now_ = clock64();                   \
      prof_acc[slot] += now_ - prof_t;                    \
      prof_t = now_;                                      \
    }                                                     \
  } while (0)

// Shared-memory layout of one CTA (floats), the same in all CTAs of the cluster.
template <int N1, int N2, int BMAX>
struct Layout {
  static constexpr int RO_MAX = (BMAX + C - 1) / C;                 // rows a CTA finishes per step
  static constexpr int NS = N2 > 0 ? N1 * N2 + N2 + N1 + 2 : N1 + 2;   // small gradients: [W2][b2][b1][loss][correct] or [b][loss][correct]
  static constexpr int P4_SLICES = N2 > 0 ? 2 : 16;                 // row slices of the weight-gradient pass
  static constexpr int x = 0;                                       // [2][BMAX][XS]
  static constexpr int w1 = x + 2 * BMAX * XS;                      // [KS][N1]
  static constexpr int dz = w1 + KS * N1;                           // [BMAX][N1]   gradient w.r.t. the first layer's pre-activations
  static constexpr int small = dz + BMAX * N1;                      // replicated small parameters [W2][b2][b1] or [b]
  static constexpr int spart = small + ((NS + 3) & ~3);             // this CTA's partial small gradients [NS]
  static constexpr int ssum = spart + ((NS + 3) & ~3);              // this rank's small gradients, kept until the rank exchange
  static constexpr int a2 = ssum + ((NS + 3) & ~3);                 // [RO_MAX][N1]  activations of my rows
  static constexpr int dzown = a2 + RO_MAX * N1;                    // [RO_MAX][N1]  my rows of dz
  static constexpr int z3 = dzown + RO_MAX * N1;                    // [RO_MAX][16]  logits of my rows
  static constexpr int dz3 = z3 + RO_MAX * 16;                      // [RO_MAX][16]
  static constexpr int ys = dz3 + RO_MAX * 16;                      // [2][RO_MAX][NC] labels of my rows, double buffered
  static constexpr int rowl = ys + 2 * RO_MAX * NC;                 // [RO_MAX] row loss, [RO_MAX] hit
  static constexpr int p4 = (rowl + 2 * RO_MAX + 3) & ~3;           // [P4_SLICES][KS][N1] partial weight gradients (none when
  static constexpr int total = p4 + (P4_SLICES > 2 ? P4_SLICES * KS * N1 : 0);   // the two slices sit on the halves of a warp)
};

// Shared-memory loads as volatile asm: they stay where the source puts them -- all operand loads of a tile step first,
// then the FMAs.  Left to itself the compiler (at 128 registers) sinks each load next to its first use, and every FMA
// group then waits out a shared-memory round trip (ncu: 46 % of the loop's stall samples were short-scoreboard waits).
__device__ __forceinline__ float4 lds_f4(const float* p) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"((unsigned)__cvta_generic_to_shared(p)));
  return v;
}
__device__ __forceinline__ float2 lds_f2(const float* p) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"((unsigned)__cvta_generic_to_shared(p)));
  return v;
}
__device__ __forceinline__ ulonglong2 lds_u64x2(const float* p) {
  ulonglong2 v;
  asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "r"((unsigned)__cvta_generic_to_shared(p)));
  return v;
}
__device__ __forceinline__ u64 lds_u64(const float* p) {
  u64 v;
  asm volatile("ld.shared.b64 %0, [%1];" : "=l"(v) : "r"((unsigned)__cvta_generic_to_shared(p)));
  return v;
}

__device__ __forceinline__ void cp_async8(void* dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}

// ---- step 1: zpart[r][n] = sum_k x[r][k] W1[k][n] over this CTA's 56 inputs, written to this CTA's slot of the exchange
// buffer in global memory.  Thread tile TR rows x TC columns, packed FMAs over column pairs (the W pair is the vector
// operand, the x value the scalar one).  Consecutive lanes take consecutive column groups of one row group: the x loads
// are broadcasts and the W loads contiguous.  The tile is tall (TR = 10 for N1 = 100) because every warp reads whole rows
// of W: with short tiles the pass is bound by shared-memory wavefronts, not by the FMA pipe.
template <int N1, int TR, int TC>
__device__ __forceinline__ void forward_partial(const float* __restrict__ xs, const float* __restrict__ w1, float* __restrict__ zout,
                                                int rows, int t, int kbeg, int kend) {
  constexpr int CG = N1 / TC;
  static_assert(N1 % TC == 0 && TC % 2 == 0, "column tiling");
  const int rg = t / CG, cgi = t - rg * CG;
  const int r0 = rg * TR, n0 = cgi * TC;
  if (r0 >= rows) return;
  u64 acc[TR][TC / 2];
#pragma unroll
  for (int i = 0; i < TR; ++i)
#pragma unroll
    for (int p = 0; p < TC / 2; ++p) acc[i][p] = 0ull;
  // Software pipeline over steps of two inputs, two register buffers: the loads of the next step are issued before the
  // FMAs of the current one, so the shared-memory pipe and the FMA pipe work at the same time inside one warp (with 2-4
  // warps per sub-partition there are not enough other warps to do that overlapping for us).
  struct Ops { float2 x[TR]; u64 w[2][TC / 2]; };
  auto load = [&](Ops& o, int k) {
    k = min(k, kend - 2);                                               // past the end: a valid address, result unused
#pragma unroll
    for (int i = 0; i < TR; ++i) o.x[i] = lds_f2(xs + (r0 + i) * XS + k);   // rows >= `rows` hold zeros
#pragma unroll
    for (int kk = 0; kk < 2; ++kk) {
      if (TC % 4 == 0) {
#pragma unroll
        for (int p = 0; p < TC / 4; ++p) {
          const ulonglong2 ww = lds_u64x2(w1 + (k + kk) * N1 + n0 + 4 * p);
          o.w[kk][2 * p] = ww.x;
          o.w[kk][(2 * p + 1) % (TC / 2)] = ww.y;
        }
      } else {
#pragma unroll
        for (int p = 0; p < TC / 2; ++p) o.w[kk][p] = lds_u64(w1 + (k + kk) * N1 + n0 + 2 * p);
      }
    }
  };
  auto fmas = [&](const Ops& o) {
#pragma unroll
    for (int kk = 0; kk < 2; ++kk)
#pragma unroll
      for (int i = 0; i < TR; ++i) {
        const float xs_ = kk == 0 ? o.x[i].x : o.x[i].y;
        const u64 xx = pack2(xs_, xs_);
#pragma unroll
        for (int p = 0; p < TC / 2; ++p) acc[i][p] = ffma2(xx, o.w[kk][p], acc[i][p]);
      }
  };
  Ops A, Bf;
  load(A, kbeg);
#pragma unroll 1
  for (int k = kbeg; k < kend; k += 4) {
    load(Bf, k + 2);
    fmas(A);
    load(A, k + 4);
    fmas(Bf);
  }
#pragma unroll
  for (int i = 0; i < TR; ++i) {
    if (r0 + i < rows) {
      float* o = zout + (size_t)(r0 + i) * N1 + n0;
      if (TC % 4 == 0) {
#pragma unroll
        for (int p = 0; p < TC / 4; ++p) {
          ulonglong2 v;
          v.x = acc[i][2 * p];
          v.y = acc[i][(2 * p + 1) % (TC / 2)];
          *reinterpret_cast<ulonglong2*>(o + 4 * p) = v;
        }
      } else {
#pragma unroll
        for (int p = 0; p < TC / 2; ++p) *reinterpret_cast<u64*>(o + 2 * p) = acc[i][p];
      }
    }
  }
}

// ---- step 4: dW1[k][n] = sum_r x[r][k] dz[r][n] for this CTA's 56 inputs, then W1 -= lr dW1.  Thread tile TK input
// pairs x TCC columns over one of SLICES row slices (summed in slice order afterwards); the x pair is the vector operand,
// dz the scalar one.  Consecutive lanes: consecutive column groups of one input group (x broadcast, dz contiguous).
template <int N1, int TK, int TCC, int SLICES>
__device__ __forceinline__ void weight_gradient(const float* __restrict__ xs, const float* __restrict__ dz, float* __restrict__ w1,
                                                float* __restrict__ p4, int rows, float lr, int t, float* gout) {
  constexpr int KG = KS / (2 * TK), CG = N1 / TCC, TILES = KG * CG;
  static_assert(KS % (2 * TK) == 0 && N1 % TCC == 0 && TCC % 2 == 0, "tiling");
  // SLICES == 2: the two row slices of a tile sit on the two halves of a warp (lanes l and l ^ 16) and are added with
  // shuffles -- no trip through shared memory.  More slices: one thread group per slice, partials summed through p4.
  constexpr bool LANE_SLICES = SLICES == 2;
  const int lane = t & 31;
  const int tile = LANE_SLICES ? (t >> 5) * 16 + (lane & 15) : t % TILES;
  const int sl = LANE_SLICES ? (lane >> 4) : t / TILES;
  const bool has_tile = tile < TILES && sl < SLICES;
  const int rem = has_tile ? tile : 0;
  const int kg = rem / CG, cgi = rem - kg * CG;
  const int k0 = kg * 2 * TK, n0 = cgi * TCC;
  const int per = (rows + SLICES - 1) / SLICES;
  const int ra = has_tile ? min(rows, sl * per) : 0, rb = has_tile ? min(rows, ra + per) : 0;   // no tile: no rows
  u64 acc[TK][TCC];
#pragma unroll
  for (int a = 0; a < TK; ++a)
#pragma unroll
    for (int c = 0; c < TCC; ++c) acc[a][c] = 0ull;
  // Software pipeline over pairs of rows, two register buffers (see forward_partial).
  constexpr int RB = 2;
  struct Ops { u64 xp[RB][TK]; float d[RB][TCC]; };
  auto load = [&](Ops& o, int r) {
#pragma unroll
    for (int q = 0; q < RB; ++q) {
      const int rr = max(min(r + q, rb - 1), 0);                      // rows past the slice: a valid address, masked in fmas()
      if (TK % 2 == 0) {
#pragma unroll
        for (int a = 0; a < TK / 2; ++a) {
          const ulonglong2 v = lds_u64x2(xs + rr * XS + k0 + 4 * a);
          o.xp[q][2 * a] = v.x;
          o.xp[q][(2 * a + 1) % TK] = v.y;
        }
      } else {
#pragma unroll
        for (int a = 0; a < TK; ++a) o.xp[q][a] = lds_u64(xs + rr * XS + k0 + 2 * a);
      }
      if (TCC % 4 == 0) {
#pragma unroll
        for (int c4 = 0; c4 < TCC / 4; ++c4) {
          const float4 v = lds_f4(dz + rr * N1 + n0 + 4 * c4);
          o.d[q][4 * c4] = v.x; o.d[q][(4 * c4 + 1) % TCC] = v.y; o.d[q][(4 * c4 + 2) % TCC] = v.z; o.d[q][(4 * c4 + 3) % TCC] = v.w;
        }
      } else {
#pragma unroll
        for (int c2 = 0; c2 < TCC / 2; ++c2) {
          const float2 v = lds_f2(dz + rr * N1 + n0 + 2 * c2);
          o.d[q][2 * c2] = v.x; o.d[q][2 * c2 + 1] = v.y;
        }
      }
    }
  };
  auto fmas = [&](const Ops& o, int r) {
#pragma unroll
    for (int q = 0; q < RB; ++q) {
      const bool on = r + q < rb;
#pragma unroll
      for (int c = 0; c < TCC; ++c) {
        const float dv = on ? o.d[q][c] : 0.f;                        // a masked row adds exact zeros
        const u64 dd = pack2(dv, dv);
#pragma unroll
        for (int a = 0; a < TK; ++a) acc[a][c] = ffma2(o.xp[q][a], dd, acc[a][c]);
      }
    }
  };
  Ops A, Bf;
  load(A, ra);
#pragma unroll 1
  for (int r = ra; r < rb; r += 2 * RB) {
    load(Bf, r + RB);
    fmas(A, r);
    load(A, r + 2 * RB);
    fmas(Bf, r + RB);
  }
  if (LANE_SLICES) {
    float g[TK][TCC][2];                                              // slice 0 + slice 1 of my tile's 2*TK x TCC elements
#pragma unroll
    for (int a = 0; a < TK; ++a)
#pragma unroll
      for (int c = 0; c < TCC; ++c) {
        float lo, hi;
        unpack2(acc[a][c], lo, hi);
        g[a][c][0] = lo + __shfl_xor_sync(0xffffffffu, lo, 16);
        g[a][c][1] = hi + __shfl_xor_sync(0xffffffffu, hi, 16);
      }
    if (gout) __syncthreads();                                        // several ranks: every warp is done reading dz, gout may alias it
    if (has_tile && sl == 0) {
#pragma unroll
      for (int a = 0; a < TK; ++a)
#pragma unroll
        for (int c = 0; c < TCC; ++c) {
          const int i0 = (k0 + 2 * a) * N1 + n0 + c;
          if (gout) {                                                 // handed to the rank exchange instead of applied here
            gout[i0] = g[a][c][0];
            gout[i0 + N1] = g[a][c][1];
          } else {
            w1[i0] = fmaf(-lr, g[a][c][0], w1[i0]);
            w1[i0 + N1] = fmaf(-lr, g[a][c][1], w1[i0 + N1]);
          }
        }
    }
  } else {
    if (has_tile) {
#pragma unroll
      for (int a = 0; a < TK; ++a)
#pragma unroll
        for (int c = 0; c < TCC; ++c) {
          float lo, hi;
          unpack2(acc[a][c], lo, hi);
          const int i0 = (k0 + 2 * a) * N1 + n0 + c;
          p4[sl * KS * N1 + i0] = lo;
          p4[sl * KS * N1 + i0 + N1] = hi;
        }
    }
    __syncthreads();
    for (int i = t; i < KS * N1; i += THREADS) {
      float g = 0.f;
#pragma unroll
      for (int s2 = 0; s2 < SLICES; ++s2) g += p4[s2 * KS * N1 + i];    // row slices in order
      if (gout) gout[i] = g;                                         // (gout aliases dz: all reads of dz ended before the barrier above)
      else w1[i] = fmaf(-lr, g, w1[i]);
    }
  }
}

// Several workers: every element of buf[0..n) (this rank's gradient) is replaced by the sum over all ranks, in rank order.
// All `nthr` threads take part, EPT elements each: push them all to the peers' cells, then collect peer by peer, all of a
// peer's cells polled together -- the NVLink trips of the elements overlap.  Cell index of element i = cell0 + i; elements
// [npush_all, n) are pushed by CTAs with `push_tail` set only (values every CTA of the rank holds identically).
template <int EPT>
__device__ __forceinline__ void exchange_block(const b2_fx_args& fx, unsigned step1, int parity, long cell0, long tail_cell0,
                                               float* __restrict__ buf, int n, int npush_all, bool push_tail, int t, int nthr) {
  const size_t slot = ((size_t)parity * B2_MAX_PEERS) * (size_t)fx.stride;
  float mine[EPT], acc[EPT];
#pragma unroll
  for (int e = 0; e < EPT; ++e) {
    const int i = t + e * nthr;
    mine[e] = i < n ? buf[i] : 0.f;
    acc[e] = 0.f;
    if (i < n && (i < npush_all || push_tail)) {
      const long ci = i < npush_all ? cell0 + i : tail_cell0 + (i - npush_all);
      b2_fx_push(fx, step1, parity, ci, mine[e]);
    }
  }
  for (int p = 0; p < fx.nranks; ++p) {
    if (p == fx.rank) {
#pragma unroll
      for (int e = 0; e < EPT; ++e) acc[e] += mine[e];
      continue;
    }
    const unsigned long long* cells = fx.region[fx.rank] + slot + (size_t)p * fx.stride;
    unsigned pending = 0;
#pragma unroll
    for (int e = 0; e < EPT; ++e)
      if (t + e * nthr < n) pending |= 1u << e;
    unsigned spins = 0;
    while (pending) {
      unsigned long long got[EPT];
#pragma unroll
      for (int e = 0; e < EPT; ++e)
        if ((pending >> e) & 1u) {
          const int i = t + e * nthr;
          const long ci = i < npush_all ? cell0 + i : tail_cell0 + (i - npush_all);
          asm volatile("ld.relaxed.sys.global.b64 %0, [%1];" : "=l"(got[e]) : "l"(cells + ci) : "memory");
        }
#pragma unroll
      for (int e = 0; e < EPT; ++e)
        if (((pending >> e) & 1u) && (unsigned)(got[e] >> 32) == step1) {
          acc[e] += __uint_as_float((unsigned)got[e]);
          pending &= ~(1u << e);
        }
      if (++spins > (1u << 26)) __trap();   // a dead peer must not hang this GPU for ever
    }
  }
#pragma unroll
  for (int e = 0; e < EPT; ++e) {
    const int i = t + e * nthr;
    if (i < n) buf[i] = acc[e];
  }
}

template <int N1, int N2, int BMAX>
__global__ void __launch_bounds__(THREADS, 1) cluster_steps_kernel(const args a) {
  using L = Layout<N1, N2, BMAX>;
  constexpr bool TWO = N2 > 0;
  constexpr int NS = L::NS, RO_MAX = L::RO_MAX;
  constexpr int NOUT = TWO ? N2 : N1;                   // classes
  constexpr int V = (N1 % 4 == 0) ? 4 : 2;              // floats per exchange load (rows of N1 floats)
  constexpr int KSPLIT = 1;                             // slots of partial pre-activations per CTA
  constexpr int NSLOT = C * KSPLIT;
  extern __shared__ __align__(16) float sm[];
  cg::cluster_group cluster = cg::this_cluster();
  const int j = (int)cluster.block_rank();
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int B = a.rows;
  const int RO = (B + C - 1) / C;                       // rows per owner CTA
  const int rbeg = min(B, j * RO), rcnt = min(RO, B - rbeg);
  float* xs = sm + L::x;
  float* w1 = sm + L::w1;
  float* dz = sm + L::dz;
  float* small = sm + L::small;
  float* spart = sm + L::spart;
  float* ssum = sm + L::ssum;
  float* a2 = sm + L::a2;
  float* dzown = sm + L::dzown;
  float* z3 = sm + L::z3;
  float* dz3 = sm + L::dz3;
  float* ys = sm + L::ys;
  float* rowl = sm + L::rowl;
  float* p4 = sm + L::p4;
  // replicated small parameters: [W2][b2][b1] (two layers) or [b]
  float* W2s = small;
  float* b2s = small + N1 * N2;
  float* b1s = TWO ? small + N1 * N2 + N2 : small;
  // exchange buffers in global memory (they stay in L2): one slot of partial pre-activations per CTA, then dz
  float* zp_all = a.scratch;
  float* zp_mine = zp_all + (size_t)j * KSPLIT * B * N1;
  float* dzg = zp_all + (size_t)NSLOT * B * N1;
  float* sp_all = dzg + (size_t)B * N1;                  // [C][NS] partial small gradients (two-layer model)

  // ---- prologue: parameters into shared memory, x buffers zeroed (rows past the batch must read as zeros)
  for (int i = t; i < 2 * BMAX * XS; i += THREADS) xs[i] = 0.f;
  for (int i = t; i < KS * N1; i += THREADS) w1[i] = a.params[a.w1_off + (long)(j * KS) * N1 + i];
  for (int i = t; i < N1; i += THREADS) b1s[i] = a.params[a.b1_off + i];
  if (TWO) {
    for (int i = t; i < N1 * N2; i += THREADS) W2s[i] = a.params[a.w2_off + i];
    for (int i = t; i < N2; i += THREADS) b2s[i] = a.params[a.b2_off + i];
  }
  __syncthreads();                                      // the zero fill is complete before any async copy lands
  // the rows of step s into their buffer: issued by `nthr` threads starting at thread `t0` (in the two-layer model the
  // upper half of the CTA requests the next step's rows while the lower half already runs the forward pass)
  auto prefetch = [&](int s, int t0, int nthr) {
    const int tt = t - t0;
    if (tt < 0 || tt >= nthr) return;
    const long start = (a.first_row + (long)s * a.global_batch) % a.span + a.rank_off;
    float* xb = xs + (s & 1) * BMAX * XS;
    const float* src = a.x_pool + start * K0 + j * KS;
    for (int i = tt; i < B * (KS / 4); i += nthr) {
      const int r = i / (KS / 4), q = i - r * (KS / 4);
      cp_async16(xb + r * XS + 4 * q, src + (long)r * K0 + 4 * q);
    }
    float* yb = ys + (s & 1) * RO_MAX * NC;
    const float* ysrc = a.y_pool + (start + rbeg) * NC;
    for (int i = tt; i < rcnt * NC; i += nthr) cp_async4(yb + i, ysrc + i);
    cp_async_commit();
  };
  prefetch(0, 0, THREADS);
  cp_async_wait<0>();
  __syncthreads();
  float gs = 0.f;
  if (j == 0 && t == 0) gs = *a.global_step;
  long long prof_acc[NPROF] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  long long prof_t = clock64();

  for (int s = 0; s < a.n_steps; ++s) {
    // step s's rows are in shared memory (the previous step waited for them); request the next step's now
    if (s + 1 < a.n_steps) prefetch(s + 1, TWO ? THREADS / 2 : 0, TWO ? THREADS / 2 : THREADS);
    CS_TICK(0);
    const float* xb = xs + (s & 1) * BMAX * XS;
    const float* yb = ys + (s & 1) * RO_MAX * NC;
    const unsigned xstep1 = a.fx.step1 + (unsigned)s;               // this step's stamp / cell parity of the rank exchange
    const int xparity = (a.fx.parity + s) & 1;
    // ---- 1. partial pre-activations over my 56 inputs -> my slot of the exchange buffer
    if constexpr (TWO) forward_partial<N1, 10, 4>(xb, w1, zp_mine, B, t, 0, KS);      // warps 0-7
    else forward_partial<N1, 2, N1>(xb, w1, zp_mine, B, t, 0, KS);
    CS_TICK(1);
    cluster.sync();          // release / acquire at cluster scope: the partials are visible to every CTA of the cluster
    CS_TICK(2);
    // ---- 2. my rows: sum the partials of all CTAs in CTA order, finish the forward pass, loss head, backward to dz
    for (int i = t; i < rcnt * (N1 / V); i += THREADS) {
      const int r = i / (N1 / V), n = (i - r * (N1 / V)) * V;
      const float* src = zp_all + (size_t)(rbeg + r) * N1 + n;
      float acc[V];
#pragma unroll
      for (int v = 0; v < V; ++v) acc[v] = 0.f;
      if (V == 4) {
#pragma unroll
        for (int c0 = 0; c0 < NSLOT; c0 += C) {         // C loads in flight at a time, added in slot order
          float4 p[C];
#pragma unroll
          for (int c = 0; c < C; ++c) p[c] = __ldcg(reinterpret_cast<const float4*>(src + (size_t)(c0 + c) * B * N1));
#pragma unroll
          for (int c = 0; c < C; ++c) { acc[0] += p[c].x; acc[1] += p[c].y; acc[2 % V] += p[c].z; acc[3 % V] += p[c].w; }
        }
      } else {
        float2 p[NSLOT];
#pragma unroll
        for (int c = 0; c < NSLOT; ++c) p[c] = __ldcg(reinterpret_cast<const float2*>(src + (size_t)c * B * N1));
#pragma unroll
        for (int c = 0; c < NSLOT; ++c) { acc[0] += p[c].x; acc[1] += p[c].y; }
      }
#pragma unroll
      for (int v = 0; v < V; ++v) {
        const float z = acc[v] + b1s[n + v];
        if (TWO) a2[r * N1 + n + v] = a.act == B2_ACT_SIGMOID ? 1.0f / (1.0f + expf(-z)) : z;
        else z3[r * 16 + n + v] = z;
      }
    }
    __syncthreads();
    CS_TICK(3);
    if constexpr (TWO) {
      // output layer z3 = a2 W2 + b2 (distributed.py:98)
      for (int i = t; i < rcnt * N2; i += THREADS) {
        const int r = i / N2, c = i - r * N2;
        float z = 0.f;
        for (int h = 0; h < N1; ++h) z = fmaf(a2[r * N1 + h], W2s[h * N2 + c], z);
        z3[r * 16 + c] = z + b2s[c];
      }
      __syncthreads();
    }
    CS_TICK(4);
    {
      // loss head (distributed.py:99-102,117-118 / simple.py:20): 16 lanes per row, lane = class; two rows per warp
      const int r = 2 * warp + (lane >> 4), c = lane & 15;
      const bool act_lane = r < rcnt && c < NOUT;
      const float z = act_lane ? z3[r * 16 + c] : -INFINITY;
      const float tt = act_lane ? yb[r * NC + c] : 0.f;
      float l, d, hit = 0.f;
      if (a.loss == B2_LOSS_SSE_SUM) {
        const float e = act_lane ? z - tt : 0.f;
        l = e * e;
        d = 2.0f * e;
      } else {
        float m = z;
#pragma unroll
        for (int o = 8; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        const float e = act_lane ? expf(z - m) : 0.f;
        float ssum = e;
#pragma unroll
        for (int o = 8; o >= 1; o >>= 1) ssum += __shfl_xor_sync(0xffffffffu, ssum, o);
        const float p = e * (1.0f / ssum);
        l = act_lane ? -(tt * logf(p)) : 0.f;          // un-fused log(softmax): 0 * log(0) = NaN, as in the reference
        float sy = tt;
#pragma unroll
        for (int o = 8; o >= 1; o >>= 1) sy += __shfl_xor_sync(0xffffffffu, sy, o);
        // arg max of the prediction and of the label, first index wins ties (tf.argmax)
        float bv = act_lane ? p : -INFINITY, lv = act_lane ? tt : -INFINITY;
        int bi = c, li = c;
#pragma unroll
        for (int o = 8; o >= 1; o >>= 1) {
          const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
          const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
          if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
          const float ov2 = __shfl_xor_sync(0xffffffffu, lv, o);
          const int oi2 = __shfl_xor_sync(0xffffffffu, li, o);
          if (ov2 > lv || (ov2 == lv && oi2 < li)) { lv = ov2; li = oi2; }
        }
        hit = bi == li ? 1.0f : 0.0f;
        d = (p * sy - tt) * a.inv_B;
      }
#pragma unroll
      for (int o = 8; o >= 1; o >>= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
      if (act_lane) {
        dz3[r * 16 + c] = d;
        if (!TWO) {                                     // one layer: this IS the first layer's dz
          dzown[r * N1 + c] = d;
          dzg[(size_t)(rbeg + r) * N1 + c] = d;
        }
      }
      if (r < rcnt && c == 0) { rowl[r] = l; rowl[RO_MAX + r] = hit; }
    }
    __syncthreads();
    CS_TICK(5);
    if constexpr (TWO) {
      // dz2 = (dz3 W2^T) * a2 (1 - a2)   (MatMul grad transpose_b + SigmoidGrad): my rows, also to the exchange buffer
      for (int i = t; i < rcnt * N1; i += THREADS) {
        const int r = i / N1, h = i - r * N1;
        float d = 0.f;
#pragma unroll
        for (int c = 0; c < N2; ++c) d = fmaf(dz3[r * 16 + c], W2s[h * N2 + c], d);
        const float av = a2[r * N1 + h];
        d = a.act == B2_ACT_SIGMOID ? d * av * (1.0f - av) : d;
        dzown[r * N1 + h] = d;
        dzg[(size_t)(rbeg + r) * N1 + h] = d;
      }
      __syncthreads();
    }
    // partial sums of the small gradients over my rows: [W2][b2][b1][loss][correct] or [b][loss][correct]
    for (int i = t; i < NS; i += THREADS) {
      float g = 0.f;
      constexpr int N2D = TWO ? N2 : 1;                 // (keeps the index arithmetic of the two-layer branches legal)
      if (TWO && i < N1 * N2) {
        const int h = i / N2D, c = i - h * N2D;
        for (int r = 0; r < rcnt; ++r) g = fmaf(a2[r * N1 + h], dz3[r * 16 + c], g);
      } else if (TWO && i < N1 * N2 + N2) {
        for (int r = 0; r < rcnt; ++r) g += dz3[r * 16 + (i - N1 * N2)];
      } else if (i < NS - 2) {
        const int h = TWO ? i - N1 * N2 - N2 : i;
        for (int r = 0; r < rcnt; ++r) g += dzown[r * N1 + h];
      } else {
        const float* src = rowl + (i == NS - 2 ? 0 : RO_MAX);
        for (int r = 0; r < rcnt; ++r) g += src[r];
      }
      spart[i] = g;
      if (NS > 64) sp_all[(size_t)j * NS + i] = g;       // many of them: exchanged through L2 like the bulk data
    }
    CS_TICK(6);
    cluster.sync();
    CS_TICK(7);
    // ---- 3. the whole dz into my shared memory (asynchronous copy from L2); meanwhile the small gradients
    if (V == 4) {
      for (int i = t; i < B * N1 / 4; i += THREADS) cp_async16(dz + 4 * i, dzg + 4 * i);
    } else {
      for (int i = t; i < B * N1 / 2; i += THREADS) cp_async8(dz + 2 * i, dzg + 2 * i);
    }
    cp_async_commit();
    float step_loss = 0.f, step_correct = 0.f;
    if (NS <= 64) {
      // few of them (one layer: the bias): every CTA adds all partials itself
      if (t < NS) {
        float g = 0.f;
#pragma unroll
        for (int c = 0; c < C; ++c) g += cluster.map_shared_rank(spart, c)[t];
        if (a.exchange) ssum[t] = g;              // several ranks: summed over the ranks and applied after step 4
        else if (t < NS - 2) b1s[t] = fmaf(-a.lr, g, b1s[t]);
        else if (t == NS - 2) step_loss = g;
        else step_correct = g;
      }
    } else {
      // every CTA adds all C partials of every element itself (C loads from L2 in flight per element, CTA order): no
      // reduce-scatter, no third barrier
      for (int i = t; i < NS; i += THREADS) {
        float p[C];
#pragma unroll
        for (int c = 0; c < C; ++c) p[c] = __ldcg(sp_all + (size_t)c * NS + i);
        float g = 0.f;
#pragma unroll
        for (int c = 0; c < C; ++c) g += p[c];
        if (a.exchange) ssum[i] = g;              // several ranks: summed over the ranks and applied after step 4
        else if (i < NS - 2) small[i] = fmaf(-a.lr, g, small[i]); // [W2][b2][b1] in the order of the gradient vector
        else if (i == NS - 2) step_loss = g;
        else step_correct = g;
      }
    }
    if (j == 0 && !a.exchange) {
      // the thread that holds the step's loss / correct count writes the trace (distributed.py:148-151 fetches them every step)
      const int tl = (NS - 2) % THREADS, tc = (NS - 1) % THREADS;
      if (t == tl) {
        const float l = a.loss == B2_LOSS_SSE_SUM ? step_loss : step_loss * a.inv_B;
        if (a.loss_trace) a.loss_trace[s] = l;
        if (s + 1 == a.n_steps) a.grads[a.P] = l;
      }
      if (t == tc) {
        if (a.correct_trace) a.correct_trace[s] = (int)(step_correct + 0.5f);
        if (s + 1 == a.n_steps) a.grads[a.P + 1] = step_correct;
      }
    }
    if (j == 0 && t == 0) gs += 1.0f;                               // global_step += 1 (distributed.py:108)
    CS_TICK(8);
    cp_async_wait<0>();                                             // dz (and the next step's rows) have landed
    __syncthreads();
    CS_TICK(9);
    // ---- 4. weight gradient of my 56 inputs and the update, all local
    float* gout = a.exchange ? dz : nullptr;                        // several ranks: the local gradient goes to shared memory
    if constexpr (TWO) weight_gradient<N1, 4, 4, L::P4_SLICES>(xb, dz, w1, p4, B, a.lr, t, gout);
    else weight_gradient<N1, 1, N1, L::P4_SLICES>(xb, dz, w1, p4, B, a.lr, t, gout);
    if (a.exchange) {
      // ---- 5. several workers: this rank's gradients (my slice of dW1, then the small ones) summed over the ranks through
      // the NVLink peer cells, then the update.  Cell index: element of W1 in [0, 784*N1), small gradients behind it.
      __syncthreads();
      constexpr int NW = KS * N1;
      for (int i = t; i < NS; i += THREADS) dz[NW + i] = ssum[i];   // one buffer for both
      __syncthreads();
      constexpr int EPT = (NW + NS + THREADS - 1) / THREADS;
      exchange_block<EPT>(a.fx, xstep1, xparity, (long)(j * KS) * N1, (long)K0 * N1, dz, NW + NS, NW, j == 0, t, THREADS);
      __syncthreads();
      for (int i = t; i < NW; i += THREADS) w1[i] = fmaf(-a.lr, dz[i], w1[i]);
      for (int i = t; i < NS; i += THREADS) {
        const float g = dz[NW + i];
        if (i < NS - 2) small[i] = fmaf(-a.lr, g, small[i]);       // [W2][b2][b1] or [b], the order of the gradient vector
        else if (j == 0 && i == NS - 2) {
          const float l = a.loss == B2_LOSS_SSE_SUM ? g : g * a.inv_B;
          if (a.loss_trace) a.loss_trace[s] = l;
          if (s + 1 == a.n_steps) a.grads[a.P] = l;
        } else if (j == 0 && i == NS - 1) {
          if (a.correct_trace) a.correct_trace[s] = (int)(g + 0.5f);
          if (s + 1 == a.n_steps) a.grads[a.P + 1] = g;
        }
      }
    }
    __syncthreads();
    CS_TICK(10);
  }
  if (a.prof && j == 0 && t == 0)
    for (int i = 0; i < NPROF; ++i) a.prof[i] = prof_acc[i];
  // ---- epilogue: parameters back to global memory; peers may still be reading my shared memory in their step 3
  cluster.sync();
  for (int i = t; i < KS * N1; i += THREADS) a.params[a.w1_off + (long)(j * KS) * N1 + i] = w1[i];
  if (j == 0) {
    for (int i = t; i < N1; i += THREADS) a.params[a.b1_off + i] = b1s[i];
    if (TWO) {
      for (int i = t; i < N1 * N2; i += THREADS) a.params[a.w2_off + i] = W2s[i];
      for (int i = t; i < N2; i += THREADS) a.params[a.b2_off + i] = b2s[i];
    }
    if (t == 0) *a.global_step = gs;
  }
}

constexpr int B2_CLUSTER_UNAVAILABLE = -2;

template <int N1, int N2, int BMAX>
static int launch(const args& a, cudaStream_t s) {
  using L = Layout<N1, N2, BMAX>;
  const int smem = L::total * (int)sizeof(float);
  static int prepared[64] = {};                    // 0 = not yet, 1 = usable, -1 = this device cannot co-schedule the cluster
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return B2_CLUSTER_UNAVAILABLE;
  if (!prepared[dev]) {
    B2_CUDA(cudaFuncSetAttribute(cluster_steps_kernel<N1, N2, BMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    B2_CUDA(cudaFuncSetAttribute(cluster_steps_kernel<N1, N2, BMAX>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    // can 14 CTAs of this size be resident in one cluster at all (a MIG slice or a partitioned GPU may not have them)?
    cudaLaunchConfig_t q;
    memset(&q, 0, sizeof(q));
    q.gridDim = dim3(C);
    q.blockDim = dim3(THREADS);
    q.dynamicSmemBytes = smem;
    cudaLaunchAttribute qa[1];
    qa[0].id = cudaLaunchAttributeClusterDimension;
    qa[0].val.clusterDim.x = C;
    qa[0].val.clusterDim.y = 1;
    qa[0].val.clusterDim.z = 1;
    q.attrs = qa;
    q.numAttrs = 1;
    int clusters = 0;
    cudaError_t qe = cudaOccupancyMaxActiveClusters(&clusters, cluster_steps_kernel<N1, N2, BMAX>, &q);
    if (qe != cudaSuccess) (void)cudaGetLastError();
    prepared[dev] = (qe == cudaSuccess && clusters >= 1) ? 1 : -1;
  }
  if (prepared[dev] < 0) return B2_CLUSTER_UNAVAILABLE;      // the caller uses the per-step kernels instead
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(C);
  cfg.blockDim = dim3(THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = C;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  // diagnostics: B2_CS_PROF=1 prints where CTA 0 spends its clocks (synchronises the stream)
  static long long* g_prof = nullptr;
  const char* env = getenv("B2_CS_PROF");
  args b = a;
  if (env && *env == '1') {
    if (!g_prof) B2_CUDA(cudaMalloc((void**)&g_prof, NPROF * sizeof(long long)));
    b.prof = g_prof;
  }
  cudaError_t e = cudaLaunchKernelEx(&cfg, cluster_steps_kernel<N1, N2, BMAX>, b);
  if (e != cudaSuccess) {
    b2_set_error("cluster_steps launch failed: %s", cudaGetErrorString(e));
    (void)cudaGetLastError();
    return B2_ERR_CUDA;
  }
  if (b.prof) {
    long long h[NPROF];
    B2_CUDA(cudaMemcpyAsync(h, g_prof, sizeof(h), cudaMemcpyDeviceToHost, s));
    B2_CUDA(cudaStreamSynchronize(s));
    static const char* names[NPROF] = {"issue next step's loads", "1 forward partial", "fence + cluster barrier 1", "2a sum partials (L2)",
                                       "2b output layer", "2c loss head", "2d dz2 + 2e small partials", "fence + cluster barrier 2",
                                       "3 small grads (dz copy in flight)", "wait dz + sync", "4 weight gradient + apply", ""};
    fprintf(stderr, "B2_CS_PROF N1=%d N2=%d rows=%d steps=%d: CTA 0 thread 0, SM clocks per step\n", N1, N2, a.rows, a.n_steps);
    for (int i = 0; i < 11; ++i) fprintf(stderr, "  %-30s %9.0f\n", names[i], (double)h[i] / a.n_steps);
  }
  return B2_OK;
}

}  // namespace cs

// floats of global scratch the kernel exchanges through for a model whose first layer has n1 outputs: 14 slots of partial
// pre-activations + dz, at the largest batch the kernel takes
long b2_cluster_steps_scratch_floats(int n_layers) {
  return n_layers == 2 ? (long)(cs::C + 1) * 200 * 100 + (long)cs::C * (100 * 10 + 10 + 100 + 2) : (long)(cs::C + 1) * 256 * 10 + 64;
}

// which models / batch sizes the cluster kernel takes: 784-100-10 up to 200 rows per step (distributed.py ships 100; 200 is
// what fits the shared memory with x double-buffered),
// 784 -> 10 up to 128 rows (measured: 5.6 / 6.4 us per step at 16 / 100 rows against 7.7 for the per-step kernels; at 256
// rows the per-step kernels win, 7.8 against 9.1 us, although the kernel itself handles up to 256)
bool b2_cluster_steps_applies(int n_layers, const int* dims, int rows) {
  if (n_layers == 1) return dims[0] == cs::K0 && dims[1] == 10 && rows >= 1 && rows <= 128;
  if (n_layers == 2) return dims[0] == cs::K0 && dims[1] == 100 && dims[2] == 10 && rows >= 1 && rows <= 200;
  return false;
}

int b2_launch_cluster_steps(int n_layers, int act, int loss, const long* w_off, const long* b_off, long P, const float* x_pool,
                            const float* y_pool, long span, long first_row, long global_batch, int rows, int n_steps, float lr,
                            float* params, float* grads, float* loss_trace, int* correct_trace, float* global_step,
                            float* scratch, long rank_off, const b2_fx_args* fx, cudaStream_t s) {
  if (n_steps <= 0) return B2_OK;
  cs::args a;
  memset(&a, 0, sizeof(a));
  a.x_pool = x_pool; a.y_pool = y_pool; a.span = span; a.first_row = first_row; a.global_batch = global_batch;
  a.rows = rows; a.n_steps = n_steps; a.act = act; a.loss = loss; a.lr = lr; a.inv_B = 1.0f / (float)global_batch;
  a.params = params; a.grads = grads; a.P = P; a.loss_trace = loss_trace; a.correct_trace = correct_trace;
  a.global_step = global_step;
  a.scratch = scratch;
  a.rank_off = rank_off;
  if (fx) { a.exchange = 1; a.fx = *fx; }
  a.w1_off = w_off[0]; a.b1_off = b_off[0];
  if (n_layers == 2) { a.w2_off = w_off[1]; a.b2_off = b_off[1]; }
  return n_layers == 2 ? cs::launch<100, 10, 200>(a, s) : cs::launch<10, 0, 256>(a, s);
}
