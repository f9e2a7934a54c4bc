// Many training steps of the "784 -> 10" model in ONE launch (simple.py:34-35 / distributed.py:145-150: the
// reference's hot loop is a Python `for` around sess.run; here the loop lives inside the kernel).
//
// Same pipeline as fused_linear.cu (P producer, F forward, H loss head, G backward over 16-row tiles), plus the
// step boundary inside the kernel instead of a second kernel and a new launch:
//   1. every CTA writes its partial dW / db / loss / correct to scratch[cta]          (G + H warps)
//   2. grid barrier a (all CTAs are co-resident: grid <= SM count, one CTA per SM)
//   3. every CTA sums ITS slice of the 7852 outputs over all CTAs -- the same association as
//      fused_linear_reduce_kernel (four quarters of the grid, each in CTA order, combined 0,1,2,3), so both
//      paths give bit-identical results --, exchanges it with the other ranks (b2_fx_exchange) when there
//      are several, applies var -= lr * grad and writes the step's loss / correct-count trace entries
//   4. grid barrier b, every CTA reloads W (31 KB, from L2) and b
// The producer warp does not take part in any of this: it keeps streaming the NEXT step's rows into the ring
// (x and y do not depend on the parameters), so the data pipeline never drains between steps.
#include <stdlib.h>

#include "fused_linear.cuh"

namespace fl {

struct ms_args {
  const float* x_pool;
  const float* y_pool;
  long span;          // pool rows in use (a multiple of global_batch): batches wrap inside it
  long first_row;     // global row of step 0's batch
  long global_batch;
  long rank_off;      // rank * local rows
  int rows;           // local rows per step
  int n_steps;
  float* params;      // W [784][10] at 0, b [10] at 7840
  float* grads;       // [P + 2] reduced gradients of the last step, loss, correct
  float* scratch;     // [grid][SCRATCH_STRIDE]
  unsigned* bar;      // [2] grid-barrier counters, zero when the kernel starts
  float lr, inv_B;
  int loss_kind;
  float* loss_trace;  // [n_steps] or NULL
  int* correct_trace; // [n_steps] or NULL
  float* global_step;
  long P;
  int exchange;       // != 0: fx describes the peers (step1 / parity are those of step 0)
  unsigned long long* stamps;   // diagnostics (B2_MS_STAMPS=1): %globaltimer of CTA 0's phases in steps stamp_step, +1
  int stamp_step;
  int debug;          // diagnostics (B2_MS_DEBUG bit mask): leave parts of the step boundary out to see what they cost
  b2_fx_args fx;
};

// ---- diagnostics: where the time of a step boundary goes (CTA 0 only, two consecutive steps) ----------------------------
enum { ST_G_FIRST = 0, ST_G_LAST, ST_G_STAGED, ST_G_WRITTEN, ST_G_BAR_A, ST_SB_A, ST_SB_REDUCED, ST_SB_C0, ST_SB_BAR_B, ST_SB_C,
       ST_SB_D, ST_F_FIRST_IN, ST_F_FIRST_DONE, ST_F_LAST_DONE, ST_H_FIRST_DONE, ST_H_LAST_DONE, ST_POINTS };
__device__ __forceinline__ void stamp(const ms_args& a, int s, int point) {
  if (a.stamps && blockIdx.x == 0 && (s == a.stamp_step || s == a.stamp_step + 1)) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    a.stamps[(s - a.stamp_step) * ST_POINTS + point] = t;
  }
}

constexpr int NB_PART = 3, NB_ALL = 4;           // named barriers: G+H (256 threads), F+G+H (480 threads)
constexpr int PART_THREADS = (G_WARPS + 1) * 32;  // 256
constexpr int ALL_THREADS = THREADS - 32;         // 480: everybody but the producer warp

__device__ __forceinline__ void named_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// One thread per CTA: publish this CTA's arrival (everything the CTA wrote before the preceding CTA-level barrier
// is ordered before it by the fence) and wait until all CTAs of the grid have arrived `target` times in total.
__device__ __forceinline__ void grid_arrive_wait(unsigned* counter, unsigned target) {
  __threadfence();
  atomicAdd(counter, 1u);
  unsigned spins = 0;
  while (ld_acquire_gpu(counter) < target) {
    __nanosleep(32);
    if (++spins > (1u << 26)) __trap();   // a CTA that never became resident must not hang the GPU for ever
  }
  __threadfence();
}

// G and H warps (pt = 0..255): this CTA's partial -- dW staged in sm.W by the G warps, db / loss / correct summed in a
// fixed order over the head warp's lanes -- goes to scratch[cta], coalesced.
__device__ __forceinline__ void write_partial(const ms_args& a, const Smem& sm, int pt) {
  float* out = a.scratch + (size_t)blockIdx.x * SCRATCH_STRIDE;
  float4* dst = reinterpret_cast<float4*>(out);
  const float4* src = reinterpret_cast<const float4*>(sm.W);
  for (int i = pt; i < K * NC / 4; i += PART_THREADS) dst[i] = src[i];
  if (pt < NC) {
    const int half = pt / 5, cc = pt - 5 * half;
    float v = 0.f;
    for (int r = 0; r < 16; ++r) v += sm.red_db[(half * 16 + r) * 5 + cc];
    out[K * NC + pt] = v;
  } else if (pt == 32) {
    float v = 0.f;
    for (int r = 0; r < 16; ++r) v += sm.red_loss[r];
    out[LOSS_IDX] = a.loss_kind == B2_LOSS_SSE_SUM ? v : v * a.inv_B;
  } else if (pt == 64) {
    float v = 0.f;
    for (int r = 0; r < 16; ++r) v += sm.red_corr[r];
    out[CORRECT_IDX] = v;
  }
}

// Applies the reduced (and, with several ranks, exchanged) value of output i: gradient descent on the parameters,
// the step's trace entries, global_step (distributed.py:107-108).  Called by the one thread that owns output i.
__device__ __forceinline__ void apply_output(const ms_args& a, int s, int i, float g) {
  const float sum = a.exchange ? b2_fx_exchange_at(a.fx, a.fx.step1 + (unsigned)s, (a.fx.parity + s) & 1, i, g) : g;
  if (i < K * NC + NC) {
    a.grads[i] = sum;
    a.params[i] = fmaf(-a.lr, sum, __ldcg(a.params + i));
  } else if (i == LOSS_IDX) {
    a.grads[a.P] = sum;
    if (a.loss_trace) a.loss_trace[s] = sum;
  } else {
    a.grads[a.P + 1] = sum;
    if (a.correct_trace) a.correct_trace[s] = (int)(sum + 0.5f);
    *a.global_step += 1.0f;
  }
}

constexpr int RED_INFLIGHT = 20;   // partial values a thread requests from L2 before it starts adding (one round trip)

// The part of the step boundary that F, G and H warps run together (480 threads, u = 0..479), after this CTA's partial
// is in scratch and grid barrier a has been passed by the CTA's elected thread (which arrives at the first named
// barrier below only then).  Returns with the updated W in shared memory (unless this was the last step).
//
// Every output is summed in the association of fused_linear_reduce_kernel: four quarter-sums over the CTAs
// [q*qper, (q+1)*qper) in CTA order, combined ((s0+s1)+s2)+s3.  What differs is only who loads what, so that a thread
// waits for ONE L2 round trip: all the values it needs are requested first, then added in order.
__device__ __forceinline__ void step_boundary(const ms_args& a, Smem& sm, int s, int u) {
  const int grid = gridDim.x;
  named_sync(NB_ALL, ALL_THREADS);                                    // [A] every CTA's partial is visible
  if (u == 0) stamp(a, s, ST_SB_A);
  const int per = (RED_N + grid - 1) / grid;                          // outputs this CTA owns
  const int e0 = blockIdx.x * per;
  const int qper = (grid + RED_Q - 1) / RED_Q;
  // Two ways to spread the loads, chosen by the grid size (uniform over the CTA); one place where the result is applied.
  //   large grids : output = 4 consecutive lanes, lane q adds quarter q (<= 37 CTAs, two chunks of loads), the
  //                 quarters are combined by shuffles; 120 outputs per round
  //   <= 20 CTAs  : one thread per output adds all four quarters itself (<= 20 loads in flight); 480 outputs per round
  // Every load is issued unconditionally from a clamped (always valid) address, so that they go out back to back.
  const bool lanes4 = grid > RED_INFLIGHT;
  const int q = lanes4 ? (u & 3) : 0;
  const int c0 = lanes4 ? q * qper : 0, c1 = lanes4 ? min(grid, c0 + qper) : grid;
  const int clast = max(c1 - 1, 0);
  const int stride = lanes4 ? ALL_THREADS / RED_Q : ALL_THREADS;
  for (int base = 0; base < per; base += stride) {
    const int el = base + (lanes4 ? (u >> 2) : u);
    const int i = e0 + el;
    const bool valid = el < per && i < RED_N;
    const float* p = a.scratch + (valid ? i : 0);
    float tot = 0.f, acc = 0.f;
    for (int k0 = c0; k0 < c1; k0 += RED_INFLIGHT) {
      float v[RED_INFLIGHT];
#pragma unroll
      for (int k = 0; k < RED_INFLIGHT; ++k)
        v[k] = (a.debug & 1) ? 0.f : __ldcg(p + (size_t)min(k0 + k, clast) * SCRATCH_STRIDE);
#pragma unroll
      for (int k = 0; k < RED_INFLIGHT; ++k) {
        const int c = k0 + k;
        if (c < c1) {
          acc += v[k];
          if (!lanes4 && ((c + 1) % qper == 0 || c + 1 == grid)) {    // last CTA of its quarter: fold the quarter in
            tot = (c < qper) ? acc : tot + acc;
            acc = 0.f;
          }
        }
      }
    }
    if (lanes4) {
      const float s1 = __shfl_down_sync(0xffffffffu, acc, 1);
      const float s2 = __shfl_down_sync(0xffffffffu, acc, 2);
      const float s3 = __shfl_down_sync(0xffffffffu, acc, 3);
      tot = ((acc + s1) + s2) + s3;
    } else {
      for (int qq = (grid + qper - 1) / qper; qq < RED_Q; ++qq) tot += 0.f;   // empty quarters, as the reduction kernel adds them
    }
    if (valid && q == 0 && !(a.debug & 2)) apply_output(a, s, i, tot);
  }
  if (u == 0) stamp(a, s, ST_SB_REDUCED);
  if (s + 1 == a.n_steps) return;                                     // nothing reads the parameters after the last step
  named_sync(NB_ALL, ALL_THREADS);                                    // this CTA's slice is applied (CTA-scope order) ...
  if (u == 0) {
    stamp(a, s, ST_SB_C0);
    if (!(a.debug & 8)) grid_arrive_wait(a.bar + 1, (unsigned)(s + 1) * (unsigned)grid);  // ... and published by the fence in here
    stamp(a, s, ST_SB_BAR_B);
  }
  named_sync(NB_ALL, ALL_THREADS);                                    // [C] every CTA's slice is applied
  if (u == 0) stamp(a, s, ST_SB_C);
  if (!(a.debug & 4))
    for (int i = u; i < K * NC / 4; i += ALL_THREADS)
      reinterpret_cast<float4*>(sm.W)[i] = __ldcg(reinterpret_cast<const float4*>(a.params) + i);
  named_sync(NB_ALL, ALL_THREADS);                                    // [D] W is in shared memory
  if (u == 0) stamp(a, s, ST_SB_D);
}

__global__ void __launch_bounds__(THREADS, 1)
fused_linear_multistep_kernel(const ms_args a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int rows = a.rows;
  const int ntiles = (rows + TILE_ROWS - 1) / TILE_ROWS;
  const float* __restrict__ Wg = a.params;
  const float* __restrict__ bg = a.params + K * NC;
  const int u = tid < P_WARP * 32 ? tid : tid - 32;      // 0..479 over F, G, H (the producer warp is skipped)

  if (warp == P_WARP) {
    // =============================== P: producer ===============================
    if (lane < NSTAGES) mbar_init(&sm.full[lane], 1);
    else if (lane < 2 * NSTAGES) mbar_init(&sm.empty[lane - NSTAGES], G_WARPS);
    else if (lane < 2 * NSTAGES + NBUF) mbar_init(&sm.part_full[lane - 2 * NSTAGES], F_WARPS);
    else if (lane < 2 * NSTAGES + 2 * NBUF) mbar_init(&sm.part_empty[lane - 2 * NSTAGES - NBUF], 1);
    else if (lane < 2 * NSTAGES + 3 * NBUF) mbar_init(&sm.dz_full[lane - 2 * NSTAGES - 2 * NBUF], 1);
    else if (lane < 2 * NSTAGES + 4 * NBUF) mbar_init(&sm.dz_empty[lane - 2 * NSTAGES - 3 * NBUF], G_WARPS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();
    asm volatile("bar.arrive 2, %0;" ::"n"(THREADS) : "memory");  // barriers are live
    const bool y_aligned = (reinterpret_cast<uintptr_t>(a.y_pool) & 15u) == 0;
    int n = 0;
    for (int s = 0; s < a.n_steps; ++s) {
      const long start = (a.first_row + (long)s * a.global_batch) % a.span + a.rank_off;
      const float* __restrict__ x = a.x_pool + start * K;
      const float* __restrict__ y_ = a.y_pool + start * NC;
      const bool y_al = y_aligned && ((start * NC * 4) & 15) == 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        for (int q = 0; q < STAGES_PER_TILE; ++q, ++n) {
          const int slot = n % NSTAGES;
          const uint32_t phase = (uint32_t)(n / NSTAGES) & 1u;
          mbar_wait(&sm.empty[slot], phase ^ 1u);
          const long row0 = (long)tile * TILE_ROWS + q * STAGE_ROWS;
          long left = rows - row0;
          const int nrows = left <= 0 ? 0 : (left < STAGE_ROWS ? (int)left : STAGE_ROWS);
          // 40*nrows bytes from a 16-byte aligned address: row0 is a multiple of 8, so 40*row0 is one of 16
          const bool y_bulk = y_al && ((nrows & 1) == 0) && nrows > 0;
          if (lane == 0) {
            for (int rr = 0; rr < nrows; ++rr)
              bulk_g2s(sm.x + (slot * STAGE_ROWS + rr) * XSTRIDE, x + (row0 + rr) * K, ROW_BYTES, &sm.full[slot]);
            if (y_bulk)
              bulk_g2s(sm.ys + slot * STAGE_ROWS * NC, y_ + row0 * NC, (uint32_t)(nrows * NC * 4), &sm.full[slot]);
          }
          if (!y_bulk) {
            for (int i = lane; i < STAGE_ROWS * NC; i += 32) {
              const int rr = i / NC;
              sm.ys[slot * STAGE_ROWS * NC + i] = (rr < nrows) ? y_[row0 * NC + i] : 0.f;
            }
          }
          __syncwarp();
          if (lane == 0)
            mbar_arrive_expect_tx(&sm.full[slot], (uint32_t)(nrows * ROW_BYTES + (y_bulk ? nrows * NC * 4 : 0)));
        }
      }
    }
    return;
  }

  asm volatile("bar.sync 2, %0;" ::"n"(THREADS) : "memory");
  // step 0's parameters: plain loads by everybody but the producer (L2 -> shared memory)
  for (int i = u; i < K * NC / 4; i += ALL_THREADS)
    reinterpret_cast<float4*>(sm.W)[i] = __ldcg(reinterpret_cast<const float4*>(Wg) + i);
  named_sync(NB_ALL, ALL_THREADS);

  // One loop over the steps for every role, so that the step boundary (a lot of code) exists once in the kernel.
  int i_t = 0;                                  // tiles this CTA has processed so far: ring / buffer phases run on
  for (int s = 0; s < a.n_steps; ++s) {
  if (warp < F_WARPS) {
    // =============================== F: forward ===============================
    const int slice = warp;                       // inputs 112*slice ..
    const int ksub = lane >> 3, rp = lane & 7;    // quads 4i+ksub; rows rp and rp+8
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++i_t) {
        const int na = 2 * i_t, nb = na + 1, buf = i_t % NBUF;
        mbar_wait(&sm.full[na % NSTAGES], (uint32_t)(na / NSTAGES) & 1u);
        mbar_wait(&sm.full[nb % NSTAGES], (uint32_t)(nb / NSTAGES) & 1u);
        if (tid == 0 && tile == blockIdx.x) stamp(a, s, ST_F_FIRST_IN);
        const float* xa = sm.x + ((na % NSTAGES) * STAGE_ROWS + rp) * XSTRIDE + slice * KW + 4 * ksub;
        const float* xb = sm.x + ((nb % NSTAGES) * STAGE_ROWS + rp) * XSTRIDE + slice * KW + 4 * ksub;
        const float* wp = sm.W + (slice * KW + 4 * ksub) * NC;
        u64 acc_a[5], acc_b[5];                       // class pairs (2p, 2p+1) of rows a and b
#pragma unroll
        for (int p = 0; p < 5; ++p) { acc_a[p] = 0ull; acc_b[p] = 0ull; }
#pragma unroll
        for (int i = 0; i < 7; ++i) {
          const float4 va = *reinterpret_cast<const float4*>(xa + 16 * i);
          const float4 vb = *reinterpret_cast<const float4*>(xb + 16 * i);
          const ulonglong2* w0 = reinterpret_cast<const ulonglong2*>(wp + i * 16 * NC);   // 4 inputs x 10 classes = 10 x 16 B
          const float xsa[4] = {va.x, va.y, va.z, va.w};
          const float xsb[4] = {vb.x, vb.y, vb.z, vb.w};
#pragma unroll
          for (int kp = 0; kp < 2; ++kp) {            // inputs 2kp and 2kp+1 of the quad: 20 floats = 5 x 16 B
            const ulonglong2 w01 = w0[5 * kp + 0], w23 = w0[5 * kp + 1], w45 = w0[5 * kp + 2],
                             w67 = w0[5 * kp + 3], w89 = w0[5 * kp + 4];
            const u64 xa0 = pack2(xsa[2 * kp], xsa[2 * kp]), xb0 = pack2(xsb[2 * kp], xsb[2 * kp]);
            acc_a[0] = ffma2(xa0, w01.x, acc_a[0]); acc_b[0] = ffma2(xb0, w01.x, acc_b[0]);
            acc_a[1] = ffma2(xa0, w01.y, acc_a[1]); acc_b[1] = ffma2(xb0, w01.y, acc_b[1]);
            acc_a[2] = ffma2(xa0, w23.x, acc_a[2]); acc_b[2] = ffma2(xb0, w23.x, acc_b[2]);
            acc_a[3] = ffma2(xa0, w23.y, acc_a[3]); acc_b[3] = ffma2(xb0, w23.y, acc_b[3]);
            acc_a[4] = ffma2(xa0, w45.x, acc_a[4]); acc_b[4] = ffma2(xb0, w45.x, acc_b[4]);
            const u64 xa1 = pack2(xsa[2 * kp + 1], xsa[2 * kp + 1]), xb1 = pack2(xsb[2 * kp + 1], xsb[2 * kp + 1]);
            acc_a[0] = ffma2(xa1, w45.y, acc_a[0]); acc_b[0] = ffma2(xb1, w45.y, acc_b[0]);
            acc_a[1] = ffma2(xa1, w67.x, acc_a[1]); acc_b[1] = ffma2(xb1, w67.x, acc_b[1]);
            acc_a[2] = ffma2(xa1, w67.y, acc_a[2]); acc_b[2] = ffma2(xb1, w67.y, acc_b[2]);
            acc_a[3] = ffma2(xa1, w89.x, acc_a[3]); acc_b[3] = ffma2(xb1, w89.x, acc_b[3]);
            acc_a[4] = ffma2(xa1, w89.y, acc_a[4]); acc_b[4] = ffma2(xb1, w89.y, acc_b[4]);
          }
        }
        // the four input sub-slices (lanes xor 8, 16)
        float fa[NC], fb[NC];
#pragma unroll
        for (int p = 0; p < 5; ++p) {
          unpack2(acc_a[p], fa[2 * p], fa[2 * p + 1]);
          unpack2(acc_b[p], fb[2 * p], fb[2 * p + 1]);
        }
#pragma unroll
        for (int c = 0; c < NC; ++c) {
          fa[c] += __shfl_xor_sync(0xffffffffu, fa[c], 8);
          fb[c] += __shfl_xor_sync(0xffffffffu, fb[c], 8);
          fa[c] += __shfl_xor_sync(0xffffffffu, fa[c], 16);
          fb[c] += __shfl_xor_sync(0xffffffffu, fb[c], 16);
        }
        mbar_wait(&sm.part_empty[buf], ((uint32_t)(i_t / NBUF) & 1u) ^ 1u);   // H is done with this buffer
        if (ksub == 0) {
          float* pa = sm.part + buf * PART_BUF + slice * PART_SET + rp * PART_ROW;
          float* pb = pa + 8 * PART_ROW;
#pragma unroll
          for (int c = 0; c < NC; ++c) { pa[c] = fa[c]; pb[c] = fb[c]; }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.part_full[buf]);
        if (tid == 0 && tile == blockIdx.x) stamp(a, s, ST_F_FIRST_DONE);
      }
      if (tid == 0) stamp(a, s, ST_F_LAST_DONE);
  } else if (warp == H_WARP) {
    // =============================== H: loss head ===============================
    const int row = lane & 15, half = lane >> 4;          // classes 5*half .. 5*half+4
    const int loss_kind = a.loss_kind;
    const float inv_B = a.inv_B;
      float db[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
      float loss_acc = 0.f, corr_acc = 0.f;
      float bias5[5];
#pragma unroll
      for (int cc = 0; cc < 5; ++cc) bias5[cc] = __ldcg(bg + 5 * half + cc);   // rewritten by the previous step's boundary
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++i_t) {
        const int buf = i_t % NBUF;
        const uint32_t par = (uint32_t)(i_t / NBUF) & 1u;
        const int valid = min(TILE_ROWS, rows - tile * TILE_ROWS);
        const int n = 2 * i_t + (row >> 3);
        mbar_wait(&sm.full[n % NSTAGES], (uint32_t)(n / NSTAGES) & 1u);     // labels of my row have landed
        mbar_wait(&sm.part_full[buf], par);
        const float* pp = sm.part + buf * PART_BUF + row * PART_ROW + 5 * half;
        float z[5];
#pragma unroll
        for (int cc = 0; cc < 5; ++cc) {
          float acc = 0.f;
#pragma unroll
          for (int sl = 0; sl < KSLICES; ++sl) acc += pp[sl * PART_SET + cc];
          z[cc] = acc + bias5[cc];
        }
        const float* tp = sm.ys + ((n % NSTAGES) * STAGE_ROWS + (row & 7)) * NC + 5 * half;
        float t[5];
#pragma unroll
        for (int cc = 0; cc < 5; ++cc) t[cc] = tp[cc];
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.part_empty[buf]);                    // F may overwrite the partials
        float dval[5], rowloss;
        bool hit = false;
        if (loss_kind == B2_LOSS_SSE_SUM) {
          // simple.py:20 on this layer: loss = sum (z - y)^2, dz = 2 (z - y); no accuracy
          float l = 0.f;
#pragma unroll
          for (int cc = 0; cc < 5; ++cc) {
            const float d = z[cc] - t[cc];
            l = fmaf(d, d, l);
            dval[cc] = 2.0f * d;
          }
          const float l_o = __shfl_xor_sync(0xffffffffu, l, 16);
          rowloss = half ? (l_o + l) : (l + l_o);
        } else {
          float m = fmaxf(fmaxf(fmaxf(z[0], z[1]), fmaxf(z[2], z[3])), z[4]);
          m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 16));
          float e[5], ssum = 0.f;
#pragma unroll
          for (int cc = 0; cc < 5; ++cc) { e[cc] = expf(z[cc] - m); ssum += e[cc]; }
          ssum += __shfl_xor_sync(0xffffffffu, ssum, 16);
          const float inv = 1.0f / ssum;
          float y[5], l = 0.f, sy = 0.f;
          float bv = -INFINITY, lv = -INFINITY;
          int bi = 0, li = 0;
#pragma unroll
          for (int cc = 0; cc < 5; ++cc) {
            y[cc] = e[cc] * inv;
            l -= t[cc] * logf(y[cc]);            // un-fused log(softmax) as in the reference (distributed.py:102)
            sy += t[cc];
            if (y[cc] > bv) { bv = y[cc]; bi = 5 * half + cc; }
            if (t[cc] > lv) { lv = t[cc]; li = 5 * half + cc; }
          }
          // combine the two class halves (lane ^ 16); the l sum is formed low half first on both lanes
          const float l_o = __shfl_xor_sync(0xffffffffu, l, 16);
          const float sy_o = __shfl_xor_sync(0xffffffffu, sy, 16);
          rowloss = half ? (l_o + l) : (l + l_o);
          sy = half ? (sy_o + sy) : (sy + sy_o);
          {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, 16);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, 16);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
            const float ov2 = __shfl_xor_sync(0xffffffffu, lv, 16);
            const int oi2 = __shfl_xor_sync(0xffffffffu, li, 16);
            if (ov2 > lv || (ov2 == lv && oi2 < li)) { lv = ov2; li = oi2; }
          }
          hit = bi == li;
#pragma unroll
          for (int cc = 0; cc < 5; ++cc) dval[cc] = (y[cc] * sy - t[cc]) * inv_B;
        }
        const bool ract = row < valid;
        mbar_wait(&sm.dz_empty[buf], par ^ 1u);                            // G is done with this dz buffer
        float* dp = sm.dz + buf * DZ_BUF + row * DZ_ROW + 8 * half;
#pragma unroll
        for (int cc = 0; cc < 5; ++cc) {
          const float d = ract ? dval[cc] : 0.f;
          dp[cc] = d;
          db[cc] += d;
        }
        if (half == 0 && ract) {
          loss_acc += rowloss;
          corr_acc += hit ? 1.0f : 0.0f;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.dz_full[buf]);
        if (lane == 0 && tile == blockIdx.x) stamp(a, s, ST_H_FIRST_DONE);
      }
      // ---- step boundary: my sums into shared memory, then the CTA partial goes to scratch together with G's
#pragma unroll
      for (int cc = 0; cc < 5; ++cc) sm.red_db[lane * 5 + cc] = db[cc];
      if (half == 0) { sm.red_loss[row] = loss_acc; sm.red_corr[row] = corr_acc; }
      named_sync(NB_PART, PART_THREADS);                                   // staging complete (G's dW, my sums)
      if (lane == 0) stamp(a, s, ST_H_LAST_DONE);
      if (!(a.debug & 16)) write_partial(a, sm, G_WARPS * 32 + lane);
      named_sync(NB_PART, PART_THREADS);                                   // the CTA partial is written (published by G's fence)
  } else {
    // =============================== G: backward ===============================
    const int gt = tid - F_WARPS * 32;            // 0..223
    const int kg = gt >> 1, h = gt & 1;           // inputs 4kg..4kg+3 and 392+4kg..392+4kg+3, classes 5h..5h+4
    const bool active = kg < KGROUPS;
      u64 gacc[4][5];                               // [input pair][class]
#pragma unroll
      for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int c = 0; c < 5; ++c) gacc[p][c] = 0ull;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++i_t) {
        const int buf = i_t % NBUF;
        const int valid = min(TILE_ROWS, rows - tile * TILE_ROWS);
        mbar_wait(&sm.dz_full[buf], (uint32_t)(i_t / NBUF) & 1u);
        if (gt == 0 && tile == blockIdx.x) stamp(a, s, ST_G_FIRST);
#pragma unroll 1
        for (int q = 0; q < STAGES_PER_TILE; ++q) {
          const int n = 2 * i_t + q, slot = n % NSTAGES;
          mbar_wait(&sm.full[slot], (uint32_t)(n / NSTAGES) & 1u);          // acquire the bulk-copied rows
          int nrows = valid - q * STAGE_ROWS;
          nrows = nrows < 0 ? 0 : (nrows > STAGE_ROWS ? STAGE_ROWS : nrows);
          if (active) {
            const float* xp = sm.x + (slot * STAGE_ROWS) * XSTRIDE + 4 * kg;
            const float* dp = sm.dz + buf * DZ_BUF + (q * STAGE_ROWS) * DZ_ROW + 8 * h;
            auto bwd_row = [&](int rr) {
              const ulonglong2 x0 = *reinterpret_cast<const ulonglong2*>(xp + rr * XSTRIDE);
              const ulonglong2 x1 = *reinterpret_cast<const ulonglong2*>(xp + rr * XSTRIDE + K / 2);
              const float4 d03 = *reinterpret_cast<const float4*>(dp + rr * DZ_ROW);
              const float d4 = dp[rr * DZ_ROW + 4];
              const float ds[5] = {d03.x, d03.y, d03.z, d03.w, d4};
#pragma unroll
              for (int c = 0; c < 5; ++c) {
                const u64 dd = pack2(ds[c], ds[c]);
                gacc[0][c] = ffma2(x0.x, dd, gacc[0][c]);
                gacc[1][c] = ffma2(x0.y, dd, gacc[1][c]);
                gacc[2][c] = ffma2(x1.x, dd, gacc[2][c]);
                gacc[3][c] = ffma2(x1.y, dd, gacc[3][c]);
              }
            };
            if (nrows == STAGE_ROWS) {
#pragma unroll
              for (int rr = 0; rr < STAGE_ROWS; ++rr) bwd_row(rr);
            } else {
              for (int rr = 0; rr < nrows; ++rr) bwd_row(rr);
            }
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(&sm.empty[slot]);
        }
        if (lane == 0) mbar_arrive(&sm.dz_empty[buf]);
      }
      if (gt == 0) stamp(a, s, ST_G_LAST);
      // ---- step boundary.  The forward team has finished with the weight tile (my last tile's dz needed its last
      // partial logits), so the per-CTA weight-gradient partial is staged there and then written out coalesced.
      if (active) {
        float* out = sm.W + 5 * h;
#pragma unroll
        for (int p = 0; p < 4; ++p) {
          const int k0 = (p < 2 ? 4 * kg : K / 2 + 4 * kg) + 2 * (p & 1);
#pragma unroll
          for (int c = 0; c < 5; ++c) {
            float lo, hi;
            unpack2(gacc[p][c], lo, hi);
            out[k0 * NC + c] = lo;
            out[(k0 + 1) * NC + c] = hi;
          }
        }
      }
      named_sync(NB_PART, PART_THREADS);                                   // staging complete (my dW, H's sums)
      if (gt == 0) stamp(a, s, ST_G_STAGED);
      if (!(a.debug & 16)) write_partial(a, sm, gt);
      named_sync(NB_PART, PART_THREADS);                                   // the CTA partial is written ...
      if (gt == 0) {
        stamp(a, s, ST_G_WRITTEN);
        if (!(a.debug & 8)) grid_arrive_wait(a.bar, (unsigned)(s + 1) * gridDim.x);   // ... published (fence) + grid barrier a
        stamp(a, s, ST_G_BAR_A);
      }
  }
  step_boundary(a, sm, s, u);
  }
}

}  // namespace fl

int b2_fused_linear_ms_prepare(void) {
  B2_CUDA(cudaFuncSetAttribute(fl::fused_linear_multistep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               (int)sizeof(fl::Smem)));
  return B2_OK;
}

// n_steps consecutive steps in one launch.  `bar` (2 counters) is zeroed on the stream first.  `grid` CTAs must all be
// resident at once: the caller passes at most the SM count (one 227 KB CTA per SM).
int b2_launch_fused_linear_multistep(const float* x_pool, const float* y_pool, long span, long first_row, long global_batch,
                                     long rank_off, int rows, int n_steps, float* params, float* grads, float* scratch,
                                     unsigned* bar, int grid, float lr, int loss, float* loss_trace, int* correct_trace,
                                     float* global_step, long P, const b2_fx_args* fx, cudaStream_t s) {
  if (n_steps <= 0) return B2_OK;
  fl::ms_args a;
  memset(&a, 0, sizeof(a));
  a.x_pool = x_pool; a.y_pool = y_pool; a.span = span; a.first_row = first_row; a.global_batch = global_batch;
  a.rank_off = rank_off; a.rows = rows; a.n_steps = n_steps; a.params = params; a.grads = grads; a.scratch = scratch;
  a.bar = bar; a.lr = lr; a.inv_B = 1.0f / (float)global_batch; a.loss_kind = loss; a.loss_trace = loss_trace;
  a.correct_trace = correct_trace; a.global_step = global_step; a.P = P;
  if (fx) { a.exchange = 1; a.fx = *fx; }
  // diagnostics: B2_MS_STAMPS=1 prints CTA 0's timeline of two consecutive steps in the middle of the call (synchronises)
  static unsigned long long* g_stamps = nullptr;
  const char* dbg = getenv("B2_MS_DEBUG");
  a.debug = dbg ? atoi(dbg) : 0;
  const char* env = getenv("B2_MS_STAMPS");
  const bool stamps = env && *env == '1' && n_steps >= 4;
  if (stamps) {
    if (!g_stamps) B2_CUDA(cudaMalloc((void**)&g_stamps, 2 * fl::ST_POINTS * sizeof(unsigned long long)));
    B2_CUDA(cudaMemsetAsync(g_stamps, 0, 2 * fl::ST_POINTS * sizeof(unsigned long long), s));
    a.stamps = g_stamps;
    a.stamp_step = n_steps / 2;
  }
  B2_CUDA(cudaMemsetAsync(bar, 0, 2 * sizeof(unsigned), s));
  fl::fused_linear_multistep_kernel<<<grid, fl::THREADS, (int)sizeof(fl::Smem), s>>>(a);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    b2_set_error("fused_linear_multistep launch failed: %s", cudaGetErrorString(e));
    return B2_ERR_CUDA;
  }
  if (stamps) {
    unsigned long long h_stamps[2 * fl::ST_POINTS];
    B2_CUDA(cudaMemcpyAsync(h_stamps, g_stamps, sizeof(h_stamps), cudaMemcpyDeviceToHost, s));
    B2_CUDA(cudaStreamSynchronize(s));
    static const char* names[fl::ST_POINTS] = {"G first tile in", "G last tile done", "G staged", "partial written", "barrier a passed",
                                               "[A] all partials visible", "slice reduced+applied", "slice barrier", "barrier b passed",
                                               "[C] all slices applied", "[D] W reloaded", "F first tile in", "F first tile done",
                                               "F last tile done", "H first tile done", "H last tile done"};
    unsigned long long t0 = h_stamps[fl::ST_G_LAST];
    fprintf(stderr, "B2_MS_STAMPS rows=%d grid=%d steps=%d: CTA 0, microseconds after 'G last tile done' of step %d\n", rows, grid,
            n_steps, a.stamp_step);
    for (int k = 0; k < 2; ++k)
      for (int p = 0; p < fl::ST_POINTS; ++p)
        if (h_stamps[k * fl::ST_POINTS + p])
          fprintf(stderr, "  step+%d %-26s %9.3f\n", k, names[p], ((double)h_stamps[k * fl::ST_POINTS + p] - (double)t0) / 1e3);
  }
  return B2_OK;
}
