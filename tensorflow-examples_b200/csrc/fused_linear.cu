// Single-kernel training step for the "784 -> 10 + softmax cross-entropy" model (BASELINE.json
// configs[2]: the simple.py-style single dense layer on MNIST-shaped rows with the
// distributed.py loss head).  One pass over HBM: every x row is fetched once (3136 B) by
// cp.async.bulk into shared memory and used twice from there -- for the logits
// (distributed.py:98 z = x@W + b) and for the weight gradient (TF1 MatMul grad, dW = x^T dz).
//
// Persistent grid, one CTA per SM, 16 warps working as a pipeline over 16-row tiles; the stages
// hand tiles to each other through mbarriers, so nobody waits on a CTA-wide barrier:
//   warp 14     P  producer: cp.async.bulk of rows + labels into a 7-stage ring of 8-row stages
//   warps 0-6   F  forward of tile t+1: warp = 112-wide slice of the 784 inputs, lane = (4 input
//                  sub-slices) x (8 row pairs): 2 rows x 10 classes per thread, packed fp32x2 FMAs
//                  (FFMA2) over class pairs: the W pair comes straight from the [784][10] tile in
//                  shared memory (one bulk copy), the x value is the scalar-broadcast operand
//   warp 15     H  head of tile t+1: sums the 7 partial logits, softmax / loss / accuracy / dz
//                  (lane = row x class-half, one shuffle per reduction), dz stored as (d,d) pairs
//   warps 7-13  G  backward of tile t: thread = 8 inputs x 5 classes, 20 packed accumulators (input
//                  pairs) kept for the whole kernel; rows stream from shared memory, dz is the
//                  scalar-broadcast operand; frees ring stages as it goes
// Each CTA ends by writing its partial dW / db / loss / correct to scratch[cta][7856];
// fused_linear_reduce_kernel sums the partials in CTA order (no float atomics) and applies SGD.
#include "fused_linear.cuh"

namespace fl {

__global__ void __launch_bounds__(THREADS, 1)
fused_linear_step_kernel(const float* __restrict__ x, const float* __restrict__ y_, int rows,
                         const float* __restrict__ Wg, const float* __restrict__ bg,
                         float* __restrict__ scratch, float inv_B, int loss_kind) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int ntiles = (rows + TILE_ROWS - 1) / TILE_ROWS;
  pdl_launch_dependents();   // the reduction kernel may be queued now; it waits for this grid before reading scratch

  if (warp == P_WARP) {
    // =============================== P: producer ===============================
    // x and y do not depend on the previous step, so their loads start right away; W and b (updated by the
    // previous step's reduce + SGD kernel) are only touched after pdl_wait().
    // one barrier per lane, initialised in parallel
    if (lane < NSTAGES) mbar_init(&sm.full[lane], 1);
    else if (lane < 2 * NSTAGES) mbar_init(&sm.empty[lane - NSTAGES], G_WARPS);
    else if (lane < 2 * NSTAGES + NBUF) mbar_init(&sm.part_full[lane - 2 * NSTAGES], F_WARPS);
    else if (lane < 2 * NSTAGES + 2 * NBUF) mbar_init(&sm.part_empty[lane - 2 * NSTAGES - NBUF], 1);
    else if (lane < 2 * NSTAGES + 3 * NBUF) mbar_init(&sm.dz_full[lane - 2 * NSTAGES - 2 * NBUF], 1);
    else if (lane < 2 * NSTAGES + 4 * NBUF) mbar_init(&sm.dz_empty[lane - 2 * NSTAGES - 3 * NBUF], G_WARPS);
    else if (lane == 2 * NSTAGES + 4 * NBUF) mbar_init(&sm.w_full, 1);
    static_assert(2 * NSTAGES + 4 * NBUF + 1 <= 32, "one lane per barrier");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();
    asm volatile("bar.arrive 2, %0;" ::"n"(THREADS) : "memory");  // barriers are live
    const bool y_aligned = (reinterpret_cast<uintptr_t>(y_) & 15u) == 0;
    const bool w_bulk = (reinterpret_cast<uintptr_t>(Wg) & 15u) == 0;
    const unsigned long long x_policy = l2_evict_first_policy();
    bool w_issued = false;
    auto issue_w = [&]() {
      pdl_wait();                                            // W was written by the previous step's kernels
      if (w_bulk) {
        if (lane == 0) {
          bulk_g2s(sm.W, Wg, K * NC * 4, &sm.w_full);
          mbar_arrive_expect_tx(&sm.w_full, K * NC * 4);
        }
      } else {
        for (int i = lane; i < K * NC; i += 32) sm.W[i] = Wg[i];
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.w_full);
      }
      w_issued = true;
    };
    int n = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      for (int q = 0; q < STAGES_PER_TILE; ++q, ++n) {
        if (n == STAGES_PER_TILE && !w_issued) issue_w();   // first tile's rows are in flight: now fetch the weights
        const int slot = n % NSTAGES;
        const uint32_t phase = (uint32_t)(n / NSTAGES) & 1u;
        mbar_wait(&sm.empty[slot], phase ^ 1u);
        const long row0 = (long)tile * TILE_ROWS + q * STAGE_ROWS;
        long left = rows - row0;
        const int nrows = left <= 0 ? 0 : (left < STAGE_ROWS ? (int)left : STAGE_ROWS);
        const bool y_bulk = y_aligned && ((nrows & 1) == 0) && nrows > 0;   // 40*nrows bytes must be a 16 B multiple
        if (lane == 0) {
          // x is read once per step: evict-first keeps the stream from pushing the partials and W out of L2 (-1.4 us per
          // step); one lane issuing the rows measured better than one row per lane
          for (int rr = 0; rr < nrows; ++rr)
            bulk_g2s_hint(sm.x + (slot * STAGE_ROWS + rr) * XSTRIDE, x + (row0 + rr) * K, ROW_BYTES, &sm.full[slot], x_policy);
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
    if (!w_issued) issue_w();
  } else {
    asm volatile("bar.sync 2, %0;" ::"n"(THREADS) : "memory");

    if (warp < F_WARPS) {
      // =============================== F: forward ===============================
      const int slice = warp;                       // inputs 112*slice ..
      const int ksub = lane >> 3, rp = lane & 7;    // quads 4i+ksub; rows rp and rp+8
      mbar_wait(&sm.w_full, 0u);
      int i_t = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++i_t) {
        const int na = 2 * i_t, nb = na + 1, buf = i_t % NBUF;
        mbar_wait(&sm.full[na % NSTAGES], (uint32_t)(na / NSTAGES) & 1u);
        mbar_wait(&sm.full[nb % NSTAGES], (uint32_t)(nb / NSTAGES) & 1u);
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
            // first input: class pairs (0,1)(2,3)(4,5)(6,7)(8,9) = w01.x w01.y w23.x w23.y w45.x
            const u64 xa0 = pack2(xsa[2 * kp], xsa[2 * kp]), xb0 = pack2(xsb[2 * kp], xsb[2 * kp]);
            acc_a[0] = ffma2(xa0, w01.x, acc_a[0]); acc_b[0] = ffma2(xb0, w01.x, acc_b[0]);
            acc_a[1] = ffma2(xa0, w01.y, acc_a[1]); acc_b[1] = ffma2(xb0, w01.y, acc_b[1]);
            acc_a[2] = ffma2(xa0, w23.x, acc_a[2]); acc_b[2] = ffma2(xb0, w23.x, acc_b[2]);
            acc_a[3] = ffma2(xa0, w23.y, acc_a[3]); acc_b[3] = ffma2(xb0, w23.y, acc_b[3]);
            acc_a[4] = ffma2(xa0, w45.x, acc_a[4]); acc_b[4] = ffma2(xb0, w45.x, acc_b[4]);
            // second input: w45.y w67.x w67.y w89.x w89.y
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
      }
    } else if (warp == H_WARP) {
      // =============================== H: loss head ===============================
      const int row = lane & 15, half = lane >> 4;          // classes 5*half .. 5*half+4
      float db[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
      float loss_acc = 0.f, corr_acc = 0.f;
      float bias5[5];
      pdl_wait();                                            // b was written by the previous step's kernels
#pragma unroll
      for (int cc = 0; cc < 5; ++cc) bias5[cc] = bg[5 * half + cc];
      int i_t = 0;
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
          float a = 0.f;
#pragma unroll
          for (int s = 0; s < KSLICES; ++s) a += pp[s * PART_SET + cc];
          z[cc] = a + bias5[cc];
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
            l -= t[cc] * logf(y[cc]);            // un-fused log(softmax) as in the reference
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
      }
#pragma unroll
      for (int cc = 0; cc < 5; ++cc) sm.red_db[lane * 5 + cc] = db[cc];
      if (half == 0) { sm.red_loss[row] = loss_acc; sm.red_corr[row] = corr_acc; }
    } else {
      // =============================== G: backward ===============================
      const int gt = tid - F_WARPS * 32;            // 0..223
      const int kg = gt >> 1, h = gt & 1;           // inputs 4kg..4kg+3 and 392+4kg..392+4kg+3, classes 5h..5h+4
      const bool active = kg < KGROUPS;
      u64 gacc[4][5];                               // [input pair][class]
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < 5; ++c) gacc[a][c] = 0ull;
      int i_t = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++i_t) {
        const int buf = i_t % NBUF;
        const int valid = min(TILE_ROWS, rows - tile * TILE_ROWS);
        mbar_wait(&sm.dz_full[buf], (uint32_t)(i_t / NBUF) & 1u);
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
              // the common case, fully unrolled: addresses become immediates, no loop bookkeeping, loads hoisted freely
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
      // per-CTA weight-gradient partial: rows 4kg..4kg+3 and 392+4kg.. of dW[784][10], classes 5h..5h+4.
      // Staged through the weight tile's shared memory (the forward team has finished with it: the last
      // backward tile depends on the last forward tile) so that the global write below is fully coalesced.
      if (active) {
        float* out = sm.W + 5 * h;
#pragma unroll
        for (int a = 0; a < 4; ++a) {
          const int k0 = (a < 2 ? 4 * kg : K / 2 + 4 * kg) + 2 * (a & 1);
#pragma unroll
          for (int c = 0; c < 5; ++c) {
            float lo, hi;
            unpack2(gacc[a][c], lo, hi);
            out[k0 * NC + c] = lo;
            out[(k0 + 1) * NC + c] = hi;
          }
        }
      }
    }
  }
  __syncthreads();
  pdl_wait();   // the previous reduction kernel must have finished reading scratch before it is overwritten
  {
    float4* dst = reinterpret_cast<float4*>(scratch + (size_t)blockIdx.x * SCRATCH_STRIDE);
    const float4* src = reinterpret_cast<const float4*>(sm.W);
    for (int i = tid; i < K * NC / 4; i += THREADS) dst[i] = src[i];
  }
  // bias gradient, loss, correct count: fixed-order sums over the head warp's lanes
  if (tid < NC) {
    const int half = tid / 5, cc = tid - 5 * half;
    float v = 0.f;
    for (int r = 0; r < 16; ++r) v += sm.red_db[(half * 16 + r) * 5 + cc];
    scratch[(size_t)blockIdx.x * SCRATCH_STRIDE + K * NC + tid] = v;
  } else if (tid == 32) {
    float v = 0.f;
    for (int r = 0; r < 16; ++r) v += sm.red_loss[r];
    scratch[(size_t)blockIdx.x * SCRATCH_STRIDE + LOSS_IDX] = loss_kind == B2_LOSS_SSE_SUM ? v : v * inv_B;
  } else if (tid == 64) {
    float v = 0.f;
    for (int r = 0; r < 16; ++r) v += sm.red_corr[r];
    scratch[(size_t)blockIdx.x * SCRATCH_STRIDE + CORRECT_IDX] = v;
  }
}

// Sums the per-CTA partials in a fixed order; with apply != 0 also does var -= lr*grad, writes the
// step's loss / correct count and bumps global_step (distributed.py:107-108).
// Block = 32 outputs x 4 quarter-sums: thread (o, q) adds CTAs q*ceil(grid/4) .. in order, the four
// quarters are then combined in order 0,1,2,3 -- the same association on every run.
__global__ void __launch_bounds__(RED_OUT * RED_Q)
fused_linear_reduce_kernel(const float* __restrict__ scratch, int grid, float* __restrict__ grads,
                           float* __restrict__ params, int apply, float lr,
                           float* __restrict__ loss_trace, int* __restrict__ correct_trace,
                           float* __restrict__ global_step, long P) {
  __shared__ float sh[RED_Q][RED_OUT];
  pdl_launch_dependents();   // the next step kernel may start its prologue and its x loads
  pdl_wait();                // the step kernel's partials must be complete and visible
  const int o = threadIdx.x & (RED_OUT - 1), q = threadIdx.x / RED_OUT;
  const int i = blockIdx.x * RED_OUT + o;
  const int per = (grid + RED_Q - 1) / RED_Q;
  const int c0 = q * per, c1 = min(grid, c0 + per);
  float g = 0.f;
  if (i < RED_N) {
    // all of this quarter's partials are requested before the first add (one L2 round trip instead of five): the loads go
    // out unconditionally from clamped addresses, the adds run in CTA order over the valid ones -- same sum as before
    const float* p = scratch + i;
    const int clast = max(c1 - 1, 0);
    for (int k0 = c0; k0 < c1; k0 += RED_CHUNK) {
      float v[RED_CHUNK];
#pragma unroll
      for (int u = 0; u < RED_CHUNK; ++u) v[u] = p[(size_t)min(k0 + u, clast) * SCRATCH_STRIDE];
#pragma unroll
      for (int u = 0; u < RED_CHUNK; ++u)
        if (k0 + u < c1) g += v[u];
    }
  }
  sh[q][o] = g;
  __syncthreads();
  if (q != 0 || i >= RED_N) return;
  g = ((sh[0][o] + sh[1][o]) + sh[2][o]) + sh[3][o];
  if (i < K * NC + NC) {
    grads[i] = g;
    if (apply) params[i] -= lr * g;
  } else if (i == LOSS_IDX) {
    grads[P] = g;  // flat layout: W 7840, b padded to 12 -> loss at P = 7852
    if (apply && loss_trace) *loss_trace = g;
  } else {
    grads[P + 1] = g;
    if (apply) {
      if (correct_trace) *correct_trace = (int)(g + 0.5f);
      *global_step += 1.0f;
    }
  }
}


// ---------------------------------------------------------------------------------------------
// Several ranks: the same reduction, then the gradient exchange and the SGD apply in the same kernel (b2_fx_args /
// b2_fx_exchange in common.cuh): no block- or grid-wide step between "my partials are reduced" and "the sum is
// applied", and one kernel launch less than reduce + a separate exchange kernel.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(RED_OUT * RED_Q)
fused_linear_reduce_exchange_kernel(const float* __restrict__ scratch, int grid, b2_fx_args a,
                                    float* __restrict__ grads, float* __restrict__ params, float lr,
                                    float* __restrict__ loss_trace, int* __restrict__ correct_trace,
                                    float* __restrict__ global_step, long P) {
  __shared__ float sh[RED_Q][RED_OUT];
  pdl_launch_dependents();   // the next step kernel may start its prologue and its x loads
  pdl_wait();                // the step kernel's partials must be complete and visible
  const int o = threadIdx.x & (RED_OUT - 1), q = threadIdx.x / RED_OUT;
  const int i = blockIdx.x * RED_OUT + o;
  const int per = (grid + RED_Q - 1) / RED_Q;
  const int c0 = q * per, c1 = min(grid, c0 + per);
  float g = 0.f;
  if (i < RED_N) {
    // all of this quarter's partials are requested before the first add (one L2 round trip instead of five): the loads go
    // out unconditionally from clamped addresses, the adds run in CTA order over the valid ones -- same sum as before
    const float* p = scratch + i;
    const int clast = max(c1 - 1, 0);
    for (int k0 = c0; k0 < c1; k0 += RED_CHUNK) {
      float v[RED_CHUNK];
#pragma unroll
      for (int u = 0; u < RED_CHUNK; ++u) v[u] = p[(size_t)min(k0 + u, clast) * SCRATCH_STRIDE];
#pragma unroll
      for (int u = 0; u < RED_CHUNK; ++u)
        if (k0 + u < c1) g += v[u];
    }
  }
  sh[q][o] = g;
  __syncthreads();
  if (q != 0 || i >= RED_N) return;
  g = ((sh[0][o] + sh[1][o]) + sh[2][o]) + sh[3][o];       // this rank's value of output i
  const float sum = b2_fx_exchange(a, i, g);
  if (i < K * NC + NC) {
    grads[i] = sum;
    params[i] -= lr * sum;                                 // distributed.py:107-108
  } else if (i == LOSS_IDX) {
    grads[P] = sum;
    if (loss_trace) *loss_trace = sum;
  } else {
    grads[P + 1] = sum;
    if (correct_trace) *correct_trace = (int)(sum + 0.5f);
    *global_step += 1.0f;
  }
}

}  // namespace fl

bool b2_fused_linear_supported(int in_dim, int out_dim, int loss) {
  return in_dim == fl::K && out_dim == fl::NC && (loss == B2_LOSS_SOFTMAX_CE_MEAN || loss == B2_LOSS_SSE_SUM);
}
int b2_fused_linear_grid(int sm_count) { return sm_count; }
long b2_fused_linear_stride(void) { return fl::SCRATCH_STRIDE; }
int b2_fused_linear_smem_bytes(void) { return (int)sizeof(fl::Smem); }
int b2_fused_linear_tile_rows(void) { return fl::TILE_ROWS; }

int b2_fused_linear_prepare(void) {
  // opt in to the large dynamic shared memory carve-out on the current device (done once per context)
  B2_CUDA(cudaFuncSetAttribute(fl::fused_linear_step_kernel,
                               cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(fl::Smem)));
  return B2_OK;
}

int b2_launch_fused_linear_step(const float* x, const float* y, int rows, const float* W,
                                const float* b, float* scratch, int grid, float inv_global_batch, int loss,
                                cudaStream_t s) {
  const int smem = (int)sizeof(fl::Smem);
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(fl::THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, fl::fused_linear_step_kernel, x, y, rows, W, b, scratch, inv_global_batch, loss);
  if (e != cudaSuccess) {
    b2_set_error("fused_linear_step launch failed: %s", cudaGetErrorString(e));
    return B2_ERR_CUDA;
  }
  return B2_OK;
}


int b2_launch_fused_linear_reduce_exchange(const float* scratch, int grid, const b2_fx_args& a, float* grads, float* params,
                                           float lr, float* loss_trace, int* correct_trace, float* global_step, long P,
                                           cudaStream_t s) {
  static_assert((fl::RED_N + fl::RED_OUT - 1) / fl::RED_OUT == B2_FX_BLOCKS, "exchange region layout");
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(B2_FX_BLOCKS);
  cfg.blockDim = dim3(fl::RED_OUT * fl::RED_Q);
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, fl::fused_linear_reduce_exchange_kernel, scratch, grid, a, grads,
                                     params, lr, loss_trace, correct_trace, global_step, P);
  if (e != cudaSuccess) {
    b2_set_error("fused_linear_reduce_exchange launch failed: %s", cudaGetErrorString(e));
    return B2_ERR_CUDA;
  }
  return B2_OK;
}

int b2_launch_fused_linear_reduce(const float* scratch, int grid, float* grads, float* params,
                                  int apply, float lr, float* loss_trace, int* correct_trace,
                                  float* global_step, long P, cudaStream_t s) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((fl::RED_N + fl::RED_OUT - 1) / fl::RED_OUT);
  cfg.blockDim = dim3(fl::RED_OUT * fl::RED_Q);
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, fl::fused_linear_reduce_kernel, (const float*)scratch, grid, grads, params, apply,
                                     lr, loss_trace, correct_trace, global_step, P);
  if (e != cudaSuccess) {
    b2_set_error("fused_linear_reduce launch failed: %s", cudaGetErrorString(e));
    return B2_ERR_CUDA;
  }
  return B2_OK;
}

// =============================================================================================
// b2_linear_step_noalloc: the same 784 -> 10 training step for callers that must not allocate.
//
// Who needs this: code that runs where device allocation is off limits -- a plugin called from
// inside another framework's execution callback (which owns the allocator and forbids re-entrant
// cudaMalloc), a latency-critical loop that cannot afford the implicit device synchronisation
// cudaMalloc/cudaFree cause, or a host that has handed the whole device memory to its own pool
// and leaves nothing for cudaMalloc to find.  Those callers cannot use b2_create(), which
// allocates parameters, gradients and scratch.  Here the caller passes every input and output,
// and the one piece of scratch the step kernel needs -- the per-CTA partial sums -- lives in the
// file-scope device array below, which is part of the loaded library image.
//
// Consequences of that choice, all documented in include/b2mlp.h:
//   * the array has a fixed size (B2_NOALLOC_MAX_CTAS slots); every GPU that loads the library
//     holds these bytes until the process exits, used or not;
//   * every call on a device shares it, so two calls whose kernels overlap on the GPU (different
//     streams, or different host threads) race on it; calls on one stream are safe;
//   * it keeps the previous call's contents.  Nothing has to be cleared: the step kernel writes
//     every slot entry that the reduction of the same call reads (7852 floats x grid CTAs).
// =============================================================================================
#define B2_NOALLOC_MAX_CTAS 160
#define B2_NOALLOC_MAX_ROWS (1 << 20)
#define B2_NOALLOC_MAX_DEVICES 64

// 160 x 7856 floats = 5,027,840 bytes per GPU.
static __device__ __align__(256) float g_noalloc_scratch[B2_NOALLOC_MAX_CTAS * fl::SCRATCH_STRIDE];

namespace fl {

// Sums the per-CTA partials (read straight from the static array, which device code may name
// directly) in CTA order and applies plain gradient descent to the caller's W and b.
__global__ void __launch_bounds__(128)
noalloc_reduce_apply_kernel(int grid, float* __restrict__ W, float* __restrict__ b, float lr,
                            float* __restrict__ loss_out, int* __restrict__ correct_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 7852) return;
  const float* p = g_noalloc_scratch + i;
  float g = 0.f;
  int c = 0;
  for (; c + 8 <= grid; c += 8) {
    float v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = p[(size_t)(c + u) * SCRATCH_STRIDE];
#pragma unroll
    for (int u = 0; u < 8; ++u) g += v[u];
  }
  for (; c < grid; ++c) g += p[(size_t)c * SCRATCH_STRIDE];
  if (i < K * NC) W[i] -= lr * g;
  else if (i < K * NC + NC) b[i - K * NC] -= lr * g;
  else if (i == LOSS_IDX) { if (loss_out) *loss_out = g; }
  else { if (correct_out) *correct_out = (int)(g + 0.5f); }
}

struct noalloc_device_state {
  float* scratch;  // DEVICE address of g_noalloc_scratch on this device (0 = set-up not done yet)
  int sm_count;
};
static noalloc_device_state g_noalloc_dev[B2_NOALLOC_MAX_DEVICES];

}  // namespace fl

extern "C" int b2_linear_step_noalloc(const float* x_dev, const float* y_dev, int rows, long global_batch,
                                      float lr, float* W_dev, float* b_dev, float* loss_dev,
                                      int* correct_dev, b2_stream_t stream) {
  B2_CHECK_ARG(x_dev && y_dev && W_dev && b_dev, "b2_linear_step_noalloc: NULL x, y, W or b");
  B2_CHECK_ARG((reinterpret_cast<uintptr_t>(x_dev) & 15u) == 0,
               "b2_linear_step_noalloc: x_dev must be 16-byte aligned (its rows are fetched with bulk copies)");
  B2_CHECK_ARG(((reinterpret_cast<uintptr_t>(y_dev) | reinterpret_cast<uintptr_t>(W_dev) | reinterpret_cast<uintptr_t>(b_dev)) & 3u) == 0,
               "b2_linear_step_noalloc: y_dev, W_dev and b_dev must be float-aligned");
  B2_CHECK_ARG(rows >= 1, "b2_linear_step_noalloc: rows must be >= 1, got %d", rows);
  B2_CHECK_ARG(rows <= B2_NOALLOC_MAX_ROWS,
               "b2_linear_step_noalloc: rows %d exceeds the fixed-scratch limit of %d rows per call; "
               "split the batch or use b2_create()/b2_train_steps()", rows, B2_NOALLOC_MAX_ROWS);
  B2_CHECK_ARG(global_batch >= rows, "b2_linear_step_noalloc: global_batch %ld < rows %d", global_batch, rows);
  int dev = -1;
  B2_CUDA(cudaGetDevice(&dev));
  B2_CHECK_ARG(dev >= 0 && dev < B2_NOALLOC_MAX_DEVICES, "b2_linear_step_noalloc: device index %d not supported", dev);
  fl::noalloc_device_state& st = fl::g_noalloc_dev[dev];
  if (!st.scratch) {
    // One-time set-up for the CURRENT device.  In host code `g_noalloc_scratch` is only a stand-in
    // for the symbol, not an address the GPU can use, so ask the runtime where the array lives on
    // this device and hand that pointer to the kernel.
    int major = 0, sms = 0;
    B2_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    B2_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (major != 10) {
      b2_set_error("b2_linear_step_noalloc: device %d is not sm_100 (B200)", dev);
      return B2_ERR_CUDA;
    }
    float* ptr = nullptr;
    B2_CUDA(cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch));
    B2_CUDA(cudaFuncSetAttribute(fl::fused_linear_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)sizeof(fl::Smem)));
    st.sm_count = sms;
    st.scratch = ptr;
  }
  int grid = (rows + fl::TILE_ROWS - 1) / fl::TILE_ROWS;
  if (grid > st.sm_count) grid = st.sm_count;
  if (grid > B2_NOALLOC_MAX_CTAS) grid = B2_NOALLOC_MAX_CTAS;  // the persistent kernel loops over tiles
  cudaStream_t s = (cudaStream_t)stream;
  const float inv_B = 1.0f / (float)global_batch;
  fl::fused_linear_step_kernel<<<grid, fl::THREADS, (int)sizeof(fl::Smem), s>>>(x_dev, y_dev, rows, W_dev, b_dev,
                                                                              st.scratch, inv_B, (int)B2_LOSS_SOFTMAX_CE_MEAN);
  B2_CUDA(cudaGetLastError());
  fl::noalloc_reduce_apply_kernel<<<(7852 + 127) / 128, 128, 0, s>>>(grid, W_dev, b_dev, lr, loss_dev, correct_dev);
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}
