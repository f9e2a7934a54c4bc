// Whole training runs of TINY models in one launch: simple/simple.py as shipped is W*x+b on four points, 1005 steps
// (simple.py:11-35) -- about ten flops per step.  Launching kernels per step makes that run pure launch latency
// (1005 x 4-5 launches); here ONE warp keeps the parameters, activations and gradients on the SM and loops over all
// n_steps itself.  Any stack with <= 4 layers, every width <= 16 and <= 32 rows per step qualifies, with both loss
// heads and both activations, so small test models take the same path.
//
// Per step, exactly the graph the layered kernels evaluate (oracle/tf1_graph.py):
//   forward   a[l+1] = act(a[l] W[l] + b[l])                      simple.py:16, distributed.py:96-98
//   head      SSE-sum (simple.py:20) or softmax + un-fused log + mean (distributed.py:99-102,117-118)
//   backward  dA = dZ W^T * act'(a), dW = a^T dZ, db = colsum dZ  TF1 MatMul / BiasAdd / Sigmoid gradients
//   apply     var -= lr * grad, global_step += 1                  simple.py:22-23, distributed.py:107-108
// Every sum runs in a fixed order (ascending k / class / row) with one owner per output: bit-reproducible.
//
// Work split inside the warp (what makes a step a few hundred clocks instead of a few thousand):
//   * forward, head and dA: LANE = ROW.  A row's activations, logits and dz live in registers; W[l][k][0..15] is one
//     broadcast LDS.128 per four outputs, so the sixteen output chains of a row are independent FMA streams.
//   * dW, db and the apply: LANE = (input k, every second output n).  Activations and dz are published TRANSPOSED
//     ([feature][row], row stride 36 floats: conflict-free for both the per-row stores and the per-feature LDS.128), the
//     owner reads four rows per load and runs its chains over the rows in ascending order.
//   * the rows of the NEXT step are fetched with cp.async straight into the transposed layout while this step computes.
// All loops are unrolled to the compile-time bounds with warp-uniform guards, so there is no integer division and no
// loop bookkeeping on the critical path.
#include <math.h>

#include "common.cuh"
#ifdef B2_TINY_PROF   // per-phase SM clocks of the last step (diagnostic build only: nvcc -DB2_TINY_PROF)
#include <stdio.h>
#define TN_STAMP(i) do { __syncwarp(); if (lane == 0) prof[i] = clock64(); } while (0)
#else
#define TN_STAMP(i) do { } while (0)
#endif

namespace tn {

constexpr int MAX_L = 4, MAX_W = 16, MAX_B = 32;
constexpr int MAX_P = MAX_L * (MAX_W * MAX_W + MAX_W);
constexpr int RS = MAX_B + 4;          // row stride of the transposed [feature][row] arrays (floats)
constexpr unsigned FULL = 0xffffffffu;

struct tiny_args {
  int n_layers, act, loss;
  int dims[MAX_L + 1];
  int w_off[MAX_L], b_off[MAX_L];   // offsets in the flat parameter buffer
  int P;                            // floats in the flat parameter buffer
  const float* x_pool;
  const float* y_pool;
  long span, first_row, global_batch;
  int rows, n_steps;
  float lr, inv_B;
  float* params;
  float* grads;                     // [P + 2]: last step's gradients, loss, correct
  float* loss_trace;
  int* correct_trace;
  float* global_step;
};

__device__ __forceinline__ void cp_async4(float* dst, const float* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}

struct Smem {
  float W[MAX_L][MAX_W][MAX_W];     // W[l][k][n], rows padded to 16 floats (zeros beyond the layer's width)
  float B[MAX_L][MAX_W];
  float AT[MAX_L][MAX_W][RS];       // AT[l][k][r]: input of layer l >= 1, transposed
  float XT[2][MAX_W][RS];           // this step's / the next step's rows, transposed, filled by cp.async
  float YT[2][MAX_W][RS];           // likewise the targets
  float DT[MAX_W][RS];              // dz of the layer being differentiated, transposed
  float row_loss[MAX_B];
  unsigned short dst_x[MAX_B * MAX_W], dst_y[MAX_B * MAX_W];   // element i of a batch -> its slot in XT / YT
};

// The loops below are unrolled to compile-time bounds (register arrays need constant indices).  A single warp pays
// ~60 clocks for every TAKEN branch over unrolled code and ~30 for every shared-memory load it has to wait for, so
//   * loops run in blocks of four (inputs k, outputs n, rows r) and leave with ONE `break` per loop: a block's loads
//     are all in flight before its FMAs start, and nothing is skipped body by body;
//   * the extents are template parameters rounded UP: WQ = quads of the widest layer (1: every width <= 4, simple.py;
//     4: <= 16), NQ = output quads of a layer (1, 2, 4), BQ = row quads (1, 2, 4, 8);
//   * the rounding is exact because the padding holds zeros: W and b beyond a layer's width, dz and the row losses of
//     rows >= B, x and y of rows >= B and of features beyond the model's.  fma(a, 0, g) and g + 0 leave g unchanged
//     (a finite), so every sum keeps the value of the plain ascending chain; only an exactly-zero sum may change the
//     sign of its zero.

// z[n] += sum_k in[k] W[k][n], ascending k                                  simple.py:16, distributed.py:96,98
template <int NQ, int WQ>
__device__ __forceinline__ void forward_layer(const float (*W)[MAX_W], int K, const float* in, float* z) {
#pragma unroll
  for (int kq = 0; kq < WQ; ++kq) {
    if (kq > 0 && 4 * kq >= K) break;
    float4 w[4][NQ];
#pragma unroll
    for (int kk = 0; kk < 4; ++kk)
#pragma unroll
      for (int q = 0; q < NQ; ++q)
        w[kk][q] = *reinterpret_cast<const float4*>(&W[4 * kq + kk][4 * q]);   // the same address in every lane: a broadcast
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const float av = in[4 * kq + kk];
#pragma unroll
      for (int q = 0; q < NQ; ++q) {
        z[4 * q + 0] = fmaf(av, w[kk][q].x, z[4 * q + 0]);
        z[4 * q + 1] = fmaf(av, w[kk][q].y, z[4 * q + 1]);
        z[4 * q + 2] = fmaf(av, w[kk][q].z, z[4 * q + 2]);
        z[4 * q + 3] = fmaf(av, w[kk][q].w, z[4 * q + 3]);
      }
    }
  }
}

// dn[k] = (sum_n d[n] W[k][n], ascending n) * act'(a[k])                    TF1 MatMul / Sigmoid gradients
template <int NQ, int WQ>
__device__ __forceinline__ void dgrad_layer(const float (*W)[MAX_W], int K, const float* d, const float* a_col, bool sig,
                                            float* dn) {
#pragma unroll
  for (int kq = 0; kq < WQ; ++kq) {
    if (kq > 0 && 4 * kq >= K) break;
    float4 w[4][NQ];
    float av[4];
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
      for (int q = 0; q < NQ; ++q) w[kk][q] = *reinterpret_cast<const float4*>(&W[4 * kq + kk][4 * q]);
      av[kk] = a_col[(4 * kq + kk) * RS];
    }
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      float acc = 0.f;
#pragma unroll
      for (int q = 0; q < NQ; ++q) {
        acc = fmaf(d[4 * q + 0], w[kk][q].x, acc);
        acc = fmaf(d[4 * q + 1], w[kk][q].y, acc);
        acc = fmaf(d[4 * q + 2], w[kk][q].z, acc);
        acc = fmaf(d[4 * q + 3], w[kk][q].w, acc);
      }
      dn[4 * kq + kk] = sig ? acc * av[kk] * (1.0f - av[kk]) : acc;       // rows k >= K of W are zero: dn stays 0 there
    }
  }
}

// sum over the rows of d[r] (plain adds), ascending r                       BiasAdd gradient, reduce_sum / reduce_mean
template <int BQ>
__device__ __forceinline__ float row_sum(const float* drow) {
  float4 d4[BQ];
#pragma unroll
  for (int q = 0; q < BQ; ++q) d4[q] = *reinterpret_cast<const float4*>(drow + 4 * q);
  float g = 0.f;
#pragma unroll
  for (int q = 0; q < BQ; ++q) {
    g += d4[q].x;
    g += d4[q].y;
    g += d4[q].z;
    g += d4[q].w;
  }
  return g;
}

// The owner lanes: dW[k][n] = sum_r a[r][k] dz[r][n] (one FMA chain over ascending r), db[n] = sum_r dz[r][n], and the
// SGD apply on the shared-memory copy of the variables                      simple.py:22-23, distributed.py:107-108
// No lane branches: every lane runs the chains of "its" (k, n) -- in-bounds loads of padding or stale columns for the
// lanes that own nothing -- and only the stores are predicated.
template <int BQ, int WQ>
__device__ __forceinline__ void wgrad_apply(Smem& sm, const tiny_args& a, int l, int K, int N, const float* at_base,
                                            int lane, bool last) {
  constexpr int HALF = WQ == 1 ? 4 : 16;              // lanes per output parity: lane = (k, h), owner of W[l][k][h + 2j]
  constexpr int NJ = 2 * WQ;                          // outputs per owner
  constexpr int JB = BQ >= 8 ? 2 : (NJ < 4 ? NJ : 4); // chains run side by side (bounded by the registers their loads need)
  const int k = lane % HALF, h = (lane / HALF) & 1;
  const bool owner = lane < 2 * HALF && k < K;
  const float* at = at_base + k * RS;
  float4 a4[BQ];
#pragma unroll
  for (int q = 0; q < BQ; ++q) a4[q] = *reinterpret_cast<const float4*>(at + 4 * q);
  const float* brow = &sm.DT[lane & (MAX_W - 1)][0];  // db: lane n < N owns b[l][n]
  float4 b4[BQ];
#pragma unroll
  for (int q = 0; q < BQ; ++q) b4[q] = *reinterpret_cast<const float4*>(brow + 4 * q);
#pragma unroll
  for (int jb = 0; jb < NJ / JB; ++jb) {
    if (jb > 0 && 2 * JB * jb >= N) break;            // warp-uniform: the block's smallest output index
    float4 d4[JB][BQ];
#pragma unroll
    for (int jj = 0; jj < JB; ++jj)
#pragma unroll
      for (int q = 0; q < BQ; ++q)                    // columns >= N hold stale dz: their sums are computed and dropped
        d4[jj][q] = *reinterpret_cast<const float4*>(&sm.DT[h + 2 * (JB * jb + jj)][4 * q]);
#pragma unroll
    for (int jj = 0; jj < JB; ++jj) {
      const int n = h + 2 * (JB * jb + jj);
      float g = 0.f;
#pragma unroll
      for (int q = 0; q < BQ; ++q) {
        g = fmaf(a4[q].x, d4[jj][q].x, g);
        g = fmaf(a4[q].y, d4[jj][q].y, g);
        g = fmaf(a4[q].z, d4[jj][q].z, g);
        g = fmaf(a4[q].w, d4[jj][q].w, g);
      }
      if (owner && n < N) {
        sm.W[l][k][n] = fmaf(-a.lr, g, sm.W[l][k][n]);
        if (last) a.grads[a.w_off[l] + k * N + n] = g;
      }
    }
  }
  {
    float g = 0.f;
#pragma unroll
    for (int q = 0; q < BQ; ++q) {
      g += b4[q].x;
      g += b4[q].y;
      g += b4[q].z;
      g += b4[q].w;
    }
    if (lane < N) {
      sm.B[l][lane] = fmaf(-a.lr, g, sm.B[l][lane]);
      if (last) a.grads[a.b_off[l] + lane] = g;
    }
  }
}

template <int BQ, int WQ>   // row quads (rows per step rounded up to 4, 8, 16, 32) and quads of the widest layer (4 or 16 wide)
__global__ void __launch_bounds__(32) tiny_steps_kernel(const tiny_args a) {
  constexpr int MW = 4 * WQ;
  __shared__ __align__(16) Smem sm;
  const int lane = threadIdx.x;
  const int L = a.n_layers, B = a.rows;
  const int in0 = a.dims[0], C = a.dims[L];
  // zero everything once: the padding has to BE zero (see above), and stays zero
  for (int i = lane; i < (int)(sizeof(Smem) / 4); i += 32) reinterpret_cast<float*>(&sm)[i] = 0.f;
  __syncwarp();
  for (int l = 0; l < L; ++l) {
    const int K = a.dims[l], N = a.dims[l + 1];
    for (int i = lane; i < K * N; i += 32) sm.W[l][i / N][i % N] = a.params[a.w_off[l] + i];
    for (int i = lane; i < N; i += 32) sm.B[l][i] = a.params[a.b_off[l] + i];
  }
  for (int i = lane; i < B * in0; i += 32) sm.dst_x[i] = (unsigned short)((i % in0) * RS + i / in0);
  for (int i = lane; i < B * C; i += 32) sm.dst_y[i] = (unsigned short)((i % C) * RS + i / C);
  float gs = *a.global_step;
  const bool real_row = lane < B;
  __syncwarp();

  auto fetch = [&](int buf, long row0) {              // asynchronous, transposing copy of one step's rows
    float* xt = &sm.XT[buf][0][0];
    float* yt = &sm.YT[buf][0][0];
    const float* xs = a.x_pool + row0 * in0;
    const float* ys = a.y_pool + row0 * C;
    for (int i = lane; i < B * in0; i += 32) cp_async4(xt + sm.dst_x[i], xs + i);
    for (int i = lane; i < B * C; i += 32) cp_async4(yt + sm.dst_y[i], ys + i);
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  long start = a.first_row % a.span;                  // batches wrap inside the pool; kept incrementally (no 64-bit modulo per step)
  fetch(0, start);
#ifdef B2_TINY_PROF
  long long prof[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#endif
#pragma unroll 1
  for (int s = 0; s < a.n_steps; ++s) {
    TN_STAMP(0);
    const int buf = s & 1;
    const bool last = s + 1 == a.n_steps;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    start = start + a.global_batch >= a.span ? start + a.global_batch - a.span : start + a.global_batch;
    if (!last) fetch(buf ^ 1, start);                 // the other buffer was last read before the barrier that ended step s-1
    TN_STAMP(1);

    // ---- forward: lane = row, the row's activations in registers
    float in[MW], z[MW];
#pragma unroll
    for (int k = 0; k < MW; ++k) in[k] = 0.f;
#pragma unroll
    for (int kq = 0; kq < WQ; ++kq) {                 // features beyond in0 are zero in XT
      if (kq > 0 && 4 * kq >= in0) break;
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) in[4 * kq + kk] = sm.XT[buf][4 * kq + kk][lane];
    }
#pragma unroll 1
    for (int l = 0; l < L; ++l) {
      const int K = a.dims[l], N = a.dims[l + 1];
#pragma unroll
      for (int n = 0; n < MW; ++n) z[n] = 0.f;
      if (WQ == 1 || N <= 4) forward_layer<1, WQ>(sm.W[l], K, in, z);
      else if (N <= 8) forward_layer<2, WQ>(sm.W[l], K, in, z);
      else forward_layer<4, WQ>(sm.W[l], K, in, z);
      const bool sig = a.act == B2_ACT_SIGMOID && l < L - 1;
#pragma unroll
      for (int q = 0; q < WQ; ++q) {
        if (q > 0 && 4 * q >= N) break;
        const float4 b4 = *reinterpret_cast<const float4*>(&sm.B[l][4 * q]);
        z[4 * q + 0] += b4.x;
        z[4 * q + 1] += b4.y;
        z[4 * q + 2] += b4.z;
        z[4 * q + 3] += b4.w;
        if (sig) {                                    // only columns < N are ever read again (W[l+1] has zero rows beyond N)
#pragma unroll
          for (int c = 0; c < 4; ++c) z[4 * q + c] = 1.0f / (1.0f + expf(-z[4 * q + c]));
        }
        if (l + 1 < L) {
#pragma unroll
          for (int c = 0; c < 4; ++c)
            if (4 * q + c < N) sm.AT[l + 1][4 * q + c][lane] = z[4 * q + c];      // rows >= N of AT stay zero
        }
      }
#pragma unroll
      for (int n = 0; n < MW; ++n) in[n] = z[n];
    }
    TN_STAMP(2);

    // ---- loss head: lane = row, z = the row's logits / predictions
    float d[MW];
#pragma unroll
    for (int c = 0; c < MW; ++c) d[c] = 0.f;
    float lrow = 0.f;
    bool hit = false;
    const float* yt = &sm.YT[buf][0][lane];
    if (a.loss == B2_LOSS_SSE_SUM) {
#pragma unroll
      for (int q = 0; q < WQ; ++q) {
        if (q > 0 && 4 * q >= C) break;
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          const int c = 4 * q + cc;                   // columns >= C: prediction 0 (zero W, b), target 0: dd = 0
          const float dd = z[c] - yt[c * RS];
          lrow = fmaf(dd, dd, lrow);
          d[c] = 2.0f * dd;
        }
      }
    } else {
      float m = -INFINITY;
#pragma unroll
      for (int q = 0; q < WQ; ++q) {
        if (q > 0 && 4 * q >= C) break;
#pragma unroll
        for (int cc = 0; cc < 4; ++cc)
          if (4 * q + cc < C) m = fmaxf(m, z[4 * q + cc]);
      }
      float e[MW], t[MW];
      float ssum = 0.f;
#pragma unroll
      for (int c = 0; c < MW; ++c) { e[c] = 0.f; t[c] = 0.f; }
#pragma unroll
      for (int q = 0; q < WQ; ++q) {
        if (q > 0 && 4 * q >= C) break;
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          const int c = 4 * q + cc;
          t[c] = yt[c * RS];
          if (c < C) {
            e[c] = expf(z[c] - m);
            ssum += e[c];
          }
        }
      }
      const float inv = 1.0f / ssum;
      float sy = 0.f, bv = -INFINITY, lv = -INFINITY;
      int bi = 0, li = 0;
#pragma unroll
      for (int q = 0; q < WQ; ++q) {
        if (q > 0 && 4 * q >= C) break;
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          const int c = 4 * q + cc;
          if (c < C) {
            const float y = e[c] * inv;
            lrow -= t[c] * logf(y);                   // un-fused log(softmax), distributed.py:102
            sy += t[c];
            if (y > bv) { bv = y; bi = c; }           // first index wins ties (tf.argmax)
            if (t[c] > lv) { lv = t[c]; li = c; }
            e[c] = y;
          }
        }
      }
#pragma unroll
      for (int q = 0; q < WQ; ++q) {
        if (q > 0 && 4 * q >= C) break;
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          const int c = 4 * q + cc;
          if (c < C) d[c] = (e[c] * sy - t[c]) * a.inv_B;
        }
      }
      hit = bi == li;
    }
    TN_STAMP(3);
    sm.row_loss[lane] = real_row ? lrow : 0.f;
    const unsigned hits = __ballot_sync(FULL, hit && real_row);
    __syncwarp();
    {
      float lsum = row_sum<BQ>(sm.row_loss);          // every lane runs the same ascending-row chain
      if (a.loss != B2_LOSS_SSE_SUM) lsum *= a.inv_B;
      if (lane == 0) {
        if (a.loss_trace) a.loss_trace[s] = lsum;
        if (a.correct_trace) a.correct_trace[s] = __popc(hits);
        if (last) { a.grads[a.P] = lsum; a.grads[a.P + 1] = (float)__popc(hits); }
      }
      gs += 1.0f;
    }
    TN_STAMP(4);

    // ---- backward + apply
    float dn[MW];
#pragma unroll 1
    for (int l = L - 1; l >= 0; --l) {
      const int K = a.dims[l], N = a.dims[l + 1];
#pragma unroll
      for (int q = 0; q < WQ; ++q) {
        if (q > 0 && 4 * q >= N) break;
#pragma unroll
        for (int c = 0; c < 4; ++c)                   // rows >= B contribute exact zeros to the owners' chains
          if (4 * q + c < N) sm.DT[4 * q + c][lane] = real_row ? d[4 * q + c] : 0.f;
      }
#pragma unroll
      for (int k = 0; k < MW; ++k) dn[k] = 0.f;
      if (l > 0) {                                    // dA = dZ W^T with the pre-update W, times act'(a)
        const bool sig = a.act == B2_ACT_SIGMOID;
        const float* a_col = &sm.AT[l][0][lane];
        if (WQ == 1 || N <= 4) dgrad_layer<1, WQ>(sm.W[l], K, d, a_col, sig, dn);
        else if (N <= 8) dgrad_layer<2, WQ>(sm.W[l], K, d, a_col, sig, dn);
        else dgrad_layer<4, WQ>(sm.W[l], K, d, a_col, sig, dn);
      }
      __syncwarp();                                   // DT is complete and every row has read the pre-update W[l]
      wgrad_apply<BQ, WQ>(sm, a, l, K, N, l == 0 ? &sm.XT[buf][0][0] : &sm.AT[l][0][0], lane, last);
      __syncwarp();
#pragma unroll
      for (int k = 0; k < MW; ++k) d[k] = dn[k];
    }
    TN_STAMP(5);
  }
#ifdef B2_TINY_PROF
  if (lane == 0)
    printf("B2_TINY_PROF last step, SM clocks: wait+fetch %lld, forward %lld, head %lld, sums + traces %lld, backward + apply %lld\n",
           prof[1] - prof[0], prof[2] - prof[1], prof[3] - prof[2], prof[4] - prof[3], prof[5] - prof[4]);
#endif
  for (int l = 0; l < L; ++l) {
    const int K = a.dims[l], N = a.dims[l + 1];
    for (int i = lane; i < K * N; i += 32) a.params[a.w_off[l] + i] = sm.W[l][i / N][i % N];
    for (int i = lane; i < N; i += 32) a.params[a.b_off[l] + i] = sm.B[l][i];
  }
  if (lane == 0) *a.global_step = gs;
}

template <int WQ>
static void launch_rows(const tiny_args& a, cudaStream_t s) {
  if (a.rows <= 4) tiny_steps_kernel<1, WQ><<<1, 32, 0, s>>>(a);
  else if (a.rows <= 8) tiny_steps_kernel<2, WQ><<<1, 32, 0, s>>>(a);
  else if (a.rows <= 16) tiny_steps_kernel<4, WQ><<<1, 32, 0, s>>>(a);
  else tiny_steps_kernel<8, WQ><<<1, 32, 0, s>>>(a);
}

}  // namespace tn

bool b2_tiny_steps_applies(int n_layers, const int* dims, int rows) {
  if (n_layers < 1 || n_layers > tn::MAX_L || rows < 1 || rows > tn::MAX_B) return false;
  for (int i = 0; i <= n_layers; ++i)
    if (dims[i] > tn::MAX_W) return false;
  return true;
}

int b2_launch_tiny_steps(int n_layers, const int* dims, int act, int loss, const long* w_off, const long* b_off, long P,
                         const float* x_pool, const float* y_pool, long span, long first_row, long global_batch, int rows,
                         int n_steps, float lr, float* params, float* grads, float* loss_trace, int* correct_trace,
                         float* global_step, cudaStream_t s) {
  if (n_steps <= 0) return B2_OK;
  if (P > tn::MAX_P) { b2_set_error("tiny_steps: parameter buffer too large"); return B2_ERR_ARG; }
  tn::tiny_args a;
  memset(&a, 0, sizeof(a));
  a.n_layers = n_layers; a.act = act; a.loss = loss;
  for (int i = 0; i <= n_layers; ++i) a.dims[i] = dims[i];
  for (int l = 0; l < n_layers; ++l) { a.w_off[l] = (int)w_off[l]; a.b_off[l] = (int)b_off[l]; }
  a.P = (int)P;
  a.x_pool = x_pool; a.y_pool = y_pool; a.span = span; a.first_row = first_row; a.global_batch = global_batch;
  a.rows = rows; a.n_steps = n_steps; a.lr = lr; a.inv_B = 1.0f / (float)global_batch;
  a.params = params; a.grads = grads; a.loss_trace = loss_trace; a.correct_trace = correct_trace; a.global_step = global_step;
  int widest = 0;
  for (int i = 0; i <= n_layers; ++i) widest = dims[i] > widest ? dims[i] : widest;
  if (widest <= 4) tn::launch_rows<1>(a, s);
  else tn::launch_rows<4>(a, s);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    b2_set_error("tiny_steps launch failed: %s", cudaGetErrorString(e));
    return B2_ERR_CUDA;
  }
  return B2_OK;
}
