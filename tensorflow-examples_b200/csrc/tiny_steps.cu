// Whole training runs of TINY models in one launch: simple/simple.py as shipped is W*x+b on four points, 1005 steps
// (simple.py:11-35) -- about ten flops per step.  Launching kernels per step makes that run pure launch latency
// (1005 x 4-5 launches); here ONE warp keeps the parameters, activations and gradients in shared memory and loops
// over all n_steps itself.  Any stack with <= 4 layers, every width <= 16 and <= 32 rows per step qualifies, with both
// loss heads and both activations, so small test models take the same path.
//
// Per step, exactly the graph the layered kernels evaluate (oracle/tf1_graph.py):
//   forward   a[l+1] = act(a[l] W[l] + b[l])                      simple.py:16, distributed.py:96-98
//   head      SSE-sum (simple.py:20) or softmax + un-fused log + mean (distributed.py:99-102,117-118)
//   backward  dA = dZ W^T * act'(a), dW = a^T dZ, db = colsum dZ  TF1 MatMul / BiasAdd / Sigmoid gradients
//   apply     var -= lr * grad, global_step += 1                  simple.py:22-23, distributed.py:107-108
// Every sum runs in a fixed order (ascending k / row), one owner lane per output: bit-reproducible.
#include <math.h>

#include "common.cuh"

namespace tn {

constexpr int MAX_L = 4, MAX_W = 16, MAX_B = 32;
constexpr int MAX_P = MAX_L * (MAX_W * MAX_W + MAX_W);

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

__global__ void __launch_bounds__(32) tiny_steps_kernel(const tiny_args a) {
  __shared__ float sP[MAX_P];                       // parameters, flat layout
  __shared__ float sA[MAX_L + 1][MAX_B][MAX_W];     // sA[l] = input of layer l (l >= 1), sA[L] = logits / predictions
  __shared__ float sD[2][MAX_B][MAX_W];             // dz ping-pong
  __shared__ float sX[2][MAX_B * MAX_W];            // this step's / the next step's rows, compact ([rows][in0]), filled by cp.async
  __shared__ float sYb[2][MAX_B * MAX_W];           // likewise the targets ([rows][C])
  __shared__ float sRow[2][MAX_B];                  // per-row loss, hit
  const int lane = threadIdx.x;
  const int L = a.n_layers, B = a.rows;
  const int in0 = a.dims[0], C = a.dims[L];
  for (int i = lane; i < a.P; i += 32) sP[i] = a.params[i];
  float gs = *a.global_step;
  __syncwarp();

  long start = a.first_row % a.span;                  // batches wrap inside the pool; kept incrementally (no 64-bit modulo per step)
  auto fetch = [&](int buf, long row0) {              // asynchronous copy of one step's rows: lands while the previous step computes
    for (int i = lane; i < B * in0; i += 32) cp_async4(&sX[buf][i], a.x_pool + row0 * in0 + i);
    for (int i = lane; i < B * C; i += 32) cp_async4(&sYb[buf][i], a.y_pool + row0 * C + i);
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  fetch(0, start);
  for (int s = 0; s < a.n_steps; ++s) {
    const float* X = sX[s & 1];
    const float* Y = sYb[s & 1];
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    start = start + a.global_batch >= a.span ? start + a.global_batch - a.span : start + a.global_batch;
    if (s + 1 < a.n_steps) fetch((s + 1) & 1, start);   // the other buffer was last read in step s-1
    // ---- forward
    for (int l = 0; l < L; ++l) {
      const int Kin = a.dims[l], N = a.dims[l + 1];
      const float* W = sP + a.w_off[l];
      const float* b = sP + a.b_off[l];
      const bool sig = a.act == B2_ACT_SIGMOID && l < L - 1;
      for (int i = lane; i < B * N; i += 32) {
        const int r = i / N, n = i % N;
        float z = 0.f;
        if (l == 0) {
          for (int k = 0; k < Kin; ++k) z = fmaf(X[r * Kin + k], W[k * N + n], z);
        } else {
          for (int k = 0; k < Kin; ++k) z = fmaf(sA[l][r][k], W[k * N + n], z);
        }
        z += b[n];
        sA[l + 1][r][n] = sig ? 1.0f / (1.0f + expf(-z)) : z;
      }
      __syncwarp();
    }
    // ---- loss head: lane = row
    if (lane < B) {
      const int r = lane;
      float l = 0.f, hit = 0.f;
      if (a.loss == B2_LOSS_SSE_SUM) {
        for (int c = 0; c < C; ++c) {
          const float d = sA[L][r][c] - Y[r * C + c];
          l = fmaf(d, d, l);
          sD[0][r][c] = 2.0f * d;
        }
      } else {
        float m = -INFINITY;
        for (int c = 0; c < C; ++c) m = fmaxf(m, sA[L][r][c]);
        float ssum = 0.f;
        for (int c = 0; c < C; ++c) ssum += expf(sA[L][r][c] - m);
        const float inv = 1.0f / ssum;
        float sy = 0.f, bv = -INFINITY, lv = -INFINITY;
        int bi = 0, li = 0;
        for (int c = 0; c < C; ++c) {
          const float y = expf(sA[L][r][c] - m) * inv, t = Y[r * C + c];
          l -= t * logf(y);                          // un-fused log(softmax), distributed.py:102
          sy += t;
          if (y > bv) { bv = y; bi = c; }            // first index wins ties (tf.argmax)
          if (t > lv) { lv = t; li = c; }
        }
        for (int c = 0; c < C; ++c) {
          const float y = expf(sA[L][r][c] - m) * inv;
          sD[0][r][c] = (y * sy - Y[r * C + c]) * a.inv_B;
        }
        hit = bi == li ? 1.0f : 0.0f;
      }
      sRow[0][r] = l;
      sRow[1][r] = hit;
    }
    __syncwarp();
    if (lane == 0) {
      float l = 0.f, h = 0.f;
      for (int r = 0; r < B; ++r) { l += sRow[0][r]; h += sRow[1][r]; }
      if (a.loss != B2_LOSS_SSE_SUM) l *= a.inv_B;
      if (a.loss_trace) a.loss_trace[s] = l;
      if (a.correct_trace) a.correct_trace[s] = (int)(h + 0.5f);
      if (s + 1 == a.n_steps) { a.grads[a.P] = l; a.grads[a.P + 1] = h; }
      gs += 1.0f;
    }
    // ---- backward + apply
    int cur = 0;
    const bool last = s + 1 == a.n_steps;
    for (int l = L - 1; l >= 0; --l) {
      const int Kin = a.dims[l], N = a.dims[l + 1];
      float* W = sP + a.w_off[l];
      float* b = sP + a.b_off[l];
      if (l > 0) {
        const bool sig = a.act == B2_ACT_SIGMOID;
        for (int i = lane; i < B * Kin; i += 32) {
          const int r = i / Kin, k = i % Kin;
          float d = 0.f;
          for (int n = 0; n < N; ++n) d = fmaf(sD[cur][r][n], W[k * N + n], d);
          const float av = sA[l][r][k];
          sD[cur ^ 1][r][k] = sig ? d * av * (1.0f - av) : d;
        }
        __syncwarp();                                 // dA used the pre-update W
      }
      for (int i = lane; i < Kin * N + N; i += 32) {
        float g = 0.f;
        if (i < Kin * N) {
          const int k = i / N, n = i % N;
          if (l == 0) {
            for (int r = 0; r < B; ++r) g = fmaf(X[r * Kin + k], sD[cur][r][n], g);
          } else {
            for (int r = 0; r < B; ++r) g = fmaf(sA[l][r][k], sD[cur][r][n], g);
          }
          W[i] = fmaf(-a.lr, g, W[i]);
          if (last) a.grads[a.w_off[l] + i] = g;
        } else {
          const int n = i - Kin * N;
          for (int r = 0; r < B; ++r) g += sD[cur][r][n];
          b[n] = fmaf(-a.lr, g, b[n]);
          if (last) a.grads[a.b_off[l] + n] = g;
        }
      }
      __syncwarp();
      cur ^= 1;
    }
  }
  for (int i = lane; i < a.P; i += 32) a.params[i] = sP[i];
  if (lane == 0) *a.global_step = gs;
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
  tn::tiny_steps_kernel<<<1, 32, 0, s>>>(a);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    b2_set_error("tiny_steps launch failed: %s", cudaGetErrorString(e));
    return B2_ERR_CUDA;
  }
  return B2_OK;
}
