// Generic fp32 layer kernels (CUDA cores) for any [in, hidden..., out] stack:
//   K1 dense forward          matmul + bias + activation       distributed.py:96-98, simple.py:16   (sgemm_ffma2.cu)
//   K2 loss head              softmax-CE / SSE forward+backward  distributed.py:99-102,117-118, simple.py:20
//   K3 gradient reductions    dW = A^T dZ, db = colsum(dZ)      TF1 MatMul/BiasAdd gradients         (sgemm_ffma2.cu)
//                             dA = dZ W^T * act'                 TF1 MatMul/Sigmoid gradients        (dgrad_kernel, here)
//   K4 reduce + SGD apply     var -= lr * grad, global_step += 1  distributed.py:107-108, simple.py:22-23
// All batch-dimension reductions are split into fixed slices and summed in a fixed order: no float
// atomics, so a step is bit-reproducible.
#include <math.h>

#include "common.cuh"

namespace {

constexpr int BM = 64, BN = 64, BK = 16;
constexpr int PADT = 4;  // smem row padding (floats)


// ---------------------------------------------------------------------------------------------
// K3b: dA[M,Kin] = (dZ[M,N] @ W[Kin,N]^T) * act'(Aout)       (MatMul grad transpose_b + SigmoidGrad)
// grid = (ceil(Kin/64), ceil(M/64))
// ---------------------------------------------------------------------------------------------
template <bool VEC_N>
__global__ void __launch_bounds__(256)
dgrad_kernel(const float* __restrict__ dZ, const float* __restrict__ W,
             const float* __restrict__ Aout, float* __restrict__ dA, int M, int Kin, int N, int act) {
  __shared__ __align__(16) float Zs[BK][BM + PADT];  // Zs[n][m]
  __shared__ __align__(16) float Ws[BK][BN + PADT];  // Ws[n][k]
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * BM, k0 = blockIdx.x * BN;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  const int l_r = tid >> 2, l_nq = (tid & 3) * 4;  // 64 rows x 16 n

  for (int nb = 0; nb < N; nb += BK) {
    const int n = nb + l_nq;
    {
      const int m = m0 + l_r;
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      if (m < M) {
        if (VEC_N) {
          if (n < N) {
            const float4 t = *reinterpret_cast<const float4*>(dZ + (size_t)m * N + n);
            v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (n + j < N) v[j] = dZ[(size_t)m * N + n + j];
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) Zs[l_nq + j][l_r] = v[j];
    }
    {
      const int k = k0 + l_r;
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      if (k < Kin) {
        if (VEC_N) {
          if (n < N) {
            const float4 t = *reinterpret_cast<const float4*>(W + (size_t)k * N + n);
            v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (n + j < N) v[j] = W[(size_t)k * N + n + j];
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) Ws[l_nq + j][l_r] = v[j];
    }
    __syncthreads();
#pragma unroll
    for (int nn = 0; nn < BK; ++nn) {
      const float4 a = *reinterpret_cast<const float4*>(&Zs[nn][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Ws[nn][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w};
      const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k = k0 + tx * 4 + j;
      if (k >= Kin) continue;
      float g = acc[i][j];
      if (act == B2_ACT_SIGMOID) {
        const float a = Aout[(size_t)m * Kin + k];
        g = g * a * (1.0f - a);  // SigmoidGrad: dy * y * (1 - y)
      }
      dA[(size_t)m * Kin + k] = g;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// K2: loss head.  One warp per row, rows strided over a fixed grid, warp-shuffle row reductions.
//   softmax-CE (distributed.py:99-102): y = exp(z-max) * (1/sum); loss_row = -sum(y_*log(y));
//       dz = (y * sum(y_) - y_) * (1/B)   [= TF1's (g - sum(g*y))*y with g = -y_/(B*y)]
//       correct = argmax(y) == argmax(y_), first index wins ties (distributed.py:117)
//   SSE (simple.py:20): loss = sum((z-y_)^2), dz = 2*(z-y_)
// Per-block partial sums (fixed order) go to loss_part / correct_part.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// argmax with first-index tie-break: larger value wins; equal values -> smaller index wins.
__device__ __forceinline__ void warp_argmax(float& v, int& idx) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, v, o);
    const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
    if (ov > v || (ov == v && oi < idx)) { v = ov; idx = oi; }
  }
}

__global__ void __launch_bounds__(256)
loss_head_kernel(const float* __restrict__ z, const float* __restrict__ y_, float* __restrict__ dz,
                 float* __restrict__ loss_part, float* __restrict__ correct_part, int rows, int C,
                 int loss_kind, float inv_B) {
  __shared__ float s_loss[8], s_corr[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float loss_acc = 0.f, corr_acc = 0.f;
  for (long row = (long)blockIdx.x * 8 + warp; row < rows; row += (long)gridDim.x * 8) {
    const float* zr = z + row * C;
    const float* yr = y_ + row * C;
    float* dr = dz + row * C;
    if (loss_kind == B2_LOSS_SSE_SUM) {
      float l = 0.f;
      for (int c = lane; c < C; c += 32) {
        const float d = zr[c] - yr[c];
        l = fmaf(d, d, l);
        dr[c] = 2.0f * d;
      }
      loss_acc += warp_sum(l);
    } else {
      float m = -INFINITY;
      for (int c = lane; c < C; c += 32) m = fmaxf(m, zr[c]);
      m = warp_max(m);
      float s = 0.f;
      for (int c = lane; c < C; c += 32) s += expf(zr[c] - m);
      s = warp_sum(s);
      const float inv = 1.0f / s;
      float l = 0.f, sy = 0.f;
      float bv = -INFINITY, lv = -INFINITY;
      int bi = 0x7fffffff, li = 0x7fffffff;
      for (int c = lane; c < C; c += 32) {
        const float y = expf(zr[c] - m) * inv;
        const float t = yr[c];
        l -= t * logf(y);  // un-fused log(softmax), as the reference writes it
        sy += t;
        if (y > bv) { bv = y; bi = c; }
        if (t > lv) { lv = t; li = c; }
      }
      l = warp_sum(l);
      sy = warp_sum(sy);
      warp_argmax(bv, bi);
      warp_argmax(lv, li);
      for (int c = lane; c < C; c += 32) {
        const float y = expf(zr[c] - m) * inv;
        dr[c] = (y * sy - yr[c]) * inv_B;
      }
      loss_acc += l;
      corr_acc += (bi == li) ? 1.0f : 0.0f;
    }
  }
  if (lane == 0) { s_loss[warp] = loss_acc; s_corr[warp] = corr_acc; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float l = 0.f, c = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) { l += s_loss[w]; c += s_corr[w]; }
    loss_part[blockIdx.x] = (loss_kind == B2_LOSS_SSE_SUM) ? l : l * inv_B;
    correct_part[blockIdx.x] = c;
  }
}

// ---------------------------------------------------------------------------------------------
// K4: fixed-order reduction of the split partials, optionally fused with the SGD apply
// (ApplyGradientDescent: var -= lr*grad; then global_step += 1 -- distributed.py:107-108).
// ---------------------------------------------------------------------------------------------
__device__ float block0_head_sum(const float* __restrict__ part, int n) {
  // warp 0 of the block: lane-strided partial sums, then a shuffle tree.  Fixed order.
  float v = 0.f;
  for (int i = threadIdx.x; i < n; i += 32) v += part[i];
  return warp_sum(v);
}

// Layers with many partials (the narrow-head kernel writes one per CTA) are summed by a warp per output (lanes stride
// over the partials, then a fixed shuffle tree); the others by a thread per output, eight loads in flight.
constexpr int REDUCE_WARP_MODE_SPLITS = 64;

__device__ __forceinline__ void reduce_locate(const b2_reduce_args& a, bool warp_mode, long e, int& l, long& src, long& dst) {
  long r = e;
  for (l = 0; l < a.n_layers; ++l) {
    if ((a.splits[l] > REDUCE_WARP_MODE_SPLITS) != warp_mode) continue;
    const long sz = a.w_size[l] + a.b_size[l];
    if (r < sz) break;
    r -= sz;
  }
  if (r < a.w_size[l]) { src = r; dst = a.w_off[l] + r; }
  else { src = ((a.w_size[l] + 3) & ~3L) + (r - a.w_size[l]); dst = a.b_off[l] + (r - a.w_size[l]); }
}

__global__ void __launch_bounds__(256)
reduce_kernel(b2_reduce_args a, const float* __restrict__ scratch, float* __restrict__ grads,
              float* __restrict__ params, int apply, float lr, float* __restrict__ loss_trace,
              int* __restrict__ correct_trace, float* __restrict__ global_step, long total_thread, long total_warp,
              int blocks_thread, b2_fx_args fx) {
  if ((int)blockIdx.x < blocks_thread) {
    const long e = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e < total_thread) {
      int l; long src, dst;
      reduce_locate(a, false, e, l, src, dst);
      const long pstride = a.pstride[l];
      const float* p = scratch + a.part_off[l] + src;
      const int S = a.splits[l];
      float g = 0.f;
      int s = 0;
      for (; s + 8 <= S; s += 8) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = p[(size_t)(s + u) * pstride];
        g += ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]));
      }
      for (; s < S; ++s) g += p[(size_t)s * pstride];
      if (fx.nranks > 1) g = b2_fx_exchange(fx, dst, g);
      grads[dst] = g;
      if (apply) params[dst] -= lr * g;
    }
  } else {
    const int lane = threadIdx.x & 31;
    const long e = (long)(blockIdx.x - blocks_thread) * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (e < total_warp) {
      int l; long src, dst;
      reduce_locate(a, true, e, l, src, dst);
      const long pstride = a.pstride[l];
      const float* p = scratch + a.part_off[l] + src;
      const int S = a.splits[l];
      float g = 0.f;
      for (int s = lane; s < S; s += 32) g += p[(size_t)s * pstride];
      g = warp_sum(g);
      if (lane == 0) {
        if (fx.nranks > 1) g = b2_fx_exchange(fx, dst, g);
        grads[dst] = g;
        if (apply) params[dst] -= lr * g;
      }
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < 32) {
    float loss = block0_head_sum(scratch + a.head_part_off, a.head_blocks);
    float corr = block0_head_sum(scratch + a.head_corr_off, a.head_blocks);
    if (threadIdx.x == 0) {
      if (fx.nranks > 1) {
        loss = b2_fx_exchange(fx, a.P, loss);
        corr = b2_fx_exchange(fx, a.P + 1, corr);
      }
      grads[a.P] = loss;
      grads[a.P + 1] = corr;
      if (apply) {
        if (loss_trace) *loss_trace = loss;
        if (correct_trace) *correct_trace = (int)(corr + 0.5f);
        *global_step += 1.0f;
      }
    }
  }
}

__global__ void __launch_bounds__(256)
sgd_apply_kernel(float* __restrict__ params, const float* __restrict__ grads, long P, float lr,
                 float* __restrict__ loss_trace, int* __restrict__ correct_trace,
                 float* __restrict__ global_step) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < P;
       i += (long)gridDim.x * blockDim.x)
    params[i] -= lr * grads[i];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    if (loss_trace) *loss_trace = grads[P];
    if (correct_trace) *correct_trace = (int)(grads[P + 1] + 0.5f);
    *global_step += 1.0f;
  }
}

// Sharded update: this rank owns n consecutive parameters; `lc` = {loss, correct} summed over the ranks (all-reduced on
// the side, since they live in one rank's shard of the gradient vector only).
__global__ void __launch_bounds__(256)
sharded_apply_kernel(float* __restrict__ params_shard, const float* __restrict__ grads_shard, long n, float lr,
                     const float* __restrict__ lc, float* __restrict__ loss_trace, int* __restrict__ correct_trace,
                     float* __restrict__ global_step) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
    params_shard[i] -= lr * grads_shard[i];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    if (loss_trace) *loss_trace = lc[0];
    if (correct_trace) *correct_trace = (int)(lc[1] + 0.5f);
    *global_step += 1.0f;
  }
}

__global__ void __launch_bounds__(256) scale_kernel(float* __restrict__ p, long n, float a) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) p[i] *= a;
}

__global__ void eval_accumulate_kernel(const float* __restrict__ loss_part,
                                       const float* __restrict__ correct_part, int blocks,
                                       float* __restrict__ acc) {
  const float l = block0_head_sum(loss_part, blocks);
  const float c = block0_head_sum(correct_part, blocks);
  if (threadIdx.x == 0) { acc[0] += l; acc[1] += c; }
}

__global__ void eval_finish_kernel(const float* __restrict__ acc, float* loss_out, long* correct_out) {
  if (threadIdx.x == 0) {
    if (loss_out) *loss_out = acc[0];
    if (correct_out) *correct_out = (long)(acc[1] + 0.5f);
  }
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace

#define B2_LAUNCH_CHECK()                                                               \
  do {                                                                                  \
    cudaError_t _e = cudaGetLastError();                                                \
    if (_e != cudaSuccess) {                                                            \
      b2_set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
      return B2_ERR_CUDA;                                                               \
    }                                                                                   \
  } while (0)

int b2_launch_dgrad(const float* dZ, const float* W, const float* Aout, float* dA, int M, int Kin,
                    int N, int act, cudaStream_t s) {
  dim3 grid((Kin + BN - 1) / BN, (M + BM - 1) / BM);
  const bool vn = (N % 4 == 0) && aligned16(dZ) && aligned16(W);
  if (vn) dgrad_kernel<true><<<grid, 256, 0, s>>>(dZ, W, Aout, dA, M, Kin, N, act);
  else dgrad_kernel<false><<<grid, 256, 0, s>>>(dZ, W, Aout, dA, M, Kin, N, act);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_launch_loss_head(const float* z, const float* y, float* dz, float* loss_part,
                        float* correct_part, int rows, int C, int loss, float inv_global_batch,
                        int blocks, cudaStream_t s) {
  loss_head_kernel<<<blocks, 256, 0, s>>>(z, y, dz, loss_part, correct_part, rows, C, loss,
                                          inv_global_batch);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_launch_reduce(const b2_reduce_args& a, const float* scratch, float* grads, float* params,
                     int apply, float lr, float* loss_trace, int* correct_trace,
                     float* global_step, int sm_count, const b2_fx_args* fx, cudaStream_t s) {
  (void)sm_count;
  b2_fx_args none;
  memset(&none, 0, sizeof(none));
  long total_thread = 0, total_warp = 0;
  for (int l = 0; l < a.n_layers; ++l) {
    if (a.splits[l] > REDUCE_WARP_MODE_SPLITS) total_warp += a.w_size[l] + a.b_size[l];
    else total_thread += a.w_size[l] + a.b_size[l];
  }
  const long blocks_thread = (total_thread + 255) / 256, blocks_warp = (total_warp + 7) / 8;
  long blocks = blocks_thread + blocks_warp;
  if (blocks < 1) blocks = 1;
  reduce_kernel<<<(int)blocks, 256, 0, s>>>(a, scratch, grads, params, apply, lr, loss_trace, correct_trace, global_step,
                                            total_thread, total_warp, (int)blocks_thread, fx ? *fx : none);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_launch_sgd_apply(float* params, const float* grads, long P, float lr, float* loss_trace,
                        int* correct_trace, float* global_step, int sm_count, cudaStream_t s) {
  long blocks = (P + 255) / 256;
  const long cap = (long)sm_count * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  sgd_apply_kernel<<<(int)blocks, 256, 0, s>>>(params, grads, P, lr, loss_trace, correct_trace,
                                               global_step);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_launch_sharded_apply(float* params_shard, const float* grads_shard, long n, float lr, const float* lc, float* loss_trace,
                            int* correct_trace, float* global_step, int sm_count, cudaStream_t s) {
  long blocks = (n + 255) / 256;
  const long cap = (long)sm_count * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  sharded_apply_kernel<<<(int)blocks, 256, 0, s>>>(params_shard, grads_shard, n, lr, lc, loss_trace, correct_trace, global_step);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_launch_scale(float* p, long n, float a, int sm_count, cudaStream_t s) {
  long blocks = (n + 255) / 256;
  if (blocks > 8L * sm_count) blocks = 8L * sm_count;
  if (blocks < 1) blocks = 1;
  scale_kernel<<<(int)blocks, 256, 0, s>>>(p, n, a);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_launch_eval_accumulate(const float* loss_part, const float* correct_part, int blocks,
                              float* acc, cudaStream_t s) {
  eval_accumulate_kernel<<<1, 32, 0, s>>>(loss_part, correct_part, blocks, acc);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_launch_eval_finish(const float* acc, float* loss_out, long* correct_out, cudaStream_t s) {
  eval_finish_kernel<<<1, 32, 0, s>>>(acc, loss_out, correct_out);
  B2_LAUNCH_CHECK();
  return B2_OK;
}
