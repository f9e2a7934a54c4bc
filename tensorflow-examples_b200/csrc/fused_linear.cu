// Single-kernel training step for the "784 -> 10 + softmax cross-entropy" model (BASELINE.json
// configs[2]: the simple.py-style single dense layer on MNIST-shaped rows with the
// distributed.py loss head).  One pass over HBM: every x row is fetched once (3136 B) by
// cp.async.bulk into shared memory and used twice from there -- for the logits
// (distributed.py:98 z = x@W + b) and for the weight gradient (TF1 MatMul grad, dW = x^T dz).
//
//   persistent grid, one CTA per SM, 15 warps:
//     warp 14        producer: bulk-copies rows (8-row stages, 6-stage ring) + stages the labels
//     warps 0..13    forward:  lane = row of a 32-row tile, warp = 56-wide slice of the 784 inputs,
//                              W ([784][12] fp32, zero padded) broadcast from shared memory
//                    head:     softmax / loss / accuracy / dz with half-warp shuffle reductions
//                    backward: thread = (4 inputs, 5 classes), 20 register accumulators kept for
//                              the whole kernel; rows stream from shared memory
//   each CTA ends by writing its partial dW / db / loss / correct to scratch[cta][7856];
//   fused_linear_reduce_kernel sums the partials in CTA order (no float atomics) and applies SGD.
#include <math.h>

#include "common.cuh"

namespace fl {

constexpr int K = 784, NC = 10, NCP = 12;
constexpr int TILE_ROWS = 32, STAGE_ROWS = 8, STAGES_PER_TILE = 4, NSTAGES = 6;
constexpr int XSTRIDE = 788;  // floats per staged row: 3152 B, keeps 16 B alignment, conflict-free LDS.128
constexpr int COMPUTE_WARPS = 14, COMPUTE_THREADS = COMPUTE_WARPS * 32, THREADS = COMPUTE_THREADS + 32;
constexpr int KW = K / COMPUTE_WARPS;  // 56 inputs per warp in the forward pass
constexpr int KGROUPS = K / 4;         // 196 groups of 4 inputs in the backward pass
constexpr int ROW_BYTES = K * 4;       // 3136
constexpr long SCRATCH_STRIDE = 7856;  // 7840 dW + 10 db + loss + correct, padded to 16 B
constexpr int LOSS_IDX = 7850, CORRECT_IDX = 7851;

struct __align__(16) Smem {
  float W[K * NCP];                            // 37632 B
  float x[NSTAGES * STAGE_ROWS * XSTRIDE];     // 151296 B
  float part[COMPUTE_WARPS * NC * TILE_ROWS];  // 17920 B   part[warp][class][row]
  float ys[NSTAGES * STAGE_ROWS * NC];         // 1920 B
  float dz[TILE_ROWS * 16];                    // class c of row r at r*16 + (c/5)*8 + c%5
  float bias[16];
  float red_db[COMPUTE_WARPS * 2 * 16];
  float red_loss[COMPUTE_WARPS * 2];
  float red_corr[COMPUTE_WARPS * 2];
  unsigned long long full[NSTAGES];
  unsigned long long empty[NSTAGES];
};

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
  long long t0 = 0;
  for (;;) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(addr), "r"(parity)
        : "memory");
    if (ok) break;
    // watchdog: a protocol bug must not hang the GPU
    if (t0 == 0) t0 = clock64();
    else if (clock64() - t0 > 4000000000LL) __trap();
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
__device__ __forceinline__ void compute_bar() {
  asm volatile("bar.sync 1, %0;" ::"n"(COMPUTE_THREADS) : "memory");
}

__device__ __forceinline__ float half_sum(float v) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float half_max(float v) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ void half_argmax(float& v, int& idx) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, v, o);
    const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
    if (ov > v || (ov == v && oi < idx)) { v = ov; idx = oi; }
  }
}

__global__ void __launch_bounds__(THREADS, 1)
fused_linear_step_kernel(const float* __restrict__ x, const float* __restrict__ y_, int rows,
                         const float* __restrict__ Wg, const float* __restrict__ bg,
                         float* __restrict__ scratch, float inv_B) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int ntiles = (rows + TILE_ROWS - 1) / TILE_ROWS;

  if (tid == 0) {
    for (int s = 0; s < NSTAGES; ++s) {
      mbar_init(&sm.full[s], 1);
      mbar_init(&sm.empty[s], COMPUTE_WARPS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // stage W as [784][12] (zero padded) and the bias
  for (int i = tid; i < K * NCP; i += THREADS) {
    const int k = i / NCP, c = i - k * NCP;
    sm.W[i] = (c < NC) ? Wg[k * NC + c] : 0.f;
  }
  if (tid < 16) sm.bias[tid] = (tid < NC) ? bg[tid] : 0.f;
  __syncthreads();

  if (warp == COMPUTE_WARPS) {
    // ------------------------------ producer ------------------------------
    int n = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      for (int q = 0; q < STAGES_PER_TILE; ++q, ++n) {
        const int slot = n % NSTAGES;
        const uint32_t phase = (uint32_t)(n / NSTAGES) & 1u;
        mbar_wait(&sm.empty[slot], phase ^ 1u);
        const long row0 = (long)tile * TILE_ROWS + q * STAGE_ROWS;
        long left = rows - row0;
        const int nrows = left <= 0 ? 0 : (left < STAGE_ROWS ? (int)left : STAGE_ROWS);
        if (lane == 0) {
          for (int rr = 0; rr < nrows; ++rr)
            bulk_g2s(sm.x + (slot * STAGE_ROWS + rr) * XSTRIDE, x + (row0 + rr) * K, ROW_BYTES,
                     &sm.full[slot]);
        }
        for (int i = lane; i < STAGE_ROWS * NC; i += 32) {
          const int rr = i / NC;
          sm.ys[slot * STAGE_ROWS * NC + i] = (rr < nrows) ? y_[row0 * NC + i] : 0.f;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive_expect_tx(&sm.full[slot], (uint32_t)(nrows * ROW_BYTES));
      }
    }
  } else {
    // ------------------------------ consumers ------------------------------
    const int j = tid >> 1, h = tid & 1;  // backward ownership: inputs 4j..4j+3, classes 5h..5h+4
    const bool bwd_active = j < KGROUPS;
    float gacc[4][5];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int c = 0; c < 5; ++c) gacc[a][c] = 0.f;
    float db_acc = 0.f, loss_acc = 0.f, corr_acc = 0.f;

    int n0 = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, n0 += STAGES_PER_TILE) {
      const int valid = min(TILE_ROWS, rows - tile * TILE_ROWS);
      // ---- wait for the tile's four stages ----
#pragma unroll
      for (int q = 0; q < STAGES_PER_TILE; ++q) {
        const int n = n0 + q;
        mbar_wait(&sm.full[n % NSTAGES], (uint32_t)(n / NSTAGES) & 1u);
      }
      // ---- forward: partial logits of row `lane` over inputs [56*warp, 56*warp+56) ----
      {
        const int n = n0 + (lane >> 3);
        const float* xrow = sm.x + ((n % NSTAGES) * STAGE_ROWS + (lane & 7)) * XSTRIDE + warp * KW;
        const float* wbase = sm.W + warp * KW * NCP;
        float acc[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) acc[c] = 0.f;
#pragma unroll 2
        for (int q4 = 0; q4 < KW / 4; ++q4) {
          const float4 xv = *reinterpret_cast<const float4*>(xrow + 4 * q4);
          const float xs[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            const float4* wp = reinterpret_cast<const float4*>(wbase + (4 * q4 + kk) * NCP);
            const float4 wa = wp[0], wb = wp[1], wc = wp[2];
            acc[0] = fmaf(xs[kk], wa.x, acc[0]);
            acc[1] = fmaf(xs[kk], wa.y, acc[1]);
            acc[2] = fmaf(xs[kk], wa.z, acc[2]);
            acc[3] = fmaf(xs[kk], wa.w, acc[3]);
            acc[4] = fmaf(xs[kk], wb.x, acc[4]);
            acc[5] = fmaf(xs[kk], wb.y, acc[5]);
            acc[6] = fmaf(xs[kk], wb.z, acc[6]);
            acc[7] = fmaf(xs[kk], wb.w, acc[7]);
            acc[8] = fmaf(xs[kk], wc.x, acc[8]);
            acc[9] = fmaf(xs[kk], wc.y, acc[9]);
          }
        }
#pragma unroll
        for (int c = 0; c < NC; ++c) sm.part[(warp * NC + c) * TILE_ROWS + lane] = acc[c];
      }
      compute_bar();
      // ---- head: two rows per warp pass (one per half-warp), lane&15 = class ----
      for (int p = warp; p < TILE_ROWS / 2; p += COMPUTE_WARPS) {
        const int half = lane >> 4, c = lane & 15;
        const int row = 2 * p + half;
        const bool cact = c < NC;
        const bool ract = row < valid;
        float z = 0.f;
        if (cact) {
#pragma unroll
          for (int w = 0; w < COMPUTE_WARPS; ++w) z += sm.part[(w * NC + c) * TILE_ROWS + row];
          z += sm.bias[c];
        }
        const float m = half_max(cact ? z : -INFINITY);
        const float e = cact ? expf(z - m) : 0.f;
        const float s = half_sum(e);
        const float inv = 1.0f / s;
        const float y = e * inv;
        const int n = n0 + (row >> 3);
        const float t = cact ? sm.ys[((n % NSTAGES) * STAGE_ROWS + (row & 7)) * NC + c] : 0.f;
        const float lterm = cact ? -(t * logf(y)) : 0.f;  // un-fused log(softmax) as in the reference
        const float rowloss = half_sum(lterm);
        const float sy = half_sum(t);
        float bv = cact ? y : -INFINITY, lv = cact ? t : -INFINITY;
        int bi = c, li = c;
        half_argmax(bv, bi);
        half_argmax(lv, li);
        float dzv = (y * sy - t) * inv_B;
        if (!(cact && ract)) dzv = 0.f;
        if (cact) sm.dz[row * 16 + (c / 5) * 8 + (c % 5)] = dzv;
        db_acc += dzv;
        if (c == 0 && ract) {
          loss_acc += rowloss;
          corr_acc += (bi == li) ? 1.0f : 0.0f;
        }
      }
      compute_bar();
      // ---- backward: gacc[a][c] += x[r][4j+a] * dz[r][5h+c], rows in order; free stages as we go ----
#pragma unroll 1
      for (int q = 0; q < STAGES_PER_TILE; ++q) {
        const int slot = (n0 + q) % NSTAGES;
        int nrows = valid - q * STAGE_ROWS;
        nrows = nrows < 0 ? 0 : (nrows > STAGE_ROWS ? STAGE_ROWS : nrows);
        if (bwd_active) {
          const float* xp = sm.x + (slot * STAGE_ROWS) * XSTRIDE + 4 * j;
          const float* dp = sm.dz + (q * STAGE_ROWS) * 16 + 8 * h;
#pragma unroll 4
          for (int rr = 0; rr < nrows; ++rr) {
            const float4 xv = *reinterpret_cast<const float4*>(xp + rr * XSTRIDE);
            const float4 d4 = *reinterpret_cast<const float4*>(dp + rr * 16);
            const float d5 = dp[rr * 16 + 4];
            const float xs[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
            for (int a = 0; a < 4; ++a) {
              gacc[a][0] = fmaf(xs[a], d4.x, gacc[a][0]);
              gacc[a][1] = fmaf(xs[a], d4.y, gacc[a][1]);
              gacc[a][2] = fmaf(xs[a], d4.z, gacc[a][2]);
              gacc[a][3] = fmaf(xs[a], d4.w, gacc[a][3]);
              gacc[a][4] = fmaf(xs[a], d5, gacc[a][4]);
            }
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.empty[slot]);
      }
    }
    // ---- per-CTA partials out ----
    float* out = scratch + (size_t)blockIdx.x * SCRATCH_STRIDE;
    if (bwd_active) {
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < 5; ++c) out[(4 * j + a) * NC + 5 * h + c] = gacc[a][c];
    }
    {
      const int half = lane >> 4, c = lane & 15;
      sm.red_db[(warp * 2 + half) * 16 + c] = db_acc;
      if (c == 0) {
        sm.red_loss[warp * 2 + half] = loss_acc;
        sm.red_corr[warp * 2 + half] = corr_acc;
      }
    }
  }
  __syncthreads();
  if (tid < NC) {
    float v = 0.f;
    for (int i = 0; i < COMPUTE_WARPS * 2; ++i) v += sm.red_db[i * 16 + tid];
    scratch[(size_t)blockIdx.x * SCRATCH_STRIDE + K * NC + tid] = v;
  } else if (tid == 32) {
    float v = 0.f;
    for (int i = 0; i < COMPUTE_WARPS * 2; ++i) v += sm.red_loss[i];
    scratch[(size_t)blockIdx.x * SCRATCH_STRIDE + LOSS_IDX] = v * inv_B;
  } else if (tid == 64) {
    float v = 0.f;
    for (int i = 0; i < COMPUTE_WARPS * 2; ++i) v += sm.red_corr[i];
    scratch[(size_t)blockIdx.x * SCRATCH_STRIDE + CORRECT_IDX] = v;
  }
}

// Sums the per-CTA partials in CTA order; with apply != 0 also does var -= lr*grad, writes the
// step's loss / correct count and bumps global_step (distributed.py:107-108).
__global__ void __launch_bounds__(128)
fused_linear_reduce_kernel(const float* __restrict__ scratch, int grid, float* __restrict__ grads,
                           float* __restrict__ params, int apply, float lr,
                           float* __restrict__ loss_trace, int* __restrict__ correct_trace,
                           float* __restrict__ global_step, long P) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 7852) return;
  const float* p = scratch + i;
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

}  // namespace fl

bool b2_fused_linear_supported(int in_dim, int out_dim, int loss) {
  return in_dim == fl::K && out_dim == fl::NC && loss == B2_LOSS_SOFTMAX_CE_MEAN;
}
int b2_fused_linear_grid(int sm_count) { return sm_count; }
long b2_fused_linear_stride(void) { return fl::SCRATCH_STRIDE; }
int b2_fused_linear_smem_bytes(void) { return (int)sizeof(fl::Smem); }

int b2_fused_linear_prepare(void) {
  // opt in to the large dynamic shared memory carve-out on the current device (done once per context)
  B2_CUDA(cudaFuncSetAttribute(fl::fused_linear_step_kernel,
                               cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(fl::Smem)));
  return B2_OK;
}

int b2_launch_fused_linear_step(const float* x, const float* y, int rows, const float* W,
                                const float* b, float* scratch, int grid, float inv_global_batch,
                                cudaStream_t s) {
  const int smem = (int)sizeof(fl::Smem);
  fl::fused_linear_step_kernel<<<grid, fl::THREADS, smem, s>>>(x, y, rows, W, b, scratch,
                                                               inv_global_batch);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    b2_set_error("fused_linear_step launch failed: %s", cudaGetErrorString(e));
    return B2_ERR_CUDA;
  }
  return B2_OK;
}

int b2_launch_fused_linear_reduce(const float* scratch, int grid, float* grads, float* params,
                                  int apply, float lr, float* loss_trace, int* correct_trace,
                                  float* global_step, long P, cudaStream_t s) {
  fl::fused_linear_reduce_kernel<<<(7852 + 127) / 128, 128, 0, s>>>(
      scratch, grid, grads, params, apply, lr, loss_trace, correct_trace, global_step, P);
  cudaError_t e = cudaGetLastError();
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
                                                                              st.scratch, inv_B);
  B2_CUDA(cudaGetLastError());
  fl::noalloc_reduce_apply_kernel<<<(7852 + 127) / 128, 128, 0, s>>>(grid, W_dev, b_dev, lr, loss_dev, correct_dev);
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}
