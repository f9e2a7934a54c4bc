// The output layer of distributed.py's model in ONE kernel (fp32, CUDA cores): for a layer with at most 128 inputs
// and at most 16 classes (784-100-10: 100 -> 10) the products are tiny but the activations are not (a [rows, 100]
// is 26 MB at batch 65536), so the layer is bound by how often that matrix crosses the memory system.  Done as
// separate kernels (forward GEMM, loss head, weight gradient, input gradient) it is read three times and the logits
// and dz make two more round trips; here a tile of 128 rows is staged in shared memory once and everything that
// needs it happens there:
//   z  = a . W + b                                   distributed.py:98
//   y  = softmax(z), loss, correct, dz               distributed.py:99-102,117-118 and the TF1 gradient compositions
//   dW = a^T . dz, db = column sums of dz            MatMul / BiasAdd gradients (per-CTA partials, fixed order)
//   da = dz . W^T, dz_prev = da * a * (1 - a)        MatMul gradient (transpose_b) + SigmoidGrad -> global
// HBM traffic per row: K floats in, K floats out (+ the labels).  train == 0 stops after the loss (evaluation).
#include <cuda_bf16.h>
#include <math.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace nh {

constexpr int TR = 128;            // rows per tile
constexpr int THREADS = 256;       // two threads per row in the forward / head phase
constexpr int KMAX = 128, CMAX = 16;

struct Args {
  const float* a;        // [rows, K] activations of the last hidden layer
  const float* y;        // [rows, C] labels
  const float* W;        // [K, C]
  const float* b;        // [C]
  float* dz_prev;        // [rows, K] fp32, or NULL when only the planes are wanted   (train)
  unsigned short* dz_planes;   // three bf16 planes [3][rows_max, ld_planes] of dz_prev for the tensor-core wgrad of the layer
  long plane_stride;           // below (gemm_split3.cu), or NULL; elements between planes
  int ld_planes;
  float* partW;          // partW[cta*stride + k*C + c], partB[cta*stride + c]   (train)
  float* partB;
  long part_stride;
  float* partD;          // partD[cta*strideD + k]: column sums of dz_prev = the bias gradient of the layer below (train)
  long strideD;
  float* loss_part;      // [gridDim.x]
  float* correct_part;   // [gridDim.x]
  int rows, K, C, act, loss, train;
  float inv_B;
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// CH = classes per thread of a row pair = ceil(C / 2), a template parameter so that no issue slot is spent on padding
template <int CH>
__global__ void __launch_bounds__(THREADS, 3)   // 80 registers: three CTAs per SM overlap each other's load / compute phases
narrow_head_kernel(Args p) {
  extern __shared__ __align__(16) float smem[];
  const int K = p.K, C = p.C;
  // K % 4 == 0: the tile is one contiguous block of global memory and arrives by a single bulk copy (rows unpadded);
  // otherwise threads copy it element-wise into rows padded by one float
  const bool bulk = (K & 3) == 0 && (reinterpret_cast<uintptr_t>(p.a) & 15) == 0;
  const int KP = bulk ? K : K + 1;
  // class c sits at "position" c (c < CH) or 8 + c - CH (c >= CH) of a 16-float row of sW / sdz: each thread of a row
  // pair owns one aligned half (two LDS.128 / STS.128), and the backward loop reads a whole row as four LDS.128
  float* sa = smem;                           // [TR][KP]
  float* sW = sa + ((TR * KP + 3) & ~3);      // [K][16]    (unused positions are zero)
  float* sdz = sW + KMAX * 16;                // [TR][16]
  float* sred = sdz + TR * 16;                // [16]
  uint64_t* bar = reinterpret_cast<uint64_t*>(sred + 16);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  for (int i0 = 0; i0 < K * 16; i0 += 8 * THREADS) {
    float v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {              // eight independent loads in flight per thread
      const int i = i0 + u * THREADS + tid;
      const int k = i >> 4, pos = i & 15;
      const int j = pos & 7, c = (pos >> 3) * CH + j;
      v[u] = (i < K * 16 && j < CH && c < C) ? p.W[k * C + c] : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = i0 + u * THREADS + tid;
      if (i < K * 16) sW[i] = v[u];
    }
  }
  const int r = tid >> 1, h = tid & 1;        // forward / head: row r of the tile, classes h*CH .. h*CH+CH-1
  float bias[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) bias[j] = (j < CH && h * CH + j < C) ? p.b[h * CH + j] : 0.f;

  // backward: thread (kk, half) owns input k = kk for the rows of its parity; W[kk][.] and the dW partial sums
  // stay in registers across all the tiles of this CTA
  const int kk = tid & 127, half = tid >> 7;
  float wreg[16], accW[16], accB[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) { wreg[i] = 0.f; accW[i] = 0.f; accB[i] = 0.f; }
  float loss_acc = 0.f, corr_acc = 0.f, accD = 0.f;
  if (tid == 0) {
    tcx::mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (kk < K) {
#pragma unroll
    for (int i = 0; i < 16; ++i) wreg[i] = sW[kk * 16 + i];
  }

  const int ntiles = (p.rows + TR - 1) / TR;
  uint32_t it_parity = 0;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int row0 = tile * TR;
    __syncthreads();                          // the previous tile's readers are done with sa / sdz
    if (bulk) {
      const int valid = min(TR, p.rows - row0);
      if (tid == 0) {
        const uint32_t bytes = (uint32_t)valid * (uint32_t)K * 4u;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // earlier generic reads of sa before the async write
        tcx::mbar_expect_tx(bar, bytes);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(tcx::smem_u32(sa)), "l"(p.a + (size_t)row0 * K), "r"(bytes), "r"(tcx::smem_u32(bar))
                     : "memory");
      }
      for (int i = valid * K + tid; i < TR * K; i += THREADS) sa[i] = 0.f;   // rows past the end of the batch
      tcx::mbar_wait(bar, it_parity);
      it_parity ^= 1u;
    } else {
      for (int i = tid; i < TR * K; i += THREADS) {
        const int rr = i / K, k = i - rr * K;
        sa[rr * KP + k] = row0 + rr < p.rows ? p.a[(size_t)row0 * K + i] : 0.f;
      }
    }
    __syncthreads();

    // ---- forward + loss head: two threads per row ----
    {
      const int row = row0 + r;
      const bool ok = row < p.rows;
      float z[8], t[8], dzv[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        // the labels are requested first: their latency hides behind the K loop
        const int c = h * CH + j;
        t[j] = (ok && j < CH && c < C) ? p.y[(size_t)row * C + c] : 0.f;
        z[j] = 0.f;
        dzv[j] = 0.f;
      }
      const float* ar = sa + r * KP;
      const float* wr = sW + 8 * h;
#pragma unroll 4
      for (int k = 0; k < K; ++k) {
        const float av = ar[k];
        float w[8];
        {
          const float4 w0 = *reinterpret_cast<const float4*>(wr + k * 16);
          w[0] = w0.x; w[1] = w0.y; w[2] = w0.z; w[3] = w0.w;
          if (CH > 4) {
            const float4 w1 = *reinterpret_cast<const float4*>(wr + k * 16 + 4);
            w[4] = w1.x; w[5] = w1.y; w[6] = w1.z; w[7] = w1.w;
          }
        }
#pragma unroll
        for (int j = 0; j < CH; ++j) z[j] = fmaf(av, w[j], z[j]);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) z[j] += bias[j];
      if (p.loss == B2_LOSS_SSE_SUM) {
        float l = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (j < CH && h * CH + j < C) {
            const float d = z[j] - t[j];
            l = fmaf(d, d, l);
            dzv[j] = 2.0f * d;                                  // Square gradient, simple.py:20
          }
        }
        const float other = __shfl_xor_sync(0xffffffffu, l, 1);
        l = h == 0 ? l + other : other + l;
        if (ok && h == 0) loss_acc += l;
      } else {
        float m = -INFINITY;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (j < CH && h * CH + j < C) m = fmaxf(m, z[j]);
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
        float e[8], s = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          e[j] = (j < CH && h * CH + j < C) ? expf(z[j] - m) : 0.f;
          s += e[j];
        }
        {
          // both threads of the pair add the two halves in class order (low half first): identical sums
          const float other = __shfl_xor_sync(0xffffffffu, s, 1);
          s = h == 0 ? s + other : other + s;
        }
        const float inv = 1.0f / s;                              // softmax = exp(z - max) * (1 / sum)   [TF1]
        float l = 0.f, sy = 0.f, bv = -INFINITY, lv = -INFINITY;
        int bi = 0x7fffffff, li = 0x7fffffff;
        float yv[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          yv[j] = e[j] * inv;
          if (j < CH && h * CH + j < C) {
            const int c = h * CH + j;
            l -= t[j] * logf(yv[j]);                             // un-fused log(softmax), distributed.py:102
            sy += t[j];
            if (yv[j] > bv) { bv = yv[j]; bi = c; }
            if (t[j] > lv) { lv = t[j]; li = c; }
          }
        }
        {
          const float ol = __shfl_xor_sync(0xffffffffu, l, 1), osy = __shfl_xor_sync(0xffffffffu, sy, 1);
          l = h == 0 ? l + ol : ol + l;
          sy = h == 0 ? sy + osy : osy + sy;
          const float obv = __shfl_xor_sync(0xffffffffu, bv, 1), olv = __shfl_xor_sync(0xffffffffu, lv, 1);
          const int obi = __shfl_xor_sync(0xffffffffu, bi, 1), oli = __shfl_xor_sync(0xffffffffu, li, 1);
          if (obv > bv || (obv == bv && obi < bi)) { bv = obv; bi = obi; }   // first index wins ties (argmax, :117)
          if (olv > lv || (olv == lv && oli < li)) { lv = olv; li = oli; }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (j < CH && h * CH + j < C) dzv[j] = (yv[j] * sy - t[j]) * p.inv_B;
        if (ok && h == 0) {
          loss_acc += l;
          corr_acc += (bi == li) ? 1.0f : 0.0f;
        }
      }
      if (p.train) {
        float* d = sdz + r * 16 + 8 * h;
        *reinterpret_cast<float4*>(d) = ok ? make_float4(dzv[0], dzv[1], dzv[2], dzv[3]) : make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<float4*>(d + 4) = ok ? make_float4(dzv[4], dzv[5], dzv[6], dzv[7]) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    if (!p.train) continue;                    // uniform
    __syncthreads();

    // ---- backward: one pass over the tile's rows per thread (its parity), dz row broadcast from shared memory ----
    //   dz_prev[row][kk] = (sum_c dz[c] W[kk][c]) * act'(a)      lanes -> consecutive kk: coalesced stores
    //   dW[kk][c] += a[row][kk] * dz[row][c] ;  db[c] += dz[row][c] (thread kk == 0)
    if (kk < K) {
      const int rmax = min(TR, p.rows - row0);
#pragma unroll 2
      for (int rr = half; rr < rmax; rr += 2) {
        float dzr[16];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 v = *reinterpret_cast<const float4*>(sdz + rr * 16 + 4 * q);
          dzr[4 * q] = v.x; dzr[4 * q + 1] = v.y; dzr[4 * q + 2] = v.z; dzr[4 * q + 3] = v.w;
        }
        const float av = sa[rr * KP + kk];
        float g0 = 0.f, g1 = 0.f;              // the two class halves as separate chains, low half + high half
#pragma unroll
        for (int j = 0; j < CH; ++j) {
          g0 = fmaf(dzr[j], wreg[j], g0);
          g1 = fmaf(dzr[8 + j], wreg[8 + j], g1);
          accW[j] = fmaf(av, dzr[j], accW[j]);
          accW[8 + j] = fmaf(av, dzr[8 + j], accW[8 + j]);
        }
        const float g = g0 + g1;
        if ((warp & 3) == 0) {                 // warp-uniform: the warps that hold kk == 0 (the bias-gradient owner)
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
#pragma unroll
            for (int j = 0; j < CH; ++j) accB[8 * hh + j] += dzr[8 * hh + j];
          }
        }
        const float dzp = p.act == B2_ACT_SIGMOID ? g * av * (1.0f - av) : g;
        if (p.dz_prev) p.dz_prev[(size_t)(row0 + rr) * K + kk] = dzp;
        if (p.dz_planes) {
          // dzp = h1 + h2 + h3 exactly (three bf16 roundings of the running remainder), as split3_kernel writes them
          float r = dzp;
          unsigned short* pl = p.dz_planes + (size_t)(row0 + rr) * p.ld_planes + kk;
#pragma unroll
          for (int q = 0; q < 3; ++q) {
            const unsigned short h = __bfloat16_as_ushort(__float2bfloat16_rn(r));
            pl[(size_t)q * p.plane_stride] = h;
            r -= __uint_as_float((unsigned)h << 16);
          }
        }
        accD += dzp;                           // BiasAdd gradient of the layer below
      }
    }
  }

  // ---- per-CTA results ----
  if (p.train) {
    // the two row parities are added in a fixed order (even rows' sums + odd rows' sums)
    __syncthreads();
    float* sx = sa;                           // [128][17] + [17] scratch, reusing the activation tile
    if (half == 1) {
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        sx[kk * 17 + i] = accW[i];
        if (kk == 0) sx[128 * 17 + i] = accB[i];
      }
      sx[129 * 17 + kk] = accD;
    }
    __syncthreads();
    if (half == 0 && kk < K) {
      float* pw = p.partW + (size_t)blockIdx.x * p.part_stride;
      float* pb = p.partB + (size_t)blockIdx.x * p.part_stride;
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const int j = i & 7, c = (i >> 3) * CH + j;
        if (j < CH && c < C) {
          pw[kk * C + c] = accW[i] + sx[kk * 17 + i];
          if (kk == 0) pb[c] = accB[i] + sx[128 * 17 + i];
        }
      }
      if (p.partD) p.partD[(size_t)blockIdx.x * p.strideD + kk] = accD + sx[129 * 17 + kk];
    }
  }
  loss_acc = warp_sum(loss_acc);
  corr_acc = warp_sum(corr_acc);
  __syncthreads();
  if (lane == 0) { sred[warp] = loss_acc; sred[8 + warp] = corr_acc; }
  __syncthreads();
  if (tid == 0) {
    float l = 0.f, c = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) { l += sred[w]; c += sred[8 + w]; }
    p.loss_part[blockIdx.x] = (p.loss == B2_LOSS_SSE_SUM) ? l : l * p.inv_B;
    p.correct_part[blockIdx.x] = c;
  }
}

inline int smem_bytes(int K) {
  int act = (TR * (K + 1) + 3) & ~3;
  if (act < 129 * 17 + 128) act = (129 * 17 + 128 + 3) & ~3;     // the tile doubles as scratch for the final combine
  return (act + KMAX * 16 + TR * 16 + 16 + 4) * (int)sizeof(float);
}

}  // namespace nh

bool b2_narrow_head_applies(int K, int C) { return K >= 1 && K <= nh::KMAX && C >= 1 && C <= nh::CMAX; }

// number of CTAs (= number of partials written) for `rows` rows
int b2_narrow_head_grid(int rows, int sm_count) {
  // one tile per CTA while that stays within four CTAs' worth of partials per SM (three are resident at a time)
  const int tiles = (rows + nh::TR - 1) / nh::TR;
  const int cap = 4 * sm_count;
  return tiles < cap ? (tiles < 1 ? 1 : tiles) : cap;
}

int b2_launch_narrow_head(const float* a, const float* y, const float* W, const float* b, float* dz_prev, void* dz_planes,
                          long plane_stride, int ld_planes, float* partW,
                          float* partB, long part_stride, float* partD, long strideD, float* loss_part, float* correct_part, int rows, int K, int C,
                          int act, int loss, int train, float inv_B, int grid, cudaStream_t s) {
  if (!b2_narrow_head_applies(K, C)) { b2_set_error("narrow_head: unsupported shape %d -> %d", K, C); return B2_ERR_ARG; }
  const int smem = nh::smem_bytes(K);
  static unsigned long long attr_done = 0ull;
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  if (dev >= 64 || !((attr_done >> dev) & 1ull)) {
    const int mx = nh::smem_bytes(nh::KMAX);
    B2_CUDA(cudaFuncSetAttribute(nh::narrow_head_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
    B2_CUDA(cudaFuncSetAttribute(nh::narrow_head_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
    B2_CUDA(cudaFuncSetAttribute(nh::narrow_head_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
    B2_CUDA(cudaFuncSetAttribute(nh::narrow_head_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
    B2_CUDA(cudaFuncSetAttribute(nh::narrow_head_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
    B2_CUDA(cudaFuncSetAttribute(nh::narrow_head_kernel<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
    B2_CUDA(cudaFuncSetAttribute(nh::narrow_head_kernel<7>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
    B2_CUDA(cudaFuncSetAttribute(nh::narrow_head_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
    if (dev < 64) attr_done |= 1ull << dev;
  }
  nh::Args p;
  p.a = a; p.y = y; p.W = W; p.b = b; p.dz_prev = dz_prev; p.dz_planes = reinterpret_cast<unsigned short*>(dz_planes);
  p.plane_stride = plane_stride; p.ld_planes = ld_planes; p.partW = partW; p.partB = partB; p.part_stride = part_stride; p.partD = partD; p.strideD = strideD;
  p.loss_part = loss_part; p.correct_part = correct_part; p.rows = rows; p.K = K; p.C = C; p.act = act; p.loss = loss;
  p.train = train; p.inv_B = inv_B;
  switch ((C + 1) / 2) {
    case 1: nh::narrow_head_kernel<1><<<grid, nh::THREADS, smem, s>>>(p); break;
    case 2: nh::narrow_head_kernel<2><<<grid, nh::THREADS, smem, s>>>(p); break;
    case 3: nh::narrow_head_kernel<3><<<grid, nh::THREADS, smem, s>>>(p); break;
    case 4: nh::narrow_head_kernel<4><<<grid, nh::THREADS, smem, s>>>(p); break;
    case 5: nh::narrow_head_kernel<5><<<grid, nh::THREADS, smem, s>>>(p); break;
    case 6: nh::narrow_head_kernel<6><<<grid, nh::THREADS, smem, s>>>(p); break;
    case 7: nh::narrow_head_kernel<7><<<grid, nh::THREADS, smem, s>>>(p); break;
    default: nh::narrow_head_kernel<8><<<grid, nh::THREADS, smem, s>>>(p); break;
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { b2_set_error("narrow_head launch failed: %s", cudaGetErrorString(e)); return B2_ERR_CUDA; }
  return B2_OK;
}
