// Microbenchmark: issue rate of fma.rn.f32x2 (FFMA2) on one SM as a function of warps per sub-partition, independent
// accumulators per thread and operand form (vector x vector, or scalar-broadcast as pack2(x, x)).  Prints SM clocks per
// warp-level FFMA2 per sub-partition (2.0 = the packed-FMA pipe is saturated).
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 pack2(float lo, float hi) { u64 d; asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi)); return d; }

template <int NACC, int MODE>
__global__ void k(float* out, long long* clk, int iters, float seed) {
  u64 acc[NACC];
  for (int i = 0; i < NACC; ++i) acc[i] = pack2(seed + i, seed - i);
  float xs[4] = {seed, seed * 0.5f, seed * 0.25f, seed * 0.125f};
  u64 w[4] = {pack2(1.0f + seed, 0.5f), pack2(0.25f, seed), pack2(seed, seed), pack2(0.75f, 1.5f)};
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
      for (int i = 0; i < NACC; ++i) {
        if (MODE == 0) acc[i] = ffma2(w[(i + kk) & 3], w[kk], acc[i]);                       // vector x vector
        else if (MODE == 1) acc[i] = ffma2(pack2(xs[kk], xs[kk]), w[i & 3], acc[i]);         // scalar-broadcast x vector
        else { float lo, hi; asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc[i]));   // two scalar FMAs instead
               lo = fmaf(xs[kk], seed, lo); hi = fmaf(xs[kk], 0.5f, hi); acc[i] = pack2(lo, hi); }
      }
    }
  }
  long long t1 = clock64();
  __syncthreads();
  float s = 0.f;
  for (int i = 0; i < NACC; ++i) { float lo, hi; asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc[i])); s += lo + hi; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

template <int NACC, int MODE>
void run(const char* name) {
  float* out; long long* clk;
  cudaMalloc(&out, 2048 * sizeof(float)); cudaMalloc(&clk, 8 * sizeof(long long));
  const int iters = 2000;
  for (int warps = 4; warps <= 32; warps *= 2) {
    k<NACC, MODE><<<1, warps * 32>>>(out, clk, iters, 1.0001f);
    cudaDeviceSynchronize();
    k<NACC, MODE><<<1, warps * 32>>>(out, clk, iters, 1.0001f);
    cudaDeviceSynchronize();
    long long c; cudaMemcpy(&c, clk, sizeof(c), cudaMemcpyDeviceToHost);
    const double per_smsp = (double)(warps / 4) * iters * 4 * NACC * (MODE == 2 ? 2 : 1);
    printf("%-34s acc/thread %2d warps/SMSP %d : %.2f clk per warp-instr per SMSP\n", name, NACC, warps / 4, c / per_smsp);
  }
  cudaFree(out); cudaFree(clk);
}

int main() {
  run<20, 0>("FFMA2 vector x vector");
  run<20, 1>("FFMA2 scalar-broadcast x vector");
  run<10, 1>("FFMA2 scalar-broadcast x vector");
  run<4, 1>("FFMA2 scalar-broadcast x vector");
  run<20, 2>("FFMA (two scalar FMAs)");
  cudaError_t e = cudaGetLastError();
  printf("status: %s\n", cudaGetErrorString(e));
  return 0;
}
