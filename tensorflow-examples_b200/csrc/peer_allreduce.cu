// Gradient exchange fused with the SGD apply, over NVLink peer memory (one process per GPU).
//
// The reference pushes every gradient to the parameter server and pulls every variable back each
// step (distributed.py:63-65, 2 x 318 KB of gRPC per worker step).  Here each rank leaves its reduced
// gradient vector [grads..., loss, correct] in a buffer the other ranks have mapped
// (cudaIpcOpenMemHandle), raises a step flag, and ONE kernel per rank then
//   waits for every peer's flag (system-scope acquire loads over NVLink),
//   reads all N vectors through the peer mappings, adds them in rank order (so every rank forms the
//   bit-identical sum), and applies var -= lr * grad in the same pass.
// No NCCL call, no separate apply kernel: for the small models this exchange is pure latency and the
// one-shot form needs a single NVLink round trip.  Two buffers alternate by step parity so a fast rank
// can start writing step s+1 while a slow one still reads step s.
#include "common.cuh"

namespace {

__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ float ld_peer(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(256)
peer_allreduce_apply_kernel(b2_peer_args a, float* __restrict__ params, float* __restrict__ grads_out, long P,
                            unsigned step1, int buf, float lr, float* __restrict__ loss_trace,
                            int* __restrict__ correct_trace, float* __restrict__ global_step) {
  // programmatic dependent launch: the next step kernel may start its prologue and its x loads now; this kernel
  // itself may have been started early and waits here for the reduction kernel that wrote my vector
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  // publish: my vector for this step was written by the previous kernel on this stream
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    __threadfence_system();
    st_release_sys(a.flag[a.rank], step1);
  }
  // wait for the peers (own data is ordered by the stream)
  if (threadIdx.x < a.nranks && threadIdx.x != a.rank) {
    const unsigned* f = a.flag[threadIdx.x];
    unsigned spins = 0;
    while ((int)(ld_acquire_sys(f) - step1) < 0) {
      __nanosleep(64);
      if (++spins > (1u << 24)) __trap();   // a dead peer must not hang this GPU for ever
    }
  }
  __syncthreads();
  const long n = P + 2;
  const long off = (long)buf * a.stride;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    float s = 0.f;
    for (int p = 0; p < a.nranks; ++p) s += ld_peer(a.vec[p] + off + i);   // rank order: same sum everywhere
    grads_out[i] = s;
    if (i < P) {
      params[i] -= lr * s;                                                  // distributed.py:107-108
    } else if (i == P) {
      if (loss_trace) *loss_trace = s;
    } else {
      if (correct_trace) *correct_trace = (int)(s + 0.5f);
      *global_step += 1.0f;
    }
  }
}

}  // namespace

int b2_launch_peer_allreduce_apply(const b2_peer_args& a, float* params, float* grads_out, long P, unsigned step1,
                                   int buf, float lr, float* loss_trace, int* correct_trace, float* global_step,
                                   int sm_count, cudaStream_t s) {
  long blocks = (P + 2 + 255) / 256;
  if (blocks > 2L * sm_count) blocks = 2L * sm_count;   // all blocks resident: they spin on remote flags
  if (blocks < 1) blocks = 1;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)blocks);
  cfg.blockDim = dim3(256);
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, peer_allreduce_apply_kernel, a, params, grads_out, P, step1, buf, lr, loss_trace,
                                     correct_trace, global_step);
  if (e != cudaSuccess) {
    b2_set_error("peer_allreduce_apply launch failed: %s", cudaGetErrorString(e));
    return B2_ERR_CUDA;
  }
  return B2_OK;
}
