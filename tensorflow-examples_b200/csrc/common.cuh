// Shared declarations of libb2mlp: error plumbing, the context, and the launchers each
// translation unit exports to api.cu.  Not part of the public ABI (see include/b2mlp.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/b2mlp.h"

#define B2_MAX_LAYERS 8
#define B2_MAX_PEERS 16

// ---- error plumbing --------------------------------------------------------------------------
void b2_set_error(const char* fmt, ...);

#define B2_CUDA(expr)                                                                      \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess) {                                                               \
      b2_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return B2_ERR_CUDA;                                                                  \
    }                                                                                      \
  } while (0)

#define B2_CHECK_ARG(cond, ...)      \
  do {                               \
    if (!(cond)) {                   \
      b2_set_error(__VA_ARGS__);     \
      return B2_ERR_ARG;             \
    }                                \
  } while (0)

// ---- context ---------------------------------------------------------------------------------
struct b2_ctx {
  int device;
  int n_layers;
  int dims[B2_MAX_LAYERS + 1];
  int act, loss, precision;
  int max_local_batch;
  int sm_count;
  int opt_force_layered;

  long P;                          // trainable scalars
  long w_off[B2_MAX_LAYERS];       // offset of W[l] in the flat parameter / gradient layout
  long b_off[B2_MAX_LAYERS];       // offset of b[l]

  float* params;                   // [P]                      fp32 master parameters
  float* grads;                    // [P + 2]                  reduced gradients, then loss, then correct count
  float* act_buf[B2_MAX_LAYERS];   // act_buf[l] = output of layer l (l < n_layers-1), [max_local_batch, dims[l+1]]
  float* z_last;                   // [max_local_batch, dims[n_layers]] logits / predictions
  float* dz[2];                    // ping-pong [max_local_batch, max width]
  float* scratch;                  // split-batch partial gradients + per-block loss partials
  long scratch_floats;
  long part_off[B2_MAX_LAYERS];    // offset of layer l's partials inside scratch
  int splits[B2_MAX_LAYERS];       // number of batch splits of layer l's weight-gradient reduction
  long head_part_off;              // offset of the loss-head partials [2][head_blocks]
  int head_blocks;
  float* fused_scratch;            // fused single-layer path: [fused_grid][fused_stride]
  int fused_grid;
  long fused_stride;
  float* global_step_dev;          // float32 scalar, distributed.py:68-70
  float* eval_acc_dev;             // [2] accumulators for b2_eval
  float global_step_host;          // mirror advanced at enqueue time
  long launches;

  // staging for b2_session_run_train
  float* feed_x;                   // [max_local_batch, dims[0]]
  float* feed_y;                   // [max_local_batch, dims[n]]
  float* fetch_dev;                // [2] loss, correct (as int bits)
  float* fetch_host;               // pinned [2]
  cudaStream_t own_stream;

  // data parallel
  int rank, nranks;
  void* nccl_comm;
  // peer-memory exchange (optional, small models): my region of {value, stamp} cells (b2_fx_args) and the peers' mappings
  float* peer_local;               // cudaMalloc'ed, exported with cudaIpcGetMemHandle
  long peer_stride;                // cells per source in the region
  void* peer_base[B2_MAX_PEERS];   // mapped base of every rank's buffer (own entry = peer_local)
  int peer_ready;
  int opt_no_peer;                 // B2_OPT_NO_PEER_EXCHANGE
  int async_period;                // B2_OPT_ASYNC_PERIOD (0 = synchronous)
  long async_count;                // local steps since the context was created
  unsigned peer_step;              // exchanges enqueued so far
  unsigned long long* fx_prof;     // device [4]: exchange timeline sums (B2_FX_PROF=1), printed by b2_destroy
  // bucketed all-reduce of the wide layers' weight gradients, overlapped with the rest of the backward pass
  cudaStream_t comm_stream;        // created by b2_comm_init when nranks > 1
  cudaEvent_t ev_bucket, ev_comm_done;
  int opt_no_overlap;              // B2_OPT_NO_BUCKET_OVERLAP

  // bf16-operand / fp32-accumulate path (precision == B2_PREC_BF16_FP32ACC): tensor-core GEMMs for the hidden layers
  void* xb;                        // [max_local_batch, dims[0]] bf16 copy of the step's inputs
  void* ab[B2_MAX_LAYERS];         // bf16 activations of the hidden layers
  void* dzb[2];                    // bf16 gradients w.r.t. hidden pre-activations (ping-pong)
  void* wb[B2_MAX_LAYERS];         // bf16 copies of the hidden layers' weights, refreshed by the SGD apply
  float* colsum_part;              // bias-gradient partials of the hidden layers
  void* dzn;                       // bf16 [max_local_batch, 64]: the output layer's dz, zero padded
  int wb_stale;                    // parameters were set from outside: refresh the bf16 copies first

  // fp32 layers on the tensor cores (gemm_split3.cu): three bf16 planes of a wide layer's weights and dz (the layer
  // input is split inside the GEMM kernel)
  void* s3_w[B2_MAX_LAYERS];       // [3][dims[l], round8(dims[l+1])]; NULL = this layer stays on the CUDA cores
  void* s3_dz;                     // [3][max_local_batch, round8(widest eligible dims[l+1])]
  int opt_no_split3;               // B2_OPT_NO_SPLIT3
  int opt_no_narrow_head;          // B2_OPT_NO_NARROW_HEAD
  int opt_no_persistent;           // B2_OPT_NO_PERSISTENT
  int opt_multistep_fused;         // B2_OPT_MULTISTEP_FUSED
  int opt_sharded;                 // B2_OPT_SHARDED_APPLY
  float* lc_dev;                   // [4] loss / correct on the side of a sharded update
  unsigned* grid_bar;              // [2] grid-barrier counters of the persistent multi-step kernels
  float* cs_scratch;               // exchange buffers of the cluster kernel (cluster_steps.cu), when the model qualifies
  int head_used;                   // loss / correct partials written by the last forward (<= head_blocks)
  long head_db_off;                // scratch offset of the row-thread loss head's bias-gradient partials [head_blocks][16]
  long nh_db_off;                  // scratch offset of the narrow-head kernel's bias-gradient partials for layer L-1, or -1
  int nh_planes_used;              // ... and dz of layer L-1 as bf16 planes in s3_dz
  int nh_db_used;                  // the last backward pass left layer L-1's bias gradient there (one partial per CTA)

  // optional per-kernel event timing (B2_OPT_KERNEL_TIMING)
  int opt_timing;
  struct b2_timed* timed;
  int n_timed, cap_timed;
};
struct b2_timed { cudaEvent_t a, b; int cls; };

// ---- launchers (kernels_layered.cu) --------------------------------------------------------------
// dA[M,Kin] = (dZ[M,N] @ W[Kin,N]^T) * act'(Aout)
int b2_launch_dgrad(const float* dZ, const float* W, const float* Aout, float* dA, int M, int Kin,
                    int N, int act, cudaStream_t s);
// loss head: z -> dz, per-block loss / correct partials
int b2_launch_loss_head(const float* z, const float* y, float* dz, float* loss_part,
                        float* correct_part, int rows, int C, int loss, float inv_global_batch,
                        int blocks, cudaStream_t s);
// sums the split partials in fixed order into grads (+loss, +correct); when apply != 0 also does
// the SGD update, the trace writes and global_step += 1.
struct b2_reduce_args {
  int n_layers;
  long w_off[B2_MAX_LAYERS], b_off[B2_MAX_LAYERS];
  long w_size[B2_MAX_LAYERS], b_size[B2_MAX_LAYERS];
  long part_off[B2_MAX_LAYERS];
  long pstride[B2_MAX_LAYERS];   // floats between two partials of a layer
  int splits[B2_MAX_LAYERS];
  long P;
  long head_part_off;   // loss partials
  long head_corr_off;   // correct-count partials
  int head_blocks;      // how many of each
};
struct b2_fx_args;
// fx != NULL: every element is exchanged with the peers (b2_fx_args below) before the update is applied
int b2_launch_reduce(const b2_reduce_args& a, const float* scratch, float* grads, float* params,
                     int apply, float lr, float* loss_trace, int* correct_trace,
                     float* global_step, int sm_count, const b2_fx_args* fx, cudaStream_t s);
// params -= lr * grads after the all-reduce; trace writes; global_step += 1
int b2_launch_sgd_apply(float* params, const float* grads, long P, float lr, float* loss_trace,
                        int* correct_trace, float* global_step, int sm_count, cudaStream_t s);
// sharded update of the parameters one rank owns (after a reduce-scatter of the gradients); lc = {loss, correct} over all ranks
int b2_launch_sharded_apply(float* params_shard, const float* grads_shard, long n, float lr, const float* lc, float* loss_trace,
                            int* correct_trace, float* global_step, int sm_count, cudaStream_t s);
// eval accumulation: acc[0] += loss partials, acc[1] += correct partials (fixed order)
int b2_launch_eval_accumulate(const float* loss_part, const float* correct_part, int blocks,
                              float* acc, cudaStream_t s);
int b2_launch_eval_finish(const float* acc, float* loss_out, long* correct_out, cudaStream_t s);
int b2_launch_scale(float* p, long n, float a, int sm_count, cudaStream_t s);   // p[i] *= a

// ---- launchers (fused_linear.cu): single dense layer 784 -> 10 + softmax-CE head, one kernel --------
bool b2_fused_linear_supported(int in_dim, int out_dim, int loss);
int b2_fused_linear_grid(int sm_count);
long b2_fused_linear_stride(void);
int b2_fused_linear_smem_bytes(void);
int b2_fused_linear_tile_rows(void);
int b2_launch_fused_linear_step(const float* x, const float* y, int rows, const float* W,
                                const float* b, float* scratch, int grid, float inv_global_batch, int loss,
                                cudaStream_t s);
int b2_launch_fused_linear_reduce(const float* scratch, int grid, float* grads, float* params,
                                  int apply, float lr, float* loss_trace, int* correct_trace,
                                  float* global_step, long P, cudaStream_t s);
int b2_fused_linear_prepare(void);

// ---- gradient exchange over NVLink peer memory, fused into the kernel that finishes the local reduction -------------
// Every rank exports a region of 8-byte cells {float value, uint32 step stamp}:
//   cells [2 step parities][B2_MAX_PEERS sources][stride]
// The thread that owns gradient element i PUSHES {its locally reduced value, step} into cell [parity][my rank][i] of
// every peer's region (one aligned 8-byte remote store), then polls its own cells [parity][p][i] in local memory until
// their stamp is this step's and adds the N contributions in rank order (the same sum on every rank), applies SGD.
// The cell is written and read with ONE .b64 access (single-copy atomic), so there is no separate flag and no fence
// between data and flag; the critical path is one NVLink one-way trip.  A rank can be at most one step ahead of a peer (it needs the peer's
// values of step s to finish step s), so slot s & 1 is never overwritten before it was read.
#define B2_FX_SLOT 7872            /* fused 784->10 kernel: 246 blocks x 32 outputs >= 7852 */
#define B2_FX_BLOCKS 246
struct b2_fx_args {
  int rank, nranks;
  long stride;                                // cells per source
  unsigned step1;                             // stamp of this step (exchanges so far + 1)
  int parity;
  unsigned long long* region[B2_MAX_PEERS];   // every rank's region (own entry = local memory, the others = peer mappings)
  unsigned long long* prof;                   // diagnostics (B2_FX_PROF=1): [ns issuing pushes, ns waiting, exchanges] of element 0
};
#ifdef __CUDACC__
// device side of the exchange: push my value of element i, wait for the peers' values of the same element, return the
// sum in rank order.  Called by exactly one thread per element.
// (step1, parity) are passed separately so that a kernel that runs several steps can advance them itself.  The two halves
// are also available on their own (b2_fx_push / b2_fx_collect), for a thread that owns several elements: push them all,
// then collect them all -- the NVLink trips of the elements overlap.
__device__ __forceinline__ void b2_fx_push(const b2_fx_args& a, unsigned step1, int parity, long i, float mine) {
  const size_t slot = ((size_t)parity * B2_MAX_PEERS) * (size_t)a.stride;
  // one 64-bit access per cell: {value (low word), stamp (high word)}.  A .b64 access is single-copy atomic by the PTX
  // memory model, so a reader can never pair a new stamp with an old value (a .v2.u32 access gives no such guarantee)
  const unsigned long long cell = ((unsigned long long)step1 << 32) | (unsigned long long)__float_as_uint(mine);
  for (int p = 0; p < a.nranks; ++p)
    if (p != a.rank)
      asm volatile("st.relaxed.sys.global.b64 [%0], %1;" ::"l"(a.region[p] + slot + (size_t)a.rank * a.stride + i), "l"(cell)
                   : "memory");
}
// waits for the peers' values of element i and returns the sum over all ranks in rank order (the same sum on every rank)
__device__ __forceinline__ float b2_fx_collect(const b2_fx_args& a, unsigned step1, int parity, long i, float mine) {
  const size_t slot = ((size_t)parity * B2_MAX_PEERS) * (size_t)a.stride;
  // poll ALL outstanding peers' cells in each round (independent loads, issued back to back): the wait costs one round
  // trip to local memory after the last peer's store has landed, not one per peer
  const unsigned long long* cells = a.region[a.rank] + slot + i;
  unsigned pending = ((a.nranks >= 32 ? 0u : (1u << a.nranks)) - 1u) & ~(1u << a.rank);
  float vals[B2_MAX_PEERS];
  unsigned spins = 0;
  while (pending) {
    unsigned long long got[B2_MAX_PEERS];
#pragma unroll
    for (int p = 0; p < B2_MAX_PEERS; ++p)
      if ((pending >> p) & 1u)
        asm volatile("ld.relaxed.sys.global.b64 %0, [%1];" : "=l"(got[p]) : "l"(cells + (size_t)p * a.stride) : "memory");
#pragma unroll
    for (int p = 0; p < B2_MAX_PEERS; ++p)
      if (((pending >> p) & 1u) && (unsigned)(got[p] >> 32) == step1) {
        vals[p] = __uint_as_float((unsigned)got[p]);
        pending &= ~(1u << p);
      }
    if (++spins > (1u << 26)) __trap();   // a dead peer must not hang this GPU for ever
  }
  float sum = 0.f;
#pragma unroll
  for (int p = 0; p < B2_MAX_PEERS; ++p)
    if (p < a.nranks) sum += (p == a.rank) ? mine : vals[p];        // rank order: the same sum on every rank
  return sum;
}
__device__ __forceinline__ float b2_fx_exchange_at(const b2_fx_args& a, unsigned step1, int parity, long i, float mine) {
  const bool prof = a.prof != nullptr && i == 0;
  unsigned long long t0 = 0, t1 = 0;
  if (prof) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
  b2_fx_push(a, step1, parity, i, mine);
  if (prof) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
  const float sum = b2_fx_collect(a, step1, parity, i, mine);
  if (prof) {
    unsigned long long t2;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t2));
    a.prof[0] += t1 - t0;      // issuing the pushes
    a.prof[1] += t2 - t1;      // waiting for the last peer (NVLink latency + rank skew)
    a.prof[2] += 1;
  }
  return sum;
}
__device__ __forceinline__ float b2_fx_exchange(const b2_fx_args& a, long i, float mine) {
  return b2_fx_exchange_at(a, a.step1, a.parity, i, mine);
}
#endif
int b2_launch_fused_linear_reduce_exchange(const float* scratch, int grid, const b2_fx_args& a, float* grads, float* params,
                                           float lr, float* loss_trace, int* correct_trace, float* global_step, long P,
                                           cudaStream_t s);

// ---- persistent multi-step kernels: n_steps training steps per launch ------------------------------------------
// fused_linear_ms.cu: the 784 -> 10 step kernel with the step boundary (partial reduction, [exchange], SGD apply, weight
// reload) inside the kernel, behind two grid barriers; `bar` = 2 counters in device memory, `grid` <= SM count
int b2_fused_linear_ms_prepare(void);
int b2_launch_fused_linear_multistep(const float* x_pool, const float* y_pool, long span, long first_row, long global_batch,
                                     long rank_off, int rows, int n_steps, float* params, float* grads, float* scratch,
                                     unsigned* bar, int grid, float lr, int loss, float* loss_trace, int* correct_trace,
                                     float* global_step, long P, const b2_fx_args* fx, cudaStream_t s);
// cluster_steps.cu: small batches of the 784-wide models (784 -> 10 up to 128 rows, 784-100-10 up to 200 rows) on one
// cluster of 14 CTAs: W1 sliced over the CTAs' shared memory, the per-step exchanges through an L2-resident scratch
// buffer between two cluster barriers
bool b2_cluster_steps_applies(int n_layers, const int* dims, int rows);
int b2_launch_cluster_steps(int n_layers, int act, int loss, const long* w_off, const long* b_off, long P, const float* x_pool,
                            const float* y_pool, long span, long first_row, long global_batch, int rows, int n_steps, float lr,
                            float* params, float* grads, float* loss_trace, int* correct_trace, float* global_step,
                            float* scratch, long rank_off, const b2_fx_args* fx, cudaStream_t s);
long b2_cluster_steps_scratch_floats(int n_layers);
// tiny_steps.cu: whole runs of tiny models (<= 4 layers, widths <= 16, <= 32 rows per step) by one warp
bool b2_tiny_steps_applies(int n_layers, const int* dims, int rows);
int b2_launch_tiny_steps(int n_layers, const int* dims, int act, int loss, const long* w_off, const long* b_off, long P,
                         const float* x_pool, const float* y_pool, long span, long first_row, long global_batch, int rows,
                         int n_steps, float lr, float* params, float* grads, float* loss_trace, int* correct_trace,
                         float* global_step, cudaStream_t s);

// ---- launcher (gemm_tcgen05.cu): bf16 tensor-core GEMM with fused epilogues ------------------------------
// local_operands: the caller vouches that A and B lie in THIS GPU's memory (the context's own buffers).  Only then may
// CTA pairs share a B tile by TMA multicast; TMA multicast loads here read only this GPU's memory, since on B200 a TMA
// multicast load from another GPU's memory hung the GPU.  The public b2_matmul_bf16, whose pointers come from outside,
// passes 0 and every CTA fetches its own tiles.
int b2_launch_gemm_bf16(const void* A, int a_mn, const void* B, int b_mn, int M, int N, int K, int epi,
                        const float* bias, const void* act, void* out, int sm_count, cudaStream_t s, int local_operands = 1);
int b2_launch_gemm_bf16_ex(const void* A, int a_mn, const void* B, int b_mn, int M, int N, int K, int epi,
                           const float* bias, const void* act, void* out, int ldc, int n_store, int ksplits,
                           long split_stride, float* colsum_part, int sm_count, cudaStream_t s, int local_operands = 1);
// second stage of the bias gradients whose first stage is fused into the dgrad epilogue: out[c] = sum_z part[z][c]
// narrow_head.cu: output layer (<= 128 inputs, <= 16 classes) forward + loss head + backward in one kernel
bool b2_narrow_head_applies(int K, int C);
int b2_narrow_head_grid(int rows, int sm_count);
int b2_launch_narrow_head(const float* a, const float* y, const float* W, const float* b, float* dz_prev, void* dz_planes,
                          long plane_stride, int ld_planes, float* partW,
                          float* partB, long part_stride, float* partD, long strideD, float* loss_part, float* correct_part, int rows, int K, int C,
                          int act, int loss, int train, float inv_B, int grid, cudaStream_t s);
// gemm_split3.cu: fp32-accurate GEMM as six bf16 tensor-core products over three-plane splits of the operands
int b2_launch_split3(const float* in, void* planes, long rows, int cols, int ld_in, int ld_out, long plane_stride,
                     int sm_count, cudaStream_t s);
int b2_launch_colsum_f32(const float* dz, float* partB, long part_stride, int rows, int N, int slices, cudaStream_t s);
void b2_gemm_split3_plan(int M, int N, int K, int max_splits, int sm_count, int* block_q, int* splits);
int b2_launch_gemm_split3(const float* A, int a_mn, long lda, const void* B, int b_mn, long ldb, long b_plane,
                          int M, int N, int K, int epi, int act, const float* bias, float* out, int ldc, int block_q,
                          int ksplits, long split_stride, int sm_count, cudaStream_t s);
int b2_launch_colsum_finish(const float* part, float* out, int N, int slices, cudaStream_t s);

// ---- launchers (mixed_aux.cu): CUDA-core kernels of the bf16 path ------------------------------------------
int b2_launch_cast_bf16(const float* src, void* dst, long n, int sm_count, cudaStream_t s);
// fp32 [rows, C] -> bf16 [rows, 64], zero padded: makes the narrow output layer a legal tensor-core operand
int b2_launch_pad_cast64(const float* src, void* dst, long rows, int C, int sm_count, cudaStream_t s);
// bias gradient of the output layer: per-slice column sums of fp32 dz [rows, C] -> partB[slice*stride + c]
// loss head, one thread per row (<= 16 classes): dz as padded bf16 [rows,64], loss / correct / db partials per block
int b2_launch_loss_head_rows(const float* z, const float* y, void* dzn, float* loss_part, float* correct_part, float* db_part,
                              int rows, int C, int loss, float inv_global_batch, int blocks, cudaStream_t s);
int b2_narrow_dgrad_slices(int rows);
int b2_launch_narrow_dgrad_bf16(const void* dzn, const void* wpad, const void* act, void* out, float* colsum_part, int rows,
                                int K, int C, cudaStream_t s);
int b2_colsum_slices(int rows, int N, int sm_count);
int b2_launch_colsum_bf16(const void* m, float* part, float* out, int rows, int N, int slices, cudaStream_t s);
int b2_launch_sgd_apply_mixed(float* params, const float* grads, long P, float lr, int n_seg, const long* off,
                              const long* size, void* const* wb, float* loss_trace, int* correct_trace,
                              float* global_step, int sm_count, cudaStream_t s);

// ---- launchers (sgemm_ffma2.cu): fp32 FFMA2 GEMM micro-kernel for the generic layer path --------------------
int b2_launch_sgemm_fwd(const float* A, const float* W, const float* bias, float* C, int M, int N, int K, int act,
                        cudaStream_t s);
int b2_launch_sgemm_wgrad(const float* A, const float* dZ, float* partW, float* partB, long part_stride, int rows,
                          int Kin, int N, int splits, cudaStream_t s);
// forward for small batches (M <= 256): column-sliced, K split over the lanes of a warp
bool b2_small_m_fwd_applies(int M, int K);
int b2_launch_small_m_fwd(const float* A, const float* W, const float* bias, float* C, int M, int N, int K, int act,
                          cudaStream_t s);
