/*
 * b2mlp.h -- C ABI of the B200-native training-step library (libb2mlp.so).
 *
 * This is the drop-in boundary for the ONE hot path of manigoswami/tensorflow-examples:
 * the `sess.run(train_op, feed_dict)` call inside the Python training loops of
 *   simple/simple.py:34-35,37          (scalar linear regression, sum-of-squares loss)
 *   distributed/distributed.py:145-150 (784-100-10 sigmoid MLP, softmax cross-entropy)
 * and, for the distributed form, the parameter-server pull/push that
 * `replica_device_setter` (distributed/distributed.py:63-65) implies, which is
 * replaced here by a synchronous gradient all-reduce before the SGD apply.
 *
 * The reference has no FFI of its own (it is pure Python on top of TensorFlow 1.x);
 * each entry point below cites the reference lines whose work it takes over.
 * Plain pointers and sizes only: no torch, no C++ types.  Every function that can
 * fail returns an int: 0 = ok, non-zero = error class (B2_ERR_*); the message is
 * available from b2_last_error() on the calling thread.  No C++ exception crosses
 * this boundary.  There is NO CPU fallback: without a CUDA device every compute
 * entry point fails with B2_ERR_CUDA.
 *
 * Memory layouts (all row-major, all float32 unless stated):
 *   x pool   [pool_rows, dims[0]]        distributed.py:77  placeholder x  [None,784]
 *   y pool   [pool_rows, dims[n_layers]] distributed.py:80  placeholder y_ [None,10] (one-hot or soft)
 *   W[l]     [dims[l], dims[l+1]]        distributed.py:85-86 ([in,out], as tf.matmul(x, W) wants)
 *   b[l]     [dims[l+1]]                 distributed.py:90-91
 * Threading: one context per GPU, used from one host thread at a time.
 */
#ifndef B2MLP_H_
#define B2MLP_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b2_ctx b2_ctx;          /* opaque */
typedef void* b2_stream_t;             /* a cudaStream_t (0 = legacy default stream) */

enum {
  B2_OK = 0,
  B2_ERR_ARG = 1,        /* bad argument / unsupported configuration */
  B2_ERR_CUDA = 2,       /* CUDA runtime error (no device, launch failure, OOM ...) */
  B2_ERR_NCCL = 3,       /* NCCL error or NCCL library not found */
  B2_ERR_STATE = 4       /* call not valid in the context's current state */
};

enum { B2_ACT_IDENTITY = 0, B2_ACT_SIGMOID = 1 };          /* simple.py:16 / distributed.py:97 */
enum { B2_LOSS_SSE_SUM = 0, B2_LOSS_SOFTMAX_CE_MEAN = 1 }; /* simple.py:20 / distributed.py:99-102 */
enum { B2_PREC_FP32 = 0, B2_PREC_BF16_FP32ACC = 1 };       /* BASELINE.json configs[4] uses 1 */

/* Engine selection knobs for b2_set_option (tests cross-check the paths against each other). */
enum {
  B2_OPT_FORCE_LAYERED = 1,  /* value 1: never use the single-kernel fused step, even when the shape allows it */
  B2_OPT_KERNEL_TIMING = 2,  /* value 1: bracket every kernel launch with CUDA events (see b2_kernel_times) */
  B2_OPT_NO_PEER_EXCHANGE = 3, /* value 1: use the NCCL all-reduce even when peer memory is mapped (b2_peer_import) */
  B2_OPT_NO_BUCKET_OVERLAP = 4, /* value 1 (bf16 path, nranks > 1): one all-reduce after the backward pass instead of
                                 per-layer buckets on a second stream that overlap the remaining backward GEMMs */
  B2_OPT_NO_SPLIT3 = 6,      /* value 1 (fp32): keep the wide layers' GEMMs on the CUDA cores instead of the fp32-accurate
                                 three-plane bf16 split on the tensor cores (gemm_split3.cu) */
  B2_OPT_NO_NARROW_HEAD = 7, /* value 1 (fp32, >= 2 layers): run the output layer as separate forward / loss-head / wgrad /
                                 dgrad kernels instead of the single narrow-head kernel (narrow_head.cu) */
  B2_OPT_NO_PERSISTENT = 8,  /* value 1: b2_train_steps enqueues every step as separate kernel launches instead of running
                                 all n_steps inside one persistent kernel where such a kernel exists */
  B2_OPT_MULTISTEP_FUSED = 9, /* value 1: run the 784 -> 10 model's steps inside ONE launch of the fused step kernel with the
                                 step boundary (reduction, exchange, apply) behind in-kernel grid barriers.  Bit-identical to
                                 the default (one step kernel + one reduction kernel per step, chained by programmatic
                                 dependent launch) and measured SLOWER than it at every batch size, so it is off by default */
  B2_OPT_SHARDED_APPLY = 10, /* value 1 (nranks > 1, NCCL exchange): the B200 analogue of SEVERAL parameter servers
                                 (distributed.py:63-65: replica_device_setter spreads the variables over the ps tasks):
                                 rank r owns slice r of the flat parameter vector -- reduce-scatter of the gradients, SGD apply
                                 on the own slice only, all-gather of the updated parameters -- instead of an all-reduce and
                                 a full apply on every rank.  Same update, replicas stay bit-identical */
  B2_OPT_ASYNC_PERIOD = 5     /* value k > 0: emulate the reference's ASYNCHRONOUS parameter-server training
                                 (distributed.py:107-113: plain GradientDescentOptimizer, SyncReplicas commented out):
                                 every worker applies its own gradients locally (mean over ITS batch) and the replicas'
                                 parameters are averaged every k steps; k = 1 reproduces the synchronous update.
                                 value 0 (default): synchronous all-reduce data parallelism */
};

/* Kernel classes reported by b2_kernel_times. */
enum {
  B2_K_DENSE_FWD = 0,     /* matmul + bias + activation            distributed.py:96-98 */
  B2_K_LOSS_HEAD = 1,     /* softmax-CE / SSE forward + backward   distributed.py:99-102,117-118; simple.py:20 */
  B2_K_WGRAD = 2,         /* dW = A^T dZ, db = colsum(dZ)          TF1 MatMul / BiasAdd gradients */
  B2_K_DGRAD = 3,         /* dA = dZ W^T * act'                    TF1 MatMul / Sigmoid gradients */
  B2_K_REDUCE_APPLY = 4,  /* fixed-order partial sums (+ SGD)      distributed.py:107-108 */
  B2_K_FUSED_STEP = 5,    /* single-kernel 784->10 step */
  B2_K_FUSED_REDUCE = 6,  /* its partial reduction (+ SGD) */
  B2_K_ALLREDUCE_APPLY = 7, /* gradient all-reduce + SGD apply (data parallel only) */
  B2_K_GEMM_FWD = 8,      /* tcgen05 GEMM: act(A W + b)            bf16 path, hidden layers */
  B2_K_GEMM_DGRAD = 9,    /* tcgen05 GEMM: dZ W^T * act'           bf16 path */
  B2_K_GEMM_WGRAD = 10,   /* tcgen05 GEMM: A^T dZ                  bf16 path */
  B2_K_MIXED_AUX = 11,    /* casts, narrow output layer, bias column sums, SGD apply of the bf16 path */
  B2_K_PERSISTENT_STEPS = 12, /* one launch = ALL steps of a b2_train_steps call (simple.py:34-35, distributed.py:145-150) */
  B2_KERNEL_CLASSES = 13
};

#define B2_UNIQUE_ID_BYTES 128
#define B2_PEER_HANDLE_BYTES 64

/* Library / build information: "b2mlp <version> sm_100a ..." (static string). */
const char* b2_version(void);

/* Message of the last failing call made by this thread (static thread-local buffer). */
const char* b2_last_error(void);

/*
 * Builds the model + optimizer state on one GPU: parameters, gradient buffer, activations and
 * scratch are allocated here, once.  Replaces graph construction + variable creation
 * (simple.py:11-23; distributed.py:68-108) and tf.Session()/Supervisor set-up
 * (simple.py:28-32; distributed.py:129-135).  Parameters start at zero; load them with
 * b2_params_set().  n_layers = number of dense layers, dims has n_layers+1 entries.
 * `act` applies to every layer but the last.  max_local_batch bounds the rows per step on this GPU.
 */
int b2_create(b2_ctx** out, int device, int n_layers, const int* dims, int act, int loss,
              int precision, int max_local_batch);
void b2_destroy(b2_ctx* ctx);

int b2_set_option(b2_ctx* ctx, int option, int value);

/* Total number of trainable scalars (sum over layers of dims[l]*dims[l+1] + dims[l+1]). */
long b2_param_count(const b2_ctx* ctx);

/*
 * Copy one layer's weights/bias from/to DEVICE buffers owned by the caller (stream-ordered).
 * Replaces the init_op run by the chief (distributed.py:126,129-135; simple.py:28,32) and the
 * fetch of W, b (simple.py:37).  Either pointer may be NULL to skip it.
 */
int b2_params_set(b2_ctx* ctx, int layer, const float* W_dev, const float* b_dev, b2_stream_t stream);
int b2_params_get(b2_ctx* ctx, int layer, float* W_dev, float* b_dev, b2_stream_t stream);

/*
 * Data-parallel set-up, replacing tf.train.ClusterSpec / tf.train.Server / replica_device_setter
 * (distributed.py:37-43,63-65): rank = worker task_index, nranks = number of worker hosts.
 * b2_get_unique_id fills a 128-byte NCCL unique id on one rank; every rank then passes the same
 * bytes to b2_comm_init.  With nranks == 1 no communicator is created.
 */
int b2_get_unique_id(void* out128);
int b2_comm_init(b2_ctx* ctx, int rank, int nranks, const void* unique_id128);

/*
 * Optional fast path of the gradient exchange for models whose parameter vector is small (<= 256 Ki
 * floats): instead of calling NCCL, the kernel that finishes a rank's local gradient reduction pushes every
 * element, tagged with the step number, straight into buffers the other ranks of the box have mapped over
 * NVLink, waits for the peers' values of the same element in its own memory, adds them in rank order and
 * applies the SGD update -- no separate collective, no extra kernel.  This is what replaces the reference's
 * per-step variable pull + gradient push through the parameter server (distributed.py:63-65) on an NVSwitch box.
 * Call after b2_comm_init: b2_peer_export writes a 64-byte CUDA IPC handle; gather the handles of all
 * ranks (rank order) and pass the nranks*64 bytes to b2_peer_import on every rank.  All ranks must be
 * processes on one node with peer access between their GPUs, one GPU per rank; otherwise b2_peer_import fails
 * and the NCCL path stays in use (b2_set_option(B2_OPT_NO_PEER_EXCHANGE, 1) on EVERY rank once any rank failed:
 * the ranks must take the same path).  The exchange kernels wait for a peer's values by polling and give up
 * (device trap -> the next call returns B2_ERR_CUDA) after 2^26 polls, a few seconds: do not let two ranks share
 * one GPU and do not stall one rank (debugger, long host-side pause between enqueued steps) while others run.
 */
int b2_peer_export(b2_ctx* ctx, void* handle_out64);
int b2_peer_import(b2_ctx* ctx, const void* handles);

/*
 * n_steps consecutive training steps on device-resident data; nothing is copied to or from the
 * host and the call returns as soon as the work is enqueued on `stream`.
 * Replaces n_steps iterations of the hot loops (simple.py:34-35; distributed.py:145-150):
 * forward, loss, accuracy, backward, [all-reduce], SGD apply, global_step += 1.
 *   step s uses global rows  start_s = (first_row + s*global_batch) mod (floor(pool_rows/global_batch)*global_batch)
 *   this rank takes rows     start_s + rank*local_batch ... + local_batch   (local_batch*nranks == global_batch)
 *   the loss head's mean divides by global_batch (distributed.py:102 reduce_mean over the whole batch).
 * loss_trace_dev[s]    <- the PRE-update loss of step s over the global batch (what the reference fetches)
 * correct_trace_dev[s] <- number of rows of the global batch whose argmax matches the label's (distributed.py:117)
 * Either trace pointer may be NULL.  x_pool_dev must be 16-byte aligned when dims[0] is a multiple of 4 (rows are
 * fetched with bulk copies / TMA / 128-bit loads); both pools must be float-aligned.  Misaligned pointers are
 * rejected with B2_ERR_ARG.
 */
int b2_train_steps(b2_ctx* ctx, const void* x_pool_dev, const void* y_pool_dev, long pool_rows,
                   int local_batch, long global_batch, int n_steps, long first_row, float lr,
                   float* loss_trace_dev, int* correct_trace_dev, b2_stream_t stream);

/*
 * Loss and correct-row count over `rows` device-resident rows without touching the parameters
 * (distributed.py:164 final accuracy; simple.py:37 final loss).  Rows are processed in chunks of
 * at most max_local_batch.  Results are written to device memory (stream-ordered).
 */
int b2_eval(b2_ctx* ctx, const void* x_dev, const void* y_dev, long rows, float* loss_dev,
            long* correct_dev, b2_stream_t stream);

/*
 * The reference's call shape: sess.run([train_op, cross_entropy, accuracy], feed_dict={x:..., y_:...})
 * (distributed.py:148-150; simple.py:35).  x_host / y_host are HOST buffers holding this rank's
 * `rows` rows; they are copied to the device, one step runs, and the pre-update loss and
 * correct-row count come back to the host.  Synchronous.  This is the end-to-end path bench.py
 * reports as `e2e`.  global_batch = rows * nranks.  The call runs on a stream the context owns and
 * synchronises it; do not interleave it with b2_train_steps work still in flight on another stream.
 */
int b2_session_run_train(b2_ctx* ctx, const float* x_host, const float* y_host, int rows, float lr,
                         float* loss_out, int* correct_out);

/*
 * One training step of the 784 -> 10 softmax cross-entropy model that allocates nothing.
 * NO COUNTERPART IN THE REFERENCE: this exists for callers that cannot call cudaMalloc -- a plugin
 * running inside another framework's execution callback (that framework owns the allocator), a
 * latency-critical loop that cannot take the device-wide synchronisation an allocation implies, or
 * a host whose own memory pool already holds all of the device memory.  b2_create() allocates, so
 * those callers use this instead and pass every input and output themselves:
 *   x_dev [rows,784], y_dev [rows,10]   this call's batch (device, float32, row-major; x_dev 16-byte aligned,
 *                                       e.g. not a view at an odd float offset -- rejected with B2_ERR_ARG)
 *   W_dev [784,10], b_dev [10]          parameters, updated IN PLACE: var -= lr * grad
 *   loss_dev [1], correct_dev [1]       pre-update mean loss and correct-row count (either may be NULL)
 *   global_batch                        the divisor of the loss mean (= rows on a single GPU)
 * The step kernel's scratch (per-CTA partial sums) is a fixed array inside the library image:
 * 160 slots x 7856 floats = 5,027,840 bytes, which every GPU that loads the library holds until the
 * process exits.  LIMITS: at most 160 CTAs are used (the kernel loops over tiles, so more SMs are not
 * an error) and rows must be <= 1,048,576 per call; a larger rows is rejected with B2_ERR_ARG.
 * The first call on each device looks up the array's device address and opts the kernel into its
 * shared-memory size; make that first call with the device current and BEFORE starting any stream
 * capture.  Later calls only launch two kernels on `stream`; no cudaMalloc/cudaFree, no library
 * allocation.  All calls on a device share the array and it keeps the last call's contents (nothing
 * needs clearing: each call overwrites every entry it later reads).  Calls whose kernels can overlap
 * on the GPU -- different streams, or different host threads -- race on it; serialise them on one
 * stream.  Not thread-safe.
 */
int b2_linear_step_noalloc(const float* x_dev, const float* y_dev, int rows, long global_batch, float lr,
                           float* W_dev, float* b_dev, float* loss_dev, int* correct_dev,
                           b2_stream_t stream);

/*
 * The layer matrix product on the tensor cores (tcgen05.mma, bf16 operands, fp32 accumulate), exposed
 * on its own: what tf.matmul does at distributed.py:96,98 and what its TF1 gradient builds
 * (transpose_a / transpose_b products), with the element-wise op that follows fused into the epilogue.
 *   a_mn = 0: A is [M,K] row-major        a_mn = 1: A is [K,M] row-major (i.e. A^T is given)
 *   b_mn = 0: B is [N,K] row-major        b_mn = 1: B is [K,N] row-major
 *   epi  = 0: out(bf16) = sigmoid(acc + bias[N])     forward of a hidden layer (distributed.py:96-97)
 *          1: out(bf16) = acc * act * (1 - act)      input gradient through the sigmoid (act: bf16 [M,N])
 *          2: out(fp32) = acc                        weight gradient
 *          3: out(fp32) = acc + bias[N]              forward of a linear layer
 * All matrices are device pointers, 16-byte aligned, leading dimensions multiples of 8.  Through this entry point
 * every CTA fetches its own operand tiles (no cluster multicast), because the pointers come from outside the library.
 */
int b2_matmul_bf16(const void* A, int a_mn, const void* B, int b_mn, int M, int N, int K, int epi,
                   const float* bias, const void* act, void* out, b2_stream_t stream);

/*
 * tf.random_normal(shape, mean, stddev) on the device, for distributed.py:85-86 (initial weights): fills out_dev[0..n)
 * with the values TensorFlow 1.x's CPU kernel produces for an op whose attributes are (seed, seed2) -- Philox4x32-10
 * keyed by `seed`, counter offset `seed2`, four Box-Muller normals per Philox block.  With tf.set_random_seed(2)
 * (distributed.py:83) `seed` is 2 and `seed2` is the graph's op counter at the time the op was created, which depends
 * on the TensorFlow version; callers that know it can start from the reference's own initial weights instead of
 * injecting them.  Restated from TensorFlow's published algorithm; not verifiable against TensorFlow on this box.
 */
int b2_random_normal(float* out_dev, long n, unsigned long long seed, unsigned long long seed2, float mean, float stddev,
                     b2_stream_t stream);

/* The float32 `global_step` variable (distributed.py:68-70,108) as of the last enqueued step. */
float b2_global_step(const b2_ctx* ctx);

/*
 * Device time per kernel class since the last call, measured with CUDA events on the stream the
 * kernels were launched on, while B2_OPT_KERNEL_TIMING is on.  Synchronises.  Replaces the
 * wall-clock deltas the reference prints (distributed.py:155-161 "Time spent (delta)") with
 * per-kernel figures for the roofline report.  ms_total and launches have B2_KERNEL_CLASSES entries.
 */
int b2_kernel_times(b2_ctx* ctx, float* ms_total, long* launches);
const char* b2_kernel_class_name(int kernel_class);

/* How many of this library's kernels the context has launched so far (bench.py's gpu_launches). */
long b2_launch_count(const b2_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* B2MLP_H_ */
