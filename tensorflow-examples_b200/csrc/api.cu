// C ABI of libb2mlp (see include/b2mlp.h): context life cycle, parameter I/O, the training step
// orchestration (which kernels run, in which order, on which stream) and the NCCL plumbing.
#include <dlfcn.h>
#include <stdarg.h>
#include <stdlib.h>

#include "common.cuh"

// ---------------------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------------------
static thread_local char g_err[1024] = "";

void b2_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* b2_last_error(void) { return g_err; }
extern "C" const char* b2_version(void) {
  return "b2mlp 0.4 sm_100a (fused 784->10 step + persistent multi-step kernels, fp32 layers incl. tcgen05 three-plane split, bf16 tcgen05 path, NVLink in-kernel gradient exchange + NCCL)";
}

// ---------------------------------------------------------------------------------------------
// NCCL, resolved at run time so the library loads (and its symbols can be checked) on a box
// without a GPU or without NCCL on the loader path.
// ---------------------------------------------------------------------------------------------
namespace {
struct nccl_uid { char internal[128]; };
typedef int (*pfn_ncclGetUniqueId)(nccl_uid*);
typedef int (*pfn_ncclCommInitRank)(void**, int, nccl_uid, int);
typedef int (*pfn_ncclCommDestroy)(void*);
typedef int (*pfn_ncclAllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*pfn_ncclReduceScatter)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*pfn_ncclAllGather)(const void*, void*, size_t, int, void*, cudaStream_t);
typedef const char* (*pfn_ncclGetErrorString)(int);
typedef int (*pfn_ncclGroup)(void);
struct nccl_api {
  void* handle;
  pfn_ncclGetUniqueId GetUniqueId;
  pfn_ncclCommInitRank CommInitRank;
  pfn_ncclCommDestroy CommDestroy;
  pfn_ncclAllReduce AllReduce;
  pfn_ncclReduceScatter ReduceScatter;
  pfn_ncclAllGather AllGather;
  pfn_ncclGetErrorString GetErrorString;
  pfn_ncclGroup GroupStart, GroupEnd;
};
nccl_api g_nccl = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
constexpr int kNcclFloat = 7;  // ncclFloat32
constexpr int kNcclSum = 0;    // ncclSum

int load_nccl() {
  if (g_nccl.handle) return B2_OK;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);  // already mapped by torch?
  if (!h) {
    const char* env = getenv("B2_NCCL_LIB");
    if (env && *env) h = dlopen(env, RTLD_NOW | RTLD_GLOBAL);
  }
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) {
    b2_set_error("NCCL not found (set B2_NCCL_LIB to libnccl.so.2): %s", dlerror());
    return B2_ERR_NCCL;
  }
  g_nccl.GetUniqueId = (pfn_ncclGetUniqueId)dlsym(h, "ncclGetUniqueId");
  g_nccl.CommInitRank = (pfn_ncclCommInitRank)dlsym(h, "ncclCommInitRank");
  g_nccl.CommDestroy = (pfn_ncclCommDestroy)dlsym(h, "ncclCommDestroy");
  g_nccl.AllReduce = (pfn_ncclAllReduce)dlsym(h, "ncclAllReduce");
  g_nccl.ReduceScatter = (pfn_ncclReduceScatter)dlsym(h, "ncclReduceScatter");
  g_nccl.AllGather = (pfn_ncclAllGather)dlsym(h, "ncclAllGather");
  g_nccl.GetErrorString = (pfn_ncclGetErrorString)dlsym(h, "ncclGetErrorString");
  g_nccl.GroupStart = (pfn_ncclGroup)dlsym(h, "ncclGroupStart");
  g_nccl.GroupEnd = (pfn_ncclGroup)dlsym(h, "ncclGroupEnd");
  if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.CommDestroy || !g_nccl.AllReduce || !g_nccl.ReduceScatter ||
      !g_nccl.AllGather ||
      !g_nccl.GetErrorString || !g_nccl.GroupStart || !g_nccl.GroupEnd) {
    b2_set_error("NCCL library lacks a required symbol");
    return B2_ERR_NCCL;
  }
  g_nccl.handle = h;
  return B2_OK;
}

#define B2_NCCL(expr)                                                                \
  do {                                                                               \
    int _r = (expr);                                                                 \
    if (_r != 0) {                                                                   \
      b2_set_error("%s failed: %s", #expr, g_nccl.GetErrorString(_r));               \
      return B2_ERR_NCCL;                                                            \
    }                                                                                \
  } while (0)

inline long round4(long v) { return (v + 3) & ~3L; }
inline long w_size(const b2_ctx* c, int l) { return (long)c->dims[l] * c->dims[l + 1]; }
inline long b_size(const b2_ctx* c, int l) { return c->dims[l + 1]; }
inline long part_stride(const b2_ctx* c, int l) { return round4(w_size(c, l)) + round4(b_size(c, l)); }

constexpr int kMaxTimed = 16384;

// Records an event pair around one launch when timing is on.  Past kMaxTimed launches it stops recording.
struct timed_scope {
  b2_ctx* c; cudaStream_t s; int idx;
  timed_scope(b2_ctx* c_, int cls, cudaStream_t s_) : c(c_), s(s_), idx(-1) {
    if (!c->opt_timing || c->n_timed >= kMaxTimed) return;
    if (!c->timed) {
      c->timed = (b2_timed*)calloc(kMaxTimed, sizeof(b2_timed));
      if (!c->timed) return;
    }
    if (c->n_timed >= c->cap_timed) {
      b2_timed& t = c->timed[c->cap_timed];
      if (cudaEventCreate(&t.a) != cudaSuccess || cudaEventCreate(&t.b) != cudaSuccess) return;
      c->cap_timed++;
    }
    idx = c->n_timed++;
    c->timed[idx].cls = cls;
    cudaEventRecord(c->timed[idx].a, s);
  }
  ~timed_scope() { if (idx >= 0) cudaEventRecord(c->timed[idx].b, s); }
};

bool use_fused_linear(const b2_ctx* c) {
  return !c->opt_force_layered && c->precision == B2_PREC_FP32 && c->n_layers == 1 &&
         b2_fused_linear_supported(c->dims[0], c->dims[1], c->loss);
}

inline int round8(int v) { return (v + 7) & ~7; }
// a wide fp32 layer goes to the tensor cores (three-plane split) once the batch is large enough to fill them
inline bool use_split3(const b2_ctx* c, int l, int rows) {
  return c->s3_w[l] != nullptr && !c->opt_no_split3 && rows >= 1024;
}

// the output layer (forward, loss head and its whole backward) as one kernel when it is narrow enough (narrow_head.cu)
inline bool use_narrow_head(const b2_ctx* c) {
  const int n = c->n_layers;
  return c->precision == B2_PREC_FP32 && n >= 2 && !c->opt_no_narrow_head && b2_narrow_head_applies(c->dims[n - 1], c->dims[n]);
}

// Enqueue forward + loss head (+ backward partials) for `rows` rows.  train == 0: forward and head only.
int enqueue_layered(b2_ctx* c, const float* x, const float* y, int rows, float inv_B, int train,
                    int* splits_used, cudaStream_t s) {
  const int n = c->n_layers;
  const float* a_in = x;
  const bool narrow = use_narrow_head(c);
  for (int l = 0; l < (narrow ? n - 1 : n); ++l) {
    float* out = (l == n - 1) ? c->z_last : c->act_buf[l];
    int rc;
    {
      timed_scope t(c, B2_K_DENSE_FWD, s);
      const int act_l = (l == n - 1) ? B2_ACT_IDENTITY : c->act;
      if (use_split3(c, l, rows)) {
        const int K = c->dims[l], N = c->dims[l + 1], ldn = round8(N);
        rc = b2_launch_split3(c->params + c->w_off[l], c->s3_w[l], K, N, N, ldn, (long)K * ldn, c->sm_count, s);
        int bq = 224, one = 1;
        b2_gemm_split3_plan(rows, N, K, 1, c->sm_count, &bq, &one);
        if (!rc) rc = b2_launch_gemm_split3(a_in, 0, K, c->s3_w[l], 1, ldn, (long)K * ldn, rows, N, K, 0, act_l,
                                            c->params + c->b_off[l], out, N, bq, 1, 0, c->sm_count, s);
        c->launches += 1;
      } else if (b2_small_m_fwd_applies(rows, c->dims[l]))
        rc = b2_launch_small_m_fwd(a_in, c->params + c->w_off[l], c->params + c->b_off[l], out, rows, c->dims[l + 1],
                                   c->dims[l], act_l, s);
      else
        rc = b2_launch_sgemm_fwd(a_in, c->params + c->w_off[l], c->params + c->b_off[l], out, rows, c->dims[l + 1],
                                 c->dims[l], act_l, s);
    }
    if (rc) return rc;
    c->launches++;
    a_in = out;
  }
  float* head = c->scratch + c->head_part_off;
  int rc;
  int cur = 0, l_top = n - 1;
  if (narrow) {
    // forward, loss head, dW / db partials and dz of the layer below, all from one pass over the activations
    const int L = n - 1;
    const int grid = b2_narrow_head_grid(rows, c->sm_count);
    float* pw = c->scratch + c->part_off[L];
    timed_scope t(c, B2_K_LOSS_HEAD, s);
    const bool emit_db = train && c->nh_db_off >= 0;
    // the layer below runs its weight gradient on the tensor cores: hand it dz as bf16 planes straight from here; the
    // fp32 copy is only needed when a further layer below needs the input gradient (dgrad) as well
    const bool planes = train && L >= 1 && use_split3(c, L - 1, rows);
    const int ldn = round8(c->dims[L]);
    c->nh_planes_used = planes ? 1 : 0;
    rc = b2_launch_narrow_head(a_in, y, c->params + c->w_off[L], c->params + c->b_off[L],
                               (planes && L - 1 == 0 && emit_db) ? nullptr : c->dz[1], planes ? c->s3_dz : nullptr,
                               (long)c->max_local_batch * ldn, ldn, pw,
                               pw + round4(w_size(c, L)), part_stride(c, L), emit_db ? c->scratch + c->nh_db_off : nullptr,
                               round4(c->dims[L]), head, head + c->head_blocks, rows, c->dims[L],
                               c->dims[n], c->act, c->loss, train, inv_B, grid, s);
    c->head_used = grid;
    c->nh_db_used = emit_db ? 1 : 0;
    if (train) splits_used[L] = grid;
    cur = 1;
    l_top = L - 1;
  } else {
    timed_scope t(c, B2_K_LOSS_HEAD, s);
    rc = b2_launch_loss_head(c->z_last, y, c->dz[0], head, head + c->head_blocks, rows, c->dims[n],
                             c->loss, inv_B, c->head_blocks, s);
    c->head_used = c->head_blocks;
    c->nh_db_used = 0;
    c->nh_planes_used = 0;
  }
  if (rc) return rc;
  c->launches++;
  if (!train) return B2_OK;
  for (int l = l_top; l >= 0; --l) {
    const float* a_prev = (l == 0) ? x : c->act_buf[l - 1];
    int sp = c->splits[l];
    const int max_sp = (rows + 63) / 64;
    if (sp > max_sp) sp = max_sp;
    int s3_bq = 224;
    if (use_split3(c, l, rows)) b2_gemm_split3_plan(c->dims[l], c->dims[l + 1], rows, sp, c->sm_count, &s3_bq, &sp);
    if (sp < 1) sp = 1;
    splits_used[l] = sp;
    float* pw = c->scratch + c->part_off[l];
    float* pb = pw + round4(w_size(c, l));
    {
      timed_scope t(c, B2_K_WGRAD, s);
      if (use_split3(c, l, rows)) {
        // dz is split into planes here (the layer input is split inside the GEMM); its column sums are the bias gradient
        const int K = c->dims[l], N = c->dims[l + 1], ldn = round8(N);
        const long dzp = (long)c->max_local_batch * ldn;
        const bool from_head = c->nh_planes_used && l == n - 2;     // the narrow-head kernel already wrote the planes
        rc = from_head ? B2_OK : b2_launch_split3(c->dz[cur], c->s3_dz, rows, N, N, ldn, dzp, c->sm_count, s);
        if (!from_head) c->launches += 1;
        if (!rc && !(c->nh_db_used && l == n - 2))
          rc = b2_launch_colsum_f32(c->dz[cur], pb, part_stride(c, l), rows, N, sp, s);
        if (!rc) rc = b2_launch_gemm_split3(a_prev, 1, K, c->s3_dz, 1, ldn, dzp, K, N, rows, 1, 0, nullptr, pw, N, s3_bq, sp,
                                            part_stride(c, l), c->sm_count, s);
      } else {
        rc = b2_launch_sgemm_wgrad(a_prev, c->dz[cur], pw, pb, part_stride(c, l), rows, c->dims[l],
                                   c->dims[l + 1], sp, s);
      }
    }
    if (rc) return rc;
    c->launches++;
    if (l > 0) {
      {
        timed_scope t(c, B2_K_DGRAD, s);
        rc = b2_launch_dgrad(c->dz[cur], c->params + c->w_off[l], c->act_buf[l - 1], c->dz[cur ^ 1], rows,
                             c->dims[l], c->dims[l + 1], c->act, s);
      }
      if (rc) return rc;
      c->launches++;
      cur ^= 1;
    }
  }
  return B2_OK;
}

// Sharded update -- the analogue of k parameter servers (distributed.py:63-65).  Rank r owns elements [r*S, (r+1)*S) of
// the flat parameter vector: reduce-scatter of the gradients (in place), SGD apply on the own slice, all-gather of the
// updated parameters (in place).  loss / correct sit in one rank's slice of the gradient vector, so they are summed on
// the side.  Both buffers are allocated with room for nranks * S elements.
inline long shard_size(const b2_ctx* c) { return round4((c->P + 4 + c->nranks - 1) / c->nranks); }

int exchange_sharded(b2_ctx* c, float lr, float* loss_trace, int* correct_trace, cudaStream_t s) {
  const long S = shard_size(c);
  B2_CUDA(cudaMemcpyAsync(c->lc_dev, c->grads + c->P, 4 * sizeof(float), cudaMemcpyDeviceToDevice, s));
  B2_NCCL(g_nccl.GroupStart());
  B2_NCCL(g_nccl.AllReduce(c->lc_dev, c->lc_dev, 4, kNcclFloat, kNcclSum, c->nccl_comm, s));
  B2_NCCL(g_nccl.ReduceScatter(c->grads, c->grads + (long)c->rank * S, (size_t)S, kNcclFloat, kNcclSum, c->nccl_comm, s));
  B2_NCCL(g_nccl.GroupEnd());
  long n = c->P - (long)c->rank * S;                       // my slice may reach past the parameters (padding, loss, correct)
  n = n < 0 ? 0 : (n > S ? S : n);
  int rc = b2_launch_sharded_apply(c->params + (long)c->rank * S, c->grads + (long)c->rank * S, n, lr, c->lc_dev, loss_trace,
                                   correct_trace, c->global_step_dev, c->sm_count, s);
  if (rc) return rc;
  c->launches++;
  B2_NCCL(g_nccl.AllGather(c->params + (long)c->rank * S, c->params, (size_t)S, kNcclFloat, c->nccl_comm, s));
  c->wb_stale = 1;                                         // bf16 path: the weight copies are refreshed before the next forward
  return B2_OK;
}

// ---- bf16-operand path: tensor-core GEMMs for the hidden layers, CUDA cores for the narrow output layer ----
inline bool is_mixed(const b2_ctx* c) { return c->precision == B2_PREC_BF16_FP32ACC; }

int refresh_bf16_weights(b2_ctx* c, cudaStream_t s) {
  for (int l = 0; l < c->n_layers - 1; ++l) {
    timed_scope t(c, B2_K_MIXED_AUX, s);
    int rc = b2_launch_cast_bf16(c->params + c->w_off[l], c->wb[l], w_size(c, l), c->sm_count, s);
    if (rc) return rc;
    c->launches++;
  }
  c->wb_stale = 0;
  return B2_OK;
}

// the dgrad GEMM that produces the dz of layer l-1 (width dims[l]) can also emit its bias-gradient partials when every
// 32-column chunk of its output is full (vector epilogue)
inline bool fused_colsum(const b2_ctx* c, int l) { return c->dims[l] % 32 == 0; }

// slices of the output layer's weight-gradient GEMM (split over the batch) and of its bias column sums
inline int narrow_slices(const b2_ctx* c, int rows) {
  const int L = c->n_layers - 1;
  const int tiles_m = (c->dims[L] + 127) / 128;
  int sp = (c->sm_count + tiles_m - 1) / tiles_m;
  const int num_kb = (rows + 63) / 64;
  if (sp > num_kb) sp = num_kb;
  if (sp > 16) sp = 16;
  return sp < 1 ? 1 : sp;
}

// forward + loss head; leaves logits in z_last, dz of the output layer (fp32) in dz[0], head partials in scratch
int enqueue_forward_mixed(b2_ctx* c, const float* x, const float* y, int rows, float inv_B, cudaStream_t s) {
  const int n = c->n_layers;
  int rc;
  if (c->wb_stale && (rc = refresh_bf16_weights(c, s))) return rc;
  {
    timed_scope t(c, B2_K_MIXED_AUX, s);
    rc = b2_launch_cast_bf16(x, c->xb, (long)rows * c->dims[0], c->sm_count, s);
  }
  if (rc) return rc;
  c->launches++;
  const void* a_in = c->xb;
  for (int l = 0; l < n - 1; ++l) {
    // a_out = sigmoid(a_in . W + b)   distributed.py:96-97   A K-major [rows, in], B = W [in, out] MN-major
    timed_scope t(c, B2_K_GEMM_FWD, s);
    rc = b2_launch_gemm_bf16(a_in, 0, c->wb[l], 1, rows, c->dims[l + 1], c->dims[l], 0, c->params + c->b_off[l],
                             nullptr, c->ab[l], c->sm_count, s);
    if (rc) return rc;
    c->launches++;
    a_in = c->ab[l];
  }
  {
    // output layer (distributed.py:98): W_out [K,10] padded to a bf16 [K,64] operand, logits = a . W_out + b
    timed_scope t(c, B2_K_MIXED_AUX, s);
    rc = b2_launch_pad_cast64(c->params + c->w_off[n - 1], c->wb[n - 1], c->dims[n - 1], c->dims[n], c->sm_count, s);
  }
  if (rc) return rc;
  c->launches++;
  {
    timed_scope t(c, B2_K_GEMM_FWD, s);
    rc = b2_launch_gemm_bf16_ex(a_in, 0, c->wb[n - 1], 1, rows, 64, c->dims[n - 1], 3, c->params + c->b_off[n - 1],
                                nullptr, c->z_last, c->dims[n], c->dims[n], 1, 0, nullptr, c->sm_count, s);
  }
  if (rc) return rc;
  c->launches++;
  float* head = c->scratch + c->head_part_off;
  {
    timed_scope t(c, B2_K_LOSS_HEAD, s);
    // one thread per row (b2_create limits this path to <= 16 classes): also writes dz as the padded bf16 operand of the
    // output layer's tensor-core products and the per-block column sums of dz (its bias gradient)
    rc = b2_launch_loss_head_rows(c->z_last, y, c->dzn, head, head + c->head_blocks, c->scratch + c->head_db_off, rows,
                                  c->dims[n], c->loss, inv_B, c->head_blocks, s);
  }
  if (rc) return rc;
  c->launches++;
  return B2_OK;
}

int enqueue_step_mixed(b2_ctx* c, const float* x, const float* y, int rows, long global_batch, float lr,
                       float* loss_trace, int* correct_trace, cudaStream_t s) {
  const int n = c->n_layers;
  const float inv_B = 1.0f / (float)global_batch;
  int rc = enqueue_forward_mixed(c, x, y, rows, inv_B, s);
  if (rc) return rc;
  // ---- output layer on the tensor cores: dz padded to bf16 [rows,64] ----
  const int L = n - 1;
  const int slices = narrow_slices(c, rows);
  float* pw = c->scratch + c->part_off[L];
  // (the loss head already left dz as the padded bf16 operand dzn and the bias-gradient partials)
  {
    // dW_out partials = a^T . dz   (MatMul grad, transpose_a), K = batch cut into `slices` fixed pieces
    timed_scope t(c, B2_K_GEMM_WGRAD, s);
    rc = b2_launch_gemm_bf16_ex(c->ab[L - 1], 1, c->dzn, 1, c->dims[L], 64, rows, 2, nullptr, nullptr, pw, c->dims[n],
                                c->dims[n], slices, part_stride(c, L), nullptr, c->sm_count, s);
  }
  if (rc) return rc;
  c->launches++;
  // dz_prev = (dz . W_out^T) * a(1-a)   (MatMul grad transpose_b + SigmoidGrad)
  // ten classes in the reduction: an element-wise pass on the CUDA cores, not a GEMM; it also leaves the column sums of
  // its output (bias gradient of layer L-1) as one partial row per 64-row block
  int colsum_slices = b2_narrow_dgrad_slices(rows);
  bool colsum_pending = true;
  {
    timed_scope t(c, B2_K_GEMM_DGRAD, s);
    rc = b2_launch_narrow_dgrad_bf16(c->dzn, c->wb[L], c->ab[L - 1], c->dzb[0], c->colsum_part, rows, c->dims[L], c->dims[n], s);
  }
  if (rc) return rc;
  c->launches++;
  const bool exchange = c->nranks > 1 && c->async_period == 0;
  const bool sharded = exchange && c->opt_sharded;
  const bool overlap = exchange && !sharded && !c->opt_no_overlap && !c->opt_timing && c->comm_stream != nullptr;
  bool bucketed[B2_MAX_LAYERS] = {};
  int cur = 0;
  for (int l = L - 1; l >= 0; --l) {
    // dz of layer l (bf16 [rows, dims[l+1]]) is in dzb[cur]
    const void* a_prev = (l == 0) ? c->xb : c->ab[l - 1];
    {
      timed_scope t(c, B2_K_MIXED_AUX, s);
      if (colsum_pending) {
        rc = b2_launch_colsum_finish(c->colsum_part, c->grads + c->b_off[l], c->dims[l + 1], colsum_slices, s);
        c->launches += 1;
      } else {
        rc = b2_launch_colsum_bf16(c->dzb[cur], c->colsum_part, c->grads + c->b_off[l], rows, c->dims[l + 1],
                                   b2_colsum_slices(rows, c->dims[l + 1], c->sm_count), s);
        c->launches += 2;
      }
    }
    if (rc) return rc;
    {
      // dW = a_prev^T . dz     both operands MN-major (global [rows, in] and [rows, out]); fp32 straight into grads
      timed_scope t(c, B2_K_GEMM_WGRAD, s);
      rc = b2_launch_gemm_bf16(a_prev, 1, c->dzb[cur], 1, c->dims[l], c->dims[l + 1], rows, 2, nullptr, nullptr,
                               c->grads + c->w_off[l], c->sm_count, s);
    }
    if (rc) return rc;
    c->launches++;
    if (overlap && l > 0) {
      // this layer's weight gradient is final -> sum it over the ranks on the second stream while the
      // dgrad / wgrad GEMMs of the layers below keep the tensor cores busy on this one
      B2_CUDA(cudaEventRecord(c->ev_bucket, s));
      B2_CUDA(cudaStreamWaitEvent(c->comm_stream, c->ev_bucket, 0));
      B2_NCCL(g_nccl.AllReduce(c->grads + c->w_off[l], c->grads + c->w_off[l], (size_t)w_size(c, l), kNcclFloat, kNcclSum,
                               c->nccl_comm, c->comm_stream));
      bucketed[l] = true;
    }
    if (l > 0) {
      // dz_prev = (dz . W^T) * a(1-a)   A = dz K-major [rows, out], B = W [in, out] K-major
      timed_scope t(c, B2_K_GEMM_DGRAD, s);
      colsum_pending = fused_colsum(c, l);
      colsum_slices = (rows + 31) / 32;
      rc = b2_launch_gemm_bf16_ex(c->dzb[cur], 0, c->wb[l], 0, rows, c->dims[l], c->dims[l + 1], 1, nullptr, c->ab[l - 1],
                                  c->dzb[cur ^ 1], c->dims[l], c->dims[l], 1, 0, colsum_pending ? c->colsum_part : nullptr,
                                  c->sm_count, s);
      if (rc) return rc;
      c->launches++;
      cur ^= 1;
    }
  }
  // ---- output-layer partials + loss / correct -> grads ----
  b2_reduce_args a;
  memset(&a, 0, sizeof(a));
  a.n_layers = 1;
  a.w_off[0] = c->w_off[L]; a.b_off[0] = c->b_off[L];
  a.w_size[0] = w_size(c, L); a.b_size[0] = b_size(c, L);
  a.part_off[0] = c->part_off[L];
  a.pstride[0] = part_stride(c, L);
  a.splits[0] = slices;
  // the bias gradient comes from the loss head's per-block partials: a reduce "layer" of its own
  a.b_size[0] = 0;
  a.n_layers = 2;
  a.w_off[1] = 0; a.w_size[1] = 0;
  a.b_off[1] = c->b_off[L]; a.b_size[1] = b_size(c, L);
  a.part_off[1] = c->head_db_off;
  a.pstride[1] = 16;
  a.splits[1] = c->head_blocks;
  a.P = c->P;
  a.head_part_off = c->head_part_off;
  a.head_blocks = c->head_blocks;
  a.head_corr_off = c->head_part_off + c->head_blocks;
  {
    timed_scope t(c, B2_K_REDUCE_APPLY, s);
    rc = b2_launch_reduce(a, c->scratch, c->grads, c->params, 0, lr, nullptr, nullptr, c->global_step_dev,
                          c->sm_count, nullptr, s);
  }
  if (rc) return rc;
  c->launches++;
  if (sharded) {
    timed_scope t(c, B2_K_ALLREDUCE_APPLY, s);
    rc = exchange_sharded(c, lr, loss_trace, correct_trace, s);
    if (rc) return rc;
    c->global_step_host += 1.0f;
    return B2_OK;
  }
  if (exchange) {
    // everything that was not sent as a bucket: the first layer's weights, all biases, the output layer, loss and correct
    timed_scope t(c, B2_K_ALLREDUCE_APPLY, s);
    B2_NCCL(g_nccl.GroupStart());
    long from = 0;
    for (int l = 0; l <= n; ++l) {
      const bool cut = l < n && bucketed[l];
      const long to = l == n ? c->P + 4 : (cut ? c->w_off[l] : -1);
      if (to < 0) continue;
      if (to > from)
        B2_NCCL(g_nccl.AllReduce(c->grads + from, c->grads + from, (size_t)(to - from), kNcclFloat, kNcclSum, c->nccl_comm, s));
      if (cut) from = c->w_off[l] + w_size(c, l);
    }
    B2_NCCL(g_nccl.GroupEnd());
    if (overlap) {
      B2_CUDA(cudaEventRecord(c->ev_comm_done, c->comm_stream));
      B2_CUDA(cudaStreamWaitEvent(s, c->ev_comm_done, 0));
    }
  }
  long off[B2_MAX_LAYERS], size[B2_MAX_LAYERS];
  for (int l = 0; l < L; ++l) { off[l] = c->w_off[l]; size[l] = w_size(c, l); }
  {
    timed_scope t(c, B2_K_MIXED_AUX, s);
    rc = b2_launch_sgd_apply_mixed(c->params, c->grads, c->P, lr, L, off, size, c->wb, loss_trace, correct_trace,
                                   c->global_step_dev, c->sm_count, s);
  }
  if (rc) return rc;
  c->launches++;
  c->global_step_host += 1.0f;
  return B2_OK;
}

// One full training step on `rows` local rows starting at x / y.
int enqueue_step(b2_ctx* c, const float* x, const float* y, int rows, long global_batch, float lr,
                 float* loss_trace, int* correct_trace, cudaStream_t s) {
  if (is_mixed(c)) return enqueue_step_mixed(c, x, y, rows, global_batch, lr, loss_trace, correct_trace, s);
  const float inv_B = 1.0f / (float)global_batch;
  const int apply = (c->nranks == 1 || c->async_period > 0) ? 1 : 0;
  const bool use_peer = !apply && c->peer_ready && !c->opt_no_peer;
  // gradient exchange inside the kernel that finishes the local reduction (b2_fx_args): push over NVLink, poll locally
  b2_fx_args fx;
  memset(&fx, 0, sizeof(fx));
  if (use_peer) {
    fx.rank = c->rank; fx.nranks = c->nranks; fx.stride = c->peer_stride;
    fx.step1 = c->peer_step + 1;
    fx.parity = (int)(c->peer_step & 1u);
    for (int p = 0; p < c->nranks; ++p) fx.region[p] = (unsigned long long*)c->peer_base[p];
    fx.prof = c->fx_prof;
  }
  int rc;
  if (use_fused_linear(c)) {
    const int tr = b2_fused_linear_tile_rows();
    int grid = (rows + tr - 1) / tr;
    if (grid > c->fused_grid) grid = c->fused_grid;
    {
      timed_scope t(c, B2_K_FUSED_STEP, s);
      rc = b2_launch_fused_linear_step(x, y, rows, c->params + c->w_off[0], c->params + c->b_off[0],
                                       c->fused_scratch, grid, inv_B, c->loss, s);
    }
    if (rc) return rc;
    if (use_peer) {
      timed_scope t(c, B2_K_ALLREDUCE_APPLY, s);
      rc = b2_launch_fused_linear_reduce_exchange(c->fused_scratch, grid, fx, c->grads, c->params, lr, loss_trace,
                                                  correct_trace, c->global_step_dev, c->P, s);
      if (rc) return rc;
      c->peer_step++;
      c->launches += 2;
      c->global_step_host += 1.0f;
      return B2_OK;
    }
    {
      timed_scope t(c, B2_K_FUSED_REDUCE, s);
      rc = b2_launch_fused_linear_reduce(c->fused_scratch, grid, c->grads, c->params, apply, lr,
                                         loss_trace, correct_trace, c->global_step_dev, c->P, s);
    }
    if (rc) return rc;
    c->launches += 2;
  } else {
    b2_reduce_args a;
    memset(&a, 0, sizeof(a));
    rc = enqueue_layered(c, x, y, rows, inv_B, 1, a.splits, s);
    if (rc) return rc;
    a.n_layers = c->n_layers;
    for (int l = 0; l < c->n_layers; ++l) {
      a.w_off[l] = c->w_off[l];
      a.b_off[l] = c->b_off[l];
      a.w_size[l] = w_size(c, l);
      a.b_size[l] = b_size(c, l);
      a.part_off[l] = c->part_off[l];
      a.pstride[l] = part_stride(c, l);
    }
    if (c->nh_db_used) {
      // the bias gradient of layer L-1 was summed per CTA by the narrow-head kernel: reduce it as a layer of its own
      const int L = c->n_layers - 1, e = c->n_layers;
      a.b_size[L - 1] = 0;
      a.w_off[e] = 0; a.w_size[e] = 0;
      a.b_off[e] = c->b_off[L - 1]; a.b_size[e] = b_size(c, L - 1);
      a.part_off[e] = c->nh_db_off;
      a.pstride[e] = round4(c->dims[L]);
      a.splits[e] = a.splits[L];
      a.n_layers = e + 1;
    }
    a.P = c->P;
    a.head_part_off = c->head_part_off;
    a.head_blocks = c->head_used;
    a.head_corr_off = c->head_part_off + c->head_blocks;
    {
      timed_scope t(c, B2_K_REDUCE_APPLY, s);
      rc = b2_launch_reduce(a, c->scratch, c->grads, c->params, use_peer ? 1 : apply, lr, loss_trace, correct_trace,
                            c->global_step_dev, c->sm_count, use_peer ? &fx : nullptr, s);
    }
    if (rc) return rc;
    c->launches++;
    if (use_peer) {
      c->peer_step++;
      c->global_step_host += 1.0f;
      return B2_OK;
    }
  }
  if (!apply) {
    // the parameter-server push/pull of the reference (distributed.py:63-65) becomes one
    // all-reduce of [gradients..., loss, correct]; every rank then applies the same update.
    timed_scope t(c, B2_K_ALLREDUCE_APPLY, s);
    if (c->opt_sharded) {
      rc = exchange_sharded(c, lr, loss_trace, correct_trace, s);
      if (rc) return rc;
    } else {
      B2_NCCL(g_nccl.AllReduce(c->grads, c->grads, (size_t)(c->P + 4), kNcclFloat, kNcclSum, c->nccl_comm, s));
      rc = b2_launch_sgd_apply(c->params, c->grads, c->P, lr, loss_trace, correct_trace,
                               c->global_step_dev, c->sm_count, s);
      if (rc) return rc;
      c->launches++;
    }
  }
  c->global_step_host += 1.0f;
  return B2_OK;
}

void free_ctx(b2_ctx* c) {
  if (!c) return;
  if (c->nccl_comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->nccl_comm);
  if (c->comm_stream) cudaStreamDestroy(c->comm_stream);
  if (c->ev_bucket) cudaEventDestroy(c->ev_bucket);
  if (c->ev_comm_done) cudaEventDestroy(c->ev_comm_done);
  for (int p = 0; p < B2_MAX_PEERS; ++p)
    if (c->peer_base[p] && c->peer_base[p] != (void*)c->peer_local) cudaIpcCloseMemHandle(c->peer_base[p]);
  cudaFree(c->peer_local);
  for (int l = 0; l < B2_MAX_LAYERS; ++l) cudaFree(c->s3_w[l]);
  cudaFree(c->s3_dz);
  cudaFree(c->xb);
  for (int l = 0; l < B2_MAX_LAYERS; ++l) { cudaFree(c->ab[l]); cudaFree(c->wb[l]); }
  cudaFree(c->dzb[0]);
  cudaFree(c->dzb[1]);
  cudaFree(c->colsum_part);
  cudaFree(c->dzn);
  cudaFree(c->params);
  cudaFree(c->grads);
  cudaFree(c->lc_dev);
  for (int l = 0; l < B2_MAX_LAYERS; ++l) cudaFree(c->act_buf[l]);
  cudaFree(c->z_last);
  cudaFree(c->dz[0]);
  cudaFree(c->dz[1]);
  cudaFree(c->scratch);
  cudaFree(c->fused_scratch);
  cudaFree(c->global_step_dev);
  cudaFree(c->eval_acc_dev);
  cudaFree(c->feed_x);
  cudaFree(c->feed_y);
  cudaFree(c->fetch_dev);
  cudaFree(c->grid_bar);
  cudaFree(c->cs_scratch);
  if (c->fetch_host) cudaFreeHost(c->fetch_host);
  if (c->own_stream) cudaStreamDestroy(c->own_stream);
  if (c->timed) {
    for (int i = 0; i < c->cap_timed; ++i) { cudaEventDestroy(c->timed[i].a); cudaEventDestroy(c->timed[i].b); }
    free(c->timed);
  }
  free(c);
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// life cycle
// ---------------------------------------------------------------------------------------------
extern "C" int b2_create(b2_ctx** out, int device, int n_layers, const int* dims, int act, int loss,
                         int precision, int max_local_batch) {
  B2_CHECK_ARG(out != nullptr, "b2_create: out is NULL");
  *out = nullptr;
  B2_CHECK_ARG(n_layers >= 1 && n_layers <= B2_MAX_LAYERS, "b2_create: n_layers must be in [1,%d], got %d",
               B2_MAX_LAYERS, n_layers);
  B2_CHECK_ARG(dims != nullptr, "b2_create: dims is NULL");
  for (int i = 0; i <= n_layers; ++i)
    B2_CHECK_ARG(dims[i] >= 1, "b2_create: dims[%d] = %d must be >= 1", i, dims[i]);
  B2_CHECK_ARG(act == B2_ACT_IDENTITY || act == B2_ACT_SIGMOID, "b2_create: unknown activation %d", act);
  B2_CHECK_ARG(loss == B2_LOSS_SSE_SUM || loss == B2_LOSS_SOFTMAX_CE_MEAN, "b2_create: unknown loss %d", loss);
  B2_CHECK_ARG(precision == B2_PREC_FP32 || precision == B2_PREC_BF16_FP32ACC, "b2_create: unknown precision %d", precision);
  if (precision == B2_PREC_BF16_FP32ACC) {
    // tensor-core path: hidden layers are tcgen05 GEMMs (TMA needs 16-byte row pitches), the output layer is narrow
    B2_CHECK_ARG(n_layers >= 2 && act == B2_ACT_SIGMOID && loss == B2_LOSS_SOFTMAX_CE_MEAN,
                 "b2_create: the bf16 path needs >= 2 layers, sigmoid hidden layers and the softmax-CE head");
    for (int i = 0; i < n_layers; ++i)
      B2_CHECK_ARG(dims[i] % 8 == 0, "b2_create: the bf16 path needs dims[%d] = %d to be a multiple of 8", i, dims[i]);
    B2_CHECK_ARG(dims[n_layers] <= 16, "b2_create: the bf16 path needs <= 16 outputs, got %d", dims[n_layers]);
  }
  B2_CHECK_ARG(max_local_batch >= 1, "b2_create: max_local_batch must be >= 1");

  int ndev = 0;
  B2_CUDA(cudaGetDeviceCount(&ndev));
  B2_CHECK_ARG(device >= 0 && device < ndev, "b2_create: device %d out of range (%d visible)", device, ndev);
  B2_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  B2_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    b2_set_error("b2_create: device %d is sm_%d%d; this library is built for sm_100a (B200) only", device,
                 prop.major, prop.minor);
    return B2_ERR_CUDA;
  }

  b2_ctx* c = (b2_ctx*)calloc(1, sizeof(b2_ctx));
  if (!c) { b2_set_error("b2_create: out of host memory"); return B2_ERR_CUDA; }
  c->device = device;
  c->n_layers = n_layers;
  for (int i = 0; i <= n_layers; ++i) c->dims[i] = dims[i];
  c->act = act; c->loss = loss; c->precision = precision;
  c->max_local_batch = max_local_batch;
  c->sm_count = prop.multiProcessorCount;
  c->rank = 0; c->nranks = 1;

  long off = 0, maxw = 0;
  for (int l = 0; l < n_layers; ++l) {
    c->w_off[l] = off; off += round4(w_size(c, l));
    c->b_off[l] = off; off += round4(b_size(c, l));
    if (dims[l + 1] > maxw) maxw = dims[l + 1];
  }
  c->P = off;

  // scratch: per-layer split partials, then the loss-head partials
  long soff = 0;
  for (int l = 0; l < n_layers; ++l) {
    const int bn = dims[l + 1] <= 64 ? 64 : (dims[l + 1] <= 112 ? 112 : 128);   // column tile of the wgrad kernel
    const long tiles = (long)((dims[l] + 127) / 128) * ((dims[l + 1] + bn - 1) / bn);
    long sp = (4L * c->sm_count + tiles - 1) / tiles;
    const long cap = (max_local_batch + 63) / 64;
    if (sp > cap) sp = cap;
    if (sp > 256) sp = 256;
    if (sp < 1) sp = 1;
    if (precision == B2_PREC_BF16_FP32ACC) {
      // only the narrow output layer writes partials (the GEMMs store dW directly)
      sp = (l == n_layers - 1) ? 16 : 0;   // narrow_slices() never exceeds 16
    }
    if (precision == B2_PREC_FP32 && n_layers >= 2 && l == n_layers - 1 && b2_narrow_head_applies(dims[l], dims[l + 1])) {
      const long g = b2_narrow_head_grid(max_local_batch, c->sm_count);   // one partial per CTA of the narrow-head kernel
      if (sp < g) sp = g;
    }
    c->splits[l] = (int)sp;
    c->part_off[l] = soff;
    soff += sp * part_stride(c, l);
  }
  c->head_blocks = 4 * c->sm_count;   // also the capacity of the narrow-head kernel's loss / correct partials (one per CTA)
  {
    const int cap = (max_local_batch + 7) / 8;
    if (c->head_blocks > cap) c->head_blocks = cap;
  }
  c->head_part_off = soff;
  soff += 2L * c->head_blocks;
  soff = (soff + 3) & ~3L;
  c->head_db_off = soff;
  soff += 16L * c->head_blocks;
  c->nh_db_off = -1;
  if (precision == B2_PREC_FP32 && n_layers >= 2 && n_layers < B2_MAX_LAYERS &&
      b2_narrow_head_applies(dims[n_layers - 1], dims[n_layers])) {
    c->nh_db_off = soff;
    soff += (long)b2_narrow_head_grid(max_local_batch, c->sm_count) * round4(dims[n_layers - 1]);
  }
  c->scratch_floats = soff;

#define B2_ALLOC(ptr, nfloats)                                                         \
  do {                                                                                 \
    cudaError_t _e = cudaMalloc((void**)&(ptr), (size_t)(nfloats) * sizeof(float));    \
    if (_e != cudaSuccess) {                                                           \
      b2_set_error("b2_create: cudaMalloc of %ld bytes failed: %s",                    \
                   (long)((size_t)(nfloats) * sizeof(float)), cudaGetErrorString(_e)); \
      free_ctx(c);                                                                     \
      return B2_ERR_CUDA;                                                              \
    }                                                                                  \
  } while (0)

  // (+ 4 * B2_MAX_PEERS: a sharded update cuts both vectors into nranks slices of round4(ceil((P+4)/nranks)) elements)
  B2_ALLOC(c->params, c->P + 4 + 4 * B2_MAX_PEERS);
  B2_ALLOC(c->grads, c->P + 4 + 4 * B2_MAX_PEERS);
  B2_ALLOC(c->lc_dev, 4);
  B2_ALLOC(c->z_last, (long)max_local_batch * dims[n_layers]);
  if (precision == B2_PREC_BF16_FP32ACC) {
    // bf16 buffers are sized in floats / 2 (dims are multiples of 8)
    B2_ALLOC(c->dz[0], (long)max_local_batch * dims[n_layers]);
    B2_ALLOC(c->xb, (long)max_local_batch * dims[0] / 2);
    long hid = 0;
    for (int l = 0; l < n_layers - 1; ++l) {
      B2_ALLOC(c->ab[l], (long)max_local_batch * dims[l + 1] / 2);
      B2_ALLOC(c->wb[l], (long)dims[l] * dims[l + 1] / 2);
      if (dims[l + 1] > hid) hid = dims[l + 1];
    }
    B2_ALLOC(c->wb[n_layers - 1], (long)dims[n_layers - 1] * 64 / 2);   // output layer padded to 64 columns
    B2_ALLOC(c->dzn, (long)max_local_batch * 64 / 2);
    B2_ALLOC(c->dzb[0], (long)max_local_batch * hid / 2);
    if (n_layers > 2) B2_ALLOC(c->dzb[1], (long)max_local_batch * hid / 2);
    {
      long sl = b2_colsum_slices(max_local_batch, (int)hid, c->sm_count);
      const long blocks32 = ((max_local_batch + 127) / 128) * 4;   // one partial row per 32-row block of every 128-row GEMM tile
      if (blocks32 > sl) sl = blocks32;
      B2_ALLOC(c->colsum_part, sl * hid + 16);
    }
    c->wb_stale = 1;
  } else {
    for (int l = 0; l < n_layers - 1; ++l) B2_ALLOC(c->act_buf[l], (long)max_local_batch * dims[l + 1]);
    // wide layers (K >= 256, N >= 32) get three-plane bf16 copies for the tensor-core path; floats = 3 planes * elements / 2
    long s3_ldn = 0;
    for (int l = 0; l < n_layers && max_local_batch >= 1024; ++l) {
      if (dims[l] < 256 || dims[l] % 4 || dims[l + 1] < 32) continue;
      B2_ALLOC(c->s3_w[l], 3L * dims[l] * round8(dims[l + 1]) / 2 + 8);
      if (round8(dims[l + 1]) > s3_ldn) s3_ldn = round8(dims[l + 1]);
    }
    if (s3_ldn) {
      B2_ALLOC(c->s3_dz, 3L * max_local_batch * s3_ldn / 2 + 8);
    }
    B2_ALLOC(c->dz[0], (long)max_local_batch * maxw);
    if (n_layers > 1) B2_ALLOC(c->dz[1], (long)max_local_batch * maxw);
  }
  B2_ALLOC(c->scratch, c->scratch_floats);
  B2_ALLOC(c->global_step_dev, 1);
  B2_ALLOC(c->eval_acc_dev, 2);
  B2_ALLOC(c->fetch_dev, 2);
  B2_ALLOC(c->grid_bar, 2);
  if (precision == B2_PREC_FP32 && b2_cluster_steps_applies(n_layers, dims, 1))
    B2_ALLOC(c->cs_scratch, b2_cluster_steps_scratch_floats(n_layers));
  if (b2_fused_linear_supported(dims[0], dims[1], loss) && n_layers == 1) {
    c->fused_grid = b2_fused_linear_grid(c->sm_count);
    c->fused_stride = b2_fused_linear_stride();
    B2_ALLOC(c->fused_scratch, (long)c->fused_grid * c->fused_stride);
    if (b2_fused_linear_prepare() != B2_OK || b2_fused_linear_ms_prepare() != B2_OK) { free_ctx(c); return B2_ERR_CUDA; }
  }
#undef B2_ALLOC
  cudaError_t e = cudaMemset(c->params, 0, (size_t)(c->P + 4 + 4 * B2_MAX_PEERS) * sizeof(float));
  if (e == cudaSuccess) e = cudaMemset(c->grads, 0, (size_t)(c->P + 4 + 4 * B2_MAX_PEERS) * sizeof(float));
  if (e == cudaSuccess) e = cudaMemset(c->global_step_dev, 0, sizeof(float));
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaMallocHost((void**)&c->fetch_host, 2 * sizeof(float));
  if (e != cudaSuccess) {
    b2_set_error("b2_create: initialisation failed: %s", cudaGetErrorString(e));
    free_ctx(c);
    return B2_ERR_CUDA;
  }
  *out = c;
  return B2_OK;
}

extern "C" void b2_destroy(b2_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  if (ctx->fx_prof) {
    unsigned long long h[4] = {0, 0, 0, 0};
    if (cudaMemcpy(h, ctx->fx_prof, sizeof(h), cudaMemcpyDeviceToHost) == cudaSuccess && h[2])
      fprintf(stderr, "B2_FX_PROF rank %d of %d: %llu exchanges of element 0: %.2f us issuing the pushes, %.2f us waiting for the "
              "last peer's cell (NVLink latency + rank skew)\n", ctx->rank, ctx->nranks, h[2], h[0] / 1e3 / h[2], h[1] / 1e3 / h[2]);
    cudaFree(ctx->fx_prof);
    ctx->fx_prof = nullptr;
  }
  free_ctx(ctx);
}

extern "C" int b2_set_option(b2_ctx* ctx, int option, int value) {
  B2_CHECK_ARG(ctx != nullptr, "b2_set_option: ctx is NULL");
  switch (option) {
    case B2_OPT_FORCE_LAYERED: ctx->opt_force_layered = value ? 1 : 0; return B2_OK;
    case B2_OPT_KERNEL_TIMING: ctx->opt_timing = value ? 1 : 0; return B2_OK;
    case B2_OPT_NO_PEER_EXCHANGE: ctx->opt_no_peer = value ? 1 : 0; return B2_OK;
    case B2_OPT_NO_NARROW_HEAD: ctx->opt_no_narrow_head = value ? 1 : 0; return B2_OK;
    case B2_OPT_NO_SPLIT3: ctx->opt_no_split3 = value ? 1 : 0; return B2_OK;
    case B2_OPT_NO_BUCKET_OVERLAP: ctx->opt_no_overlap = value ? 1 : 0; return B2_OK;
    case B2_OPT_NO_PERSISTENT: ctx->opt_no_persistent = value ? 1 : 0; return B2_OK;
    case B2_OPT_MULTISTEP_FUSED: ctx->opt_multistep_fused = value ? 1 : 0; return B2_OK;
    case B2_OPT_SHARDED_APPLY: ctx->opt_sharded = value ? 1 : 0; return B2_OK;
    case B2_OPT_ASYNC_PERIOD:
      B2_CHECK_ARG(value >= 0, "b2_set_option: async period must be >= 0");
      ctx->async_period = value;
      return B2_OK;
    default: b2_set_error("b2_set_option: unknown option %d", option); return B2_ERR_ARG;
  }
}

extern "C" long b2_param_count(const b2_ctx* ctx) {
  if (!ctx) return 0;
  long p = 0;
  for (int l = 0; l < ctx->n_layers; ++l) p += w_size(ctx, l) + b_size(ctx, l);
  return p;
}

extern "C" float b2_global_step(const b2_ctx* ctx) { return ctx ? ctx->global_step_host : 0.f; }
extern "C" long b2_launch_count(const b2_ctx* ctx) { return ctx ? ctx->launches : 0; }

extern "C" const char* b2_kernel_class_name(int k) {
  static const char* names[B2_KERNEL_CLASSES] = {"dense_fwd", "loss_head", "wgrad", "dgrad", "reduce_apply",
                                                 "fused_linear_step", "fused_linear_reduce", "allreduce_apply",
                                                 "gemm_fwd_tcgen05", "gemm_dgrad_tcgen05", "gemm_wgrad_tcgen05",
                                                 "mixed_aux", "persistent_steps"};
  return (k >= 0 && k < B2_KERNEL_CLASSES) ? names[k] : "?";
}

extern "C" int b2_kernel_times(b2_ctx* ctx, float* ms_total, long* launches) {
  B2_CHECK_ARG(ctx && ms_total && launches, "b2_kernel_times: NULL argument");
  for (int k = 0; k < B2_KERNEL_CLASSES; ++k) { ms_total[k] = 0.f; launches[k] = 0; }
  B2_CUDA(cudaSetDevice(ctx->device));
  for (int i = 0; i < ctx->n_timed; ++i) {
    float ms = 0.f;
    B2_CUDA(cudaEventSynchronize(ctx->timed[i].b));
    B2_CUDA(cudaEventElapsedTime(&ms, ctx->timed[i].a, ctx->timed[i].b));
    ms_total[ctx->timed[i].cls] += ms;
    launches[ctx->timed[i].cls] += 1;
  }
  ctx->n_timed = 0;
  return B2_OK;
}

// ---------------------------------------------------------------------------------------------
// parameters
// ---------------------------------------------------------------------------------------------
extern "C" int b2_params_set(b2_ctx* ctx, int layer, const float* W_dev, const float* b_dev,
                             b2_stream_t stream) {
  B2_CHECK_ARG(ctx != nullptr, "b2_params_set: ctx is NULL");
  B2_CHECK_ARG(layer >= 0 && layer < ctx->n_layers, "b2_params_set: layer %d out of range", layer);
  B2_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t s = (cudaStream_t)stream;
  if (W_dev)
    B2_CUDA(cudaMemcpyAsync(ctx->params + ctx->w_off[layer], W_dev, (size_t)w_size(ctx, layer) * sizeof(float),
                            cudaMemcpyDefault, s));
  if (b_dev)
    B2_CUDA(cudaMemcpyAsync(ctx->params + ctx->b_off[layer], b_dev, (size_t)b_size(ctx, layer) * sizeof(float),
                            cudaMemcpyDefault, s));
  ctx->wb_stale = 1;
  return B2_OK;
}

extern "C" int b2_params_get(b2_ctx* ctx, int layer, float* W_dev, float* b_dev, b2_stream_t stream) {
  B2_CHECK_ARG(ctx != nullptr, "b2_params_get: ctx is NULL");
  B2_CHECK_ARG(layer >= 0 && layer < ctx->n_layers, "b2_params_get: layer %d out of range", layer);
  B2_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t s = (cudaStream_t)stream;
  if (W_dev)
    B2_CUDA(cudaMemcpyAsync(W_dev, ctx->params + ctx->w_off[layer], (size_t)w_size(ctx, layer) * sizeof(float),
                            cudaMemcpyDefault, s));
  if (b_dev)
    B2_CUDA(cudaMemcpyAsync(b_dev, ctx->params + ctx->b_off[layer], (size_t)b_size(ctx, layer) * sizeof(float),
                            cudaMemcpyDefault, s));
  return B2_OK;
}

// ---------------------------------------------------------------------------------------------
// data parallel
// ---------------------------------------------------------------------------------------------
extern "C" int b2_get_unique_id(void* out128) {
  B2_CHECK_ARG(out128 != nullptr, "b2_get_unique_id: out is NULL");
  int rc = load_nccl();
  if (rc) return rc;
  nccl_uid id;
  B2_NCCL(g_nccl.GetUniqueId(&id));
  memcpy(out128, &id, sizeof(id));
  return B2_OK;
}

extern "C" int b2_comm_init(b2_ctx* ctx, int rank, int nranks, const void* unique_id128) {
  B2_CHECK_ARG(ctx != nullptr, "b2_comm_init: ctx is NULL");
  B2_CHECK_ARG(nranks >= 1 && rank >= 0 && rank < nranks, "b2_comm_init: bad rank %d / nranks %d", rank, nranks);
  if (ctx->nccl_comm) { b2_set_error("b2_comm_init: communicator already initialised"); return B2_ERR_STATE; }
  ctx->rank = rank;
  ctx->nranks = nranks;
  if (nranks == 1) return B2_OK;
  B2_CHECK_ARG(unique_id128 != nullptr, "b2_comm_init: unique id is NULL");
  int rc = load_nccl();
  if (rc) { ctx->rank = 0; ctx->nranks = 1; return rc; }
  B2_CUDA(cudaSetDevice(ctx->device));
  nccl_uid id;
  memcpy(&id, unique_id128, sizeof(id));
  int r = g_nccl.CommInitRank(&ctx->nccl_comm, nranks, id, rank);
  if (r != 0) {
    b2_set_error("ncclCommInitRank failed: %s", g_nccl.GetErrorString(r));
    ctx->nccl_comm = nullptr; ctx->rank = 0; ctx->nranks = 1;
    return B2_ERR_NCCL;
  }
  if (is_mixed(ctx)) {
    B2_CUDA(cudaStreamCreateWithFlags(&ctx->comm_stream, cudaStreamNonBlocking));
    B2_CUDA(cudaEventCreateWithFlags(&ctx->ev_bucket, cudaEventDisableTiming));
    B2_CUDA(cudaEventCreateWithFlags(&ctx->ev_comm_done, cudaEventDisableTiming));
  }
  return B2_OK;
}

extern "C" int b2_matmul_bf16(const void* A, int a_mn, const void* B, int b_mn, int M, int N, int K, int epi,
                              const float* bias, const void* act, void* out, b2_stream_t stream) {
  B2_CHECK_ARG(A && B && out, "b2_matmul_bf16: NULL matrix");
  B2_CHECK_ARG(M >= 1 && N >= 1 && K >= 1, "b2_matmul_bf16: bad shape %d x %d x %d", M, N, K);
  B2_CHECK_ARG(epi >= 0 && epi <= 3, "b2_matmul_bf16: unknown epilogue %d", epi);
  B2_CHECK_ARG((epi != 0 && epi != 3) || bias, "b2_matmul_bf16: this epilogue needs a bias vector");
  B2_CHECK_ARG(epi != 1 || act, "b2_matmul_bf16: this epilogue needs the activations");
  int dev = 0, sms = 0;
  B2_CUDA(cudaGetDevice(&dev));
  B2_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  return b2_launch_gemm_bf16(A, a_mn, B, b_mn, M, N, K, epi, bias, act, out, sms, (cudaStream_t)stream, /*local_operands=*/0);
}

extern "C" int b2_peer_export(b2_ctx* ctx, void* handle_out64) {
  B2_CHECK_ARG(ctx && handle_out64, "b2_peer_export: NULL argument");
  B2_CHECK_ARG(ctx->nranks > 1, "b2_peer_export: call b2_comm_init with nranks > 1 first");
  B2_CHECK_ARG(ctx->nranks <= B2_MAX_PEERS, "b2_peer_export: at most %d ranks", B2_MAX_PEERS);
  B2_CHECK_ARG(ctx->P + 4 <= (1L << 18), "b2_peer_export: parameter vector too large for the one-shot exchange "
               "(%ld floats); the NCCL path is used", ctx->P + 4);
  B2_CUDA(cudaSetDevice(ctx->device));
  if (!ctx->peer_local) {
    // cells [2 parities][B2_MAX_PEERS sources][stride] of 8 bytes (b2_fx_args); one stride serves the fused 784->10
    // kernel's element numbering and the flat layout of the layered path
    ctx->peer_stride = round4(ctx->P + 4) > B2_FX_SLOT ? round4(ctx->P + 4) : B2_FX_SLOT;
    const size_t bytes = (size_t)(2 * B2_MAX_PEERS) * (size_t)ctx->peer_stride * 8;
    B2_CUDA(cudaMalloc((void**)&ctx->peer_local, bytes));
    B2_CUDA(cudaMemset(ctx->peer_local, 0, bytes));
  }
  cudaIpcMemHandle_t h;
  B2_CUDA(cudaIpcGetMemHandle(&h, ctx->peer_local));
  static_assert(sizeof(h) == B2_PEER_HANDLE_BYTES, "IPC handle size");
  memcpy(handle_out64, &h, sizeof(h));
  return B2_OK;
}

extern "C" int b2_peer_import(b2_ctx* ctx, const void* handles) {
  B2_CHECK_ARG(ctx && handles, "b2_peer_import: NULL argument");
  if (!ctx->peer_local) { b2_set_error("b2_peer_import: call b2_peer_export first"); return B2_ERR_STATE; }
  B2_CUDA(cudaSetDevice(ctx->device));
  for (int p = 0; p < ctx->nranks; ++p) {
    if (p == ctx->rank) { ctx->peer_base[p] = ctx->peer_local; continue; }
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char*)handles + (size_t)p * B2_PEER_HANDLE_BYTES, sizeof(h));
    void* ptr = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      b2_set_error("b2_peer_import: cannot map rank %d's buffer: %s", p, cudaGetErrorString(e));
      (void)cudaGetLastError();
      return B2_ERR_CUDA;
    }
    ctx->peer_base[p] = ptr;
  }
  ctx->peer_ready = 1;
  const char* env = getenv("B2_FX_PROF");
  if (env && *env == '1' && !ctx->fx_prof) {
    B2_CUDA(cudaMalloc((void**)&ctx->fx_prof, 4 * sizeof(unsigned long long)));
    B2_CUDA(cudaMemset(ctx->fx_prof, 0, 4 * sizeof(unsigned long long)));
  }
  return B2_OK;
}

// One data-parallel step as every public entry point runs it (b2_train_steps and b2_session_run_train must agree):
// synchronous mode divides by the global batch and exchanges gradients inside enqueue_step; with B2_OPT_ASYNC_PERIOD
// the step is local (mean over this rank's rows) and the replicas are averaged every `async_period` local steps.
static int enqueue_step_dp(b2_ctx* ctx, const float* x, const float* y, int rows, float lr, float* loss_trace,
                           int* correct_trace, cudaStream_t s) {
  const bool async = ctx->async_period > 0 && ctx->nranks > 1;
  int rc = enqueue_step(ctx, x, y, rows, async ? (long)rows : (long)rows * ctx->nranks, lr, loss_trace, correct_trace, s);
  if (rc) return rc;
  if (async) {
    // the workers of the reference update the ps variables independently; here the replicas drift for
    // `async_period` local steps and are then averaged (parameters, not gradients)
    ctx->global_step_host += (float)(ctx->nranks - 1);       // global_step counts every worker's update (distributed.py:108)
    if (++ctx->async_count % ctx->async_period == 0) {
      B2_NCCL(g_nccl.AllReduce(ctx->params, ctx->params, (size_t)ctx->P, kNcclFloat, kNcclSum, ctx->nccl_comm, s));
      rc = b2_launch_scale(ctx->params, ctx->P, 1.0f / (float)ctx->nranks, ctx->sm_count, s);
      if (rc) return rc;
      ctx->launches++;
      ctx->wb_stale = 1;
    }
  }
  return B2_OK;
}

// All n_steps in ONE launch where a persistent kernel exists for the model: the reference's hot loops (simple.py:34-35,
// distributed.py:145-150) run inside the kernel.  Returns B2_NOT_PERSISTENT when the per-step path must be used.
constexpr int B2_NOT_PERSISTENT = -1;
static int enqueue_persistent(b2_ctx* c, const float* xp, const float* yp, long span, long first_row, int local_batch,
                              long global_batch, int n_steps, float lr, float* loss_trace, int* correct_trace, cudaStream_t s) {
  if (c->opt_no_persistent || c->opt_force_layered || c->precision != B2_PREC_FP32 || n_steps < 1) return B2_NOT_PERSISTENT;
  if (c->nranks > 1 && c->async_period > 0) return B2_NOT_PERSISTENT;
  if (c->nranks == 1 && b2_tiny_steps_applies(c->n_layers, c->dims, local_batch)) {
    timed_scope t(c, B2_K_PERSISTENT_STEPS, s);
    int rc = b2_launch_tiny_steps(c->n_layers, c->dims, c->act, c->loss, c->w_off, c->b_off, c->P, xp, yp, span, first_row,
                                  global_batch, local_batch, n_steps, lr, c->params, c->grads, loss_trace, correct_trace,
                                  c->global_step_dev, s);
    if (rc) return rc;
    c->launches += 1;
    c->global_step_host += (float)n_steps;
    return B2_OK;
  }
  if (use_fused_linear(c) && c->opt_multistep_fused) {
    const bool use_peer = c->nranks > 1 && c->peer_ready && !c->opt_no_peer;
    if (c->nranks > 1 && !use_peer) return B2_NOT_PERSISTENT;       // NCCL exchange: one launch sequence per step
    b2_fx_args fx;
    memset(&fx, 0, sizeof(fx));
    if (use_peer) {
      fx.rank = c->rank; fx.nranks = c->nranks; fx.stride = c->peer_stride;
      fx.step1 = c->peer_step + 1;
      fx.parity = (int)(c->peer_step & 1u);
      for (int p = 0; p < c->nranks; ++p) fx.region[p] = (unsigned long long*)c->peer_base[p];
      fx.prof = c->fx_prof;
    }
    const int tr = b2_fused_linear_tile_rows();
    int grid = (local_batch + tr - 1) / tr;
    if (grid > c->fused_grid) grid = c->fused_grid;
    timed_scope t(c, B2_K_PERSISTENT_STEPS, s);
    int rc = b2_launch_fused_linear_multistep(xp, yp, span, first_row, global_batch, (long)c->rank * local_batch, local_batch,
                                              n_steps, c->params, c->grads, c->fused_scratch, c->grid_bar, grid, lr, c->loss,
                                              loss_trace, correct_trace, c->global_step_dev, c->P, use_peer ? &fx : nullptr, s);
    if (rc) return rc;
    if (use_peer) c->peer_step += (unsigned)n_steps;
    c->launches += 1;
    c->global_step_host += (float)n_steps;
    return B2_OK;
  }
  {
    // small batches of the 784-wide models: one cluster of 14 CTAs runs the whole loop (cluster_steps.cu); with several
    // workers the ranks' gradients are exchanged inside the kernel through the NVLink peer cells
    const bool use_peer = c->nranks > 1 && c->peer_ready && !c->opt_no_peer;
    if ((c->nranks == 1 || use_peer) && c->cs_scratch && b2_cluster_steps_applies(c->n_layers, c->dims, local_batch)) {
      b2_fx_args fx;
      memset(&fx, 0, sizeof(fx));
      if (use_peer) {
        fx.rank = c->rank; fx.nranks = c->nranks; fx.stride = c->peer_stride;
        fx.step1 = c->peer_step + 1;
        fx.parity = (int)(c->peer_step & 1u);
        for (int p = 0; p < c->nranks; ++p) fx.region[p] = (unsigned long long*)c->peer_base[p];
      }
      timed_scope t(c, B2_K_PERSISTENT_STEPS, s);
      int rc = b2_launch_cluster_steps(c->n_layers, c->act, c->loss, c->w_off, c->b_off, c->P, xp, yp, span, first_row,
                                       global_batch, local_batch, n_steps, lr, c->params, c->grads, loss_trace, correct_trace,
                                       c->global_step_dev, c->cs_scratch, (long)c->rank * local_batch, use_peer ? &fx : nullptr, s);
      if (rc == -2) return B2_NOT_PERSISTENT;      // this device cannot co-schedule the 14-CTA cluster: per-step kernels
      if (rc) return rc;
      if (use_peer) c->peer_step += (unsigned)n_steps;
      c->launches += 1;
      c->global_step_host += (float)n_steps;
      return B2_OK;
    }
  }
  return B2_NOT_PERSISTENT;
}

// ---------------------------------------------------------------------------------------------
// the hot path
// ---------------------------------------------------------------------------------------------
extern "C" int b2_train_steps(b2_ctx* ctx, const void* x_pool_dev, const void* y_pool_dev, long pool_rows,
                              int local_batch, long global_batch, int n_steps, long first_row, float lr,
                              float* loss_trace_dev, int* correct_trace_dev, b2_stream_t stream) {
  B2_CHECK_ARG(ctx != nullptr, "b2_train_steps: ctx is NULL");
  B2_CHECK_ARG(x_pool_dev && y_pool_dev, "b2_train_steps: NULL data pool");
  // rows are fetched with bulk / TMA / 128-bit loads: a view at an odd float offset must be rejected here, not trap on the GPU
  B2_CHECK_ARG((reinterpret_cast<uintptr_t>(x_pool_dev) & 15u) == 0 || ctx->dims[0] % 4 != 0,
               "b2_train_steps: x_pool_dev must be 16-byte aligned");
  B2_CHECK_ARG(((reinterpret_cast<uintptr_t>(x_pool_dev) | reinterpret_cast<uintptr_t>(y_pool_dev)) & 3u) == 0,
               "b2_train_steps: the data pools must be float-aligned");
  B2_CHECK_ARG(local_batch >= 1 && local_batch <= ctx->max_local_batch,
               "b2_train_steps: local_batch %d outside [1, max_local_batch=%d]", local_batch, ctx->max_local_batch);
  B2_CHECK_ARG(global_batch == (long)local_batch * ctx->nranks,
               "b2_train_steps: global_batch %ld != local_batch %d * nranks %d", global_batch, local_batch, ctx->nranks);
  B2_CHECK_ARG(pool_rows >= global_batch, "b2_train_steps: pool_rows %ld < global_batch %ld", pool_rows, global_batch);
  B2_CHECK_ARG(n_steps >= 0, "b2_train_steps: n_steps must be >= 0");
  const long span = (pool_rows / global_batch) * global_batch;
  B2_CHECK_ARG(first_row >= 0 && first_row % global_batch == 0 && first_row < span,
               "b2_train_steps: first_row %ld must be a multiple of global_batch inside the pool", first_row);
  B2_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t s = (cudaStream_t)stream;
  const float* xp = (const float*)x_pool_dev;
  const float* yp = (const float*)y_pool_dev;
  const int in = ctx->dims[0], outd = ctx->dims[ctx->n_layers];
  {
    int rc = enqueue_persistent(ctx, xp, yp, span, first_row, local_batch, global_batch, n_steps, lr, loss_trace_dev,
                                correct_trace_dev, s);
    if (rc != B2_NOT_PERSISTENT) return rc;
  }
  for (int i = 0; i < n_steps; ++i) {
    const long start = (first_row + (long)i * global_batch) % span + (long)ctx->rank * local_batch;
    int rc = enqueue_step_dp(ctx, xp + start * in, yp + start * outd, local_batch, lr,
                             loss_trace_dev ? loss_trace_dev + i : nullptr,
                             correct_trace_dev ? correct_trace_dev + i : nullptr, s);
    if (rc) return rc;
  }
  return B2_OK;
}

extern "C" int b2_eval(b2_ctx* ctx, const void* x_dev, const void* y_dev, long rows, float* loss_dev,
                       long* correct_dev, b2_stream_t stream) {
  B2_CHECK_ARG(ctx != nullptr, "b2_eval: ctx is NULL");
  B2_CHECK_ARG(x_dev && y_dev && rows >= 1, "b2_eval: bad arguments");
  B2_CHECK_ARG((reinterpret_cast<uintptr_t>(x_dev) & 15u) == 0 || ctx->dims[0] % 4 != 0, "b2_eval: x_dev must be 16-byte aligned");
  B2_CHECK_ARG(((reinterpret_cast<uintptr_t>(x_dev) | reinterpret_cast<uintptr_t>(y_dev)) & 3u) == 0,
               "b2_eval: x_dev and y_dev must be float-aligned");
  B2_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t s = (cudaStream_t)stream;
  const float* xp = (const float*)x_dev;
  const float* yp = (const float*)y_dev;
  const int in = ctx->dims[0], outd = ctx->dims[ctx->n_layers];
  const float inv_B = 1.0f / (float)rows;
  B2_CUDA(cudaMemsetAsync(ctx->eval_acc_dev, 0, 2 * sizeof(float), s));
  int dummy[B2_MAX_LAYERS];
  for (long r0 = 0; r0 < rows; r0 += ctx->max_local_batch) {
    const int n = (int)((rows - r0 < ctx->max_local_batch) ? rows - r0 : ctx->max_local_batch);
    int rc = is_mixed(ctx) ? enqueue_forward_mixed(ctx, xp + r0 * in, yp + r0 * outd, n, inv_B, s)
                           : enqueue_layered(ctx, xp + r0 * in, yp + r0 * outd, n, inv_B, 0, dummy, s);
    if (rc) return rc;
    float* head = ctx->scratch + ctx->head_part_off;
    rc = b2_launch_eval_accumulate(head, head + ctx->head_blocks, is_mixed(ctx) ? ctx->head_blocks : ctx->head_used,
                                   ctx->eval_acc_dev, s);
    if (rc) return rc;
    ctx->launches++;
  }
  int rc = b2_launch_eval_finish(ctx->eval_acc_dev, loss_dev, correct_dev, s);
  if (rc) return rc;
  ctx->launches++;
  return B2_OK;
}

extern "C" int b2_session_run_train(b2_ctx* ctx, const float* x_host, const float* y_host, int rows, float lr,
                                    float* loss_out, int* correct_out) {
  B2_CHECK_ARG(ctx != nullptr, "b2_session_run_train: ctx is NULL");
  B2_CHECK_ARG(x_host && y_host, "b2_session_run_train: NULL feed");
  B2_CHECK_ARG(rows >= 1 && rows <= ctx->max_local_batch,
               "b2_session_run_train: rows %d outside [1, max_local_batch=%d]", rows, ctx->max_local_batch);
  B2_CUDA(cudaSetDevice(ctx->device));
  const int in = ctx->dims[0], outd = ctx->dims[ctx->n_layers];
  if (!ctx->feed_x) {
    B2_CUDA(cudaMalloc((void**)&ctx->feed_x, (size_t)ctx->max_local_batch * in * sizeof(float)));
    B2_CUDA(cudaMalloc((void**)&ctx->feed_y, (size_t)ctx->max_local_batch * outd * sizeof(float)));
  }
  cudaStream_t s = ctx->own_stream;
  B2_CUDA(cudaMemcpyAsync(ctx->feed_x, x_host, (size_t)rows * in * sizeof(float), cudaMemcpyHostToDevice, s));
  B2_CUDA(cudaMemcpyAsync(ctx->feed_y, y_host, (size_t)rows * outd * sizeof(float), cudaMemcpyHostToDevice, s));
  // the same kernels b2_train_steps would use for this model and batch size, so both entry points give the same bits
  int rc = ctx->nranks > 1 ? B2_NOT_PERSISTENT
                           : enqueue_persistent(ctx, ctx->feed_x, ctx->feed_y, rows, 0, rows, rows, 1, lr, ctx->fetch_dev,
                                                (int*)(ctx->fetch_dev + 1), s);
  if (rc == B2_NOT_PERSISTENT)
    rc = enqueue_step_dp(ctx, ctx->feed_x, ctx->feed_y, rows, lr, ctx->fetch_dev, (int*)(ctx->fetch_dev + 1), s);
  if (rc) return rc;
  B2_CUDA(cudaMemcpyAsync(ctx->fetch_host, ctx->fetch_dev, 2 * sizeof(float), cudaMemcpyDeviceToHost, s));
  B2_CUDA(cudaStreamSynchronize(s));
  if (loss_out) *loss_out = ctx->fetch_host[0];
  if (correct_out) memcpy(correct_out, &ctx->fetch_host[1], sizeof(int));
  return B2_OK;
}
