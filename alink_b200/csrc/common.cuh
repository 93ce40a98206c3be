// common.cuh — context, error plumbing and device scratch for libalink_b200.so
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <stdarg.h>
#include <stdlib.h>

#include "../../include/alink_b200.h"
#include "../../include/alink_math.h"

struct DevBuf {
  void *p = nullptr;
  size_t cap = 0;
};

struct HeadGroup {
  bool set = false;
  int K = 0, D_in = 0, H1 = 0, H2 = 0, C = 0, act = 0, feat_mode = 0, compute = 0;
  // fp32 master copies, one allocation per tensor kind: [K][in][out] contiguous
  float *W1 = nullptr, *b1 = nullptr, *W2 = nullptr, *b2 = nullptr, *W3 = nullptr, *b3 = nullptr;
  // device arrays of K pointers into the above
  const float **pW1 = nullptr, **pb1 = nullptr, **pW2 = nullptr, **pb2 = nullptr, **pW3 = nullptr, **pb3 = nullptr;
  // bf16 operands for the tcgen05 path, N-major ("[out][in]", K contiguous): W1t [K][H1][D_in], W2t [K][H2][H1]
  __nv_bfloat16 *W1t = nullptr, *W2t = nullptr;
  // the same images in IEEE half (FP32_REFINE's default filter format)
  __half *W1h = nullptr, *W2h = nullptr;
  // scaled hi / lo half images for the split-operand second stage of FP32_REFINE (head_tc_split.cuh)
  __half *S1hi = nullptr, *S1lo = nullptr, *S2hi = nullptr, *S2lo = nullptr;
  float sw1 = 1.0f, sw2 = 1.0f;  // their power-of-two scales
  bool split_ok = false;
  bool tc_ok = false;  // the tensor-core images exist (shape supported by head_tc.cuh)
  // bound for "no half overflow": max_j sum_k |W1[k][j]| over learners, max |b1| (host, at alink_set_heads)
  double w1_col_l1 = 0.0, b1_absmax = 0.0, w_absmax = 0.0;  // w_absmax: max |W2|
  double w1_absmax = 0.0;
  // a-priori (statistical) error model of the half-precision filter, per group (alink_set_heads; refine_apriori_eps):
  // max column L2 norms of W1 / W2, max |W1| / |W2| element, spectral norm of W1 (power iteration), ||b1||_2,
  // and max over learners of ||W3[:,1] - W3[:,0]||_2 (or ||W3||_2 for the sigmoid head)
  double w1_col_l2 = 0.0, w2_col_l2 = 0.0, w1_spec = 0.0, b1_l2 = 0.0, w3_gain = 0.0;
  double w3_gain_ms = 0.0, b1_ms = 0.0;  // mean over learners of ||w3d||^2; max over learners of mean b1^2
  bool bounds_stale = false;  // weights changed on the device (train.cuh) since those bounds were computed
  // Adadelta accumulators of the finetune step (train.cuh), same [K][...] layout as the six tensors; lazy, zeroed
  float *opt_a[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  float *opt_d[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
};

struct Table {
  const void *p = nullptr;  // device pointer (owned iff own.p == p)
  int64_t M = 0;
  int D = 0;
  int dtype = ALINK_F32;
  DevBuf own;
  double absmax = -1.0;  // max |element| (lazy, refine.cuh; reset by alink_set_table)
  double rownorm = -1.0; // max row L2 norm (same pass)
  double rowms = 0.0;    // mean row sum of squares (same pass)
  cudaEvent_t ready = nullptr;  // alink_set_table_async: recorded on the copy stream behind the upload
  bool pending = false;         // the compute stream has not yet been made to wait for `ready`
};

struct SelectState;  // topk.cuh

// batched differential evolution (de.cuh): P problems x n members x G genes, device resident
struct DeState {
  int P = 0, n = 0, G = 0;     // problems, members, genes
  int slot_left = 0, slot_right = 0;
  DevBuf pop, trial;           // double [P][n][G], in [0, 1]
  DevBuf e_pop, e_trial;       // float  [P][n]
  DevBuf viol;                 // uint32 [P][n][G] flags of out-of-range genes, then [P] counts
  DevBuf rows, ia, ib;         // perturbed left rows [P*n][D] and the identity / right-row index arrays
  DevBuf consts;               // il[P] ir[P] targets[P] (int32) | lo[G] hi[G] (double) | active[P] (uint8)
  DevBuf aux;                  // uploaded scales / samples / crossover masks / repair values
  bool evaluated = false;      // the initial population has energies
};

struct alink_ctx {
  int device = 0;
  int sm_count = 148;
  int reserved_sms = 0;               // SMs the persistent kernels leave free (alink_set_reserved_sms)
  cudaStream_t stream = nullptr;      // stream in use
  cudaStream_t own_stream = nullptr;  // private stream created by alink_create
  cudaStream_t copy_stream = nullptr; // alink_set_table_async uploads (created on first use)
  cudaEvent_t copy_gate = nullptr;    // compute-stream marker the copy stream waits for before overwriting a slot
  cudaEvent_t copy_done = nullptr;    // behind the latest asynchronous upload
  bool copy_busy = false;             // an asynchronous upload may still be in flight (cleared when copy_done has passed)
  cudaEvent_t ev[8] = {};
  char err[512] = {0};
  int64_t launches = 0;
  unsigned flags = 0;                 // ALINK_FLAG_* of alink_create

  HeadGroup groups[ALINK_MAX_GROUPS];
  Table tables[ALINK_MAX_SLOTS];

  // last scoring result (device resident)
  int64_t N = 0;
  int sK = 0, sC = 0;
  DevBuf idx_a, idx_b;   // staged pair indices (when given on host)
  DevBuf h1ws;           // fp32 hidden-1 workspace of one chunk [K][Mc][H1]
  DevBuf probs;          // [N][K][C] fp32
  DevBuf cons;           // [N][C] fp32
  DevBuf util;           // [N] fp32
  bool util_valid = false;
  int util_measure = -1;  // measure ctx->util holds
  // select scratch
  DevBuf sel_state, sel_keys, sel_idx, sel_util, sel_in, sel_hist;
  DevBuf misc0, misc1, misc2, misc3;
  // FP32_REFINE (refine.cuh)
  double refine_eps = 0.0;        // caller's bound on |filter probability - fp32 probability| (0 = per-format default)
  int64_t refine_survivors = 0;   // candidates re-scored exactly by the last alink_query
  float refine_max_err = 0.0f;    // max |approx - exact| utility over those candidates
  float refine_rms_err = 0.0f;    // rms of the same
  double refine_eps_used = 0.0;   // bound the last query ran with (after any widening)
  int refine_rounds = 0;          // mark + re-score rounds of the last query (> 1: the bound had to be widened)
  int refine_f16 = 0;             // the last query's filter ran on half (1) or bf16 (0) operands
  bool probs_full = true;         // ctx->probs holds all N pairs (false after a refine query)
  DevBuf cand, cia, cib, util_c, rstat;
  DevBuf util2, probs2, cia2, cib2, map2, src2;     // stage 2 / 3 of refine_query_v2
  unsigned int *rs_host = nullptr;          // pinned copy of the statistics words (read with the result)
  // deferred verification (alink_query_sharded): the three-stage query is enqueued without its host synchronisation, its
  // verdict is evaluated on the device (refine_verify_kernel) and travels with the keys; the report is filled afterwards
  bool defer_req = false, deferred = false;
  double defer_eps1 = 0.0, defer_eps2 = 0.0, defer_scv = 1.0;
  unsigned int *defer_ok_dev = nullptr;     // device address of the verdict word of the deferred query
  unsigned int *comm_status = nullptr;      // device: [1] = every rank's verdict ok (alink_query_sharded)
  unsigned int *comm_host = nullptr;        // pinned mirror
  double refine_eps2 = 0.0, refine_eps2_used = 0.0;  // stage-2 (split operand) bound: caller's / used
  double refine_eps_apriori = 0.0;          // a-priori statistical MODEL of the stage-1 filter error (8 sigma) for the last query
  double refine_eps_worst = 0.0;            // the same propagated with worst-case norms (reported, too loose to filter with)
  double refine_eps_hint[ALINK_MAX_GROUPS] = {0, 0, 0, 0};  // bound a previous query of the group had to widen to
  float refine_err1 = 0.0f, refine_err2 = 0.0f, refine_rms2 = 0.0f;  // max |u~ - u2| over C1 u S; max / rms |u2 - u| over C2 u S2
  int64_t refine_n1 = 0, refine_nsample = 0, refine_nsample2 = 0;
  int refine_fallback = 0;                  // why the last query left the three-stage pipeline (alink_refine_report.fallback)
  int refine_version = 0;                   // pipeline the last query ran (2 = three stages, 1 = filter + FFMA)
  // all-pairs enumeration (alink_*_allpairs): pair n = (probe row (ap_pair0 + n) / ap_G, gallery row (..) % ap_G)
  uint32_t ap_G = 0;
  int64_t ap_pair0 = 0;
  DevBuf xscratch;  // head_tc_ko.cuh: per-CTA bf16 feature-tile images
  DevBuf train_ws, train_y, train_out;  // train.cuh: activations / gradients of one minibatch
  unsigned long long *wd_host = nullptr;  // host-mapped watchdog record (ptx.cuh: mbar_timeout)
  DeState de;
  // cross-shard exchange (comm.cuh): an NCCL communicator owned by the context
  void *comm = nullptr;
  int comm_rank = 0, comm_world = 1;
  DevBuf comm_keys;
  // peer push (comm.cuh): the windows of all ranks as mapped here (IPC), this rank's own allocation, the step counter
  uint8_t *push_win[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  uint8_t *push_own = nullptr;
  bool push_ok = false;
  unsigned long long push_epoch = 0;
  unsigned int *push_status = nullptr;  // device: [0] timeout, [1] all ranks ok, [2] this rank's ok bit
  unsigned int *push_host = nullptr;    // pinned mirror of push_status
  int push_used = 0;                    // 1: the last alink_query_sharded exchanged through the push kernel
};

static inline int alink_fail(alink_ctx *ctx, int code, const char *fmt, ...) {
  if (ctx) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(ctx->err, sizeof(ctx->err), fmt, ap);
    va_end(ap);
  }
  return code;
}

#define ALINK_CUDA(ctx, expr)                                                                     \
  do {                                                                                            \
    cudaError_t e__ = (expr);                                                                     \
    if (e__ != cudaSuccess)                                                                       \
      return alink_fail((ctx), ALINK_E_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #expr,         \
                        cudaGetErrorString(e__));                                                 \
  } while (0)

#define ALINK_TRY(expr)            \
  do {                             \
    int rc__ = (expr);             \
    if (rc__ != ALINK_OK) return rc__; \
  } while (0)

// count + check a kernel launch
#define ALINK_LAUNCHED(ctx)                                                                       \
  do {                                                                                            \
    (ctx)->launches++;                                                                            \
    cudaError_t e__ = cudaGetLastError();                                                         \
    if (e__ != cudaSuccess)                                                                       \
      return alink_fail((ctx), ALINK_E_CUDA, "%s:%d kernel launch -> %s", __FILE__, __LINE__,     \
                        cudaGetErrorString(e__));                                                 \
  } while (0)

static inline int dev_reserve(alink_ctx *ctx, DevBuf &b, size_t bytes) {
  if (bytes <= b.cap && b.p) return ALINK_OK;
  if (b.p) {
    ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ALINK_CUDA(ctx, cudaFree(b.p));
    b.p = nullptr;
    b.cap = 0;
  }
  size_t want = bytes < 256 ? 256 : bytes;
  cudaError_t e = cudaMalloc(&b.p, want);
  if (e != cudaSuccess) {
    b.p = nullptr;
    return alink_fail(ctx, ALINK_E_NOMEM, "cudaMalloc(%zu) -> %s", want, cudaGetErrorString(e));
  }
  b.cap = want;
  return ALINK_OK;
}

static inline void dev_release(DevBuf &b) {
  if (b.p) cudaFree(b.p);
  b.p = nullptr;
  b.cap = 0;
}

// Bring a caller buffer onto the device: ALINK_DEVICE pointers pass through, ALINK_HOST ones are
// copied into `stage` on the context stream.
// reads page-locked HOST memory straight over PCIe (16 bytes per thread and step)
__global__ void __launch_bounds__(256) zero_copy_in_kernel(const uint4 *__restrict__ src, uint4 *__restrict__ dst, size_t n16) {
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n16; i += (size_t)gridDim.x * 256) dst[i] = src[i];
}

static inline int stage_in(alink_ctx *ctx, const void *src, size_t bytes, int space, DevBuf &stage,
                           const void **out) {
  if (!src) {
    *out = nullptr;
    return ALINK_OK;
  }
  if (space == ALINK_DEVICE) {
    *out = src;
    return ALINK_OK;
  }
  ALINK_TRY(dev_reserve(ctx, stage, bytes));
  // Page-locked inputs are read by a kernel straight over PCIe instead of a cudaMemcpyAsync: a copy would queue on the
  // host-to-device copy engine behind any table upload in flight (alink_set_table_async, or the caller's own side
  // stream: alink_b200.dist.ShardedTableFeed), serialising the step behind it; the kernel shares the link instead.
  static const int zc_env = [] { const char *e = getenv("ALINK_ZERO_COPY_IN"); return e ? atoi(e) : 1; }();
  if ((zc_env == 1 || ctx->copy_busy) && zc_env != 2 && bytes % 16 == 0 && ((uintptr_t)src & 15u) == 0 && bytes <= (64u << 20)) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, src) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer) {
      const size_t n16 = bytes / 16;
      size_t blocks = (n16 + 255) / 256;
      if (blocks > (size_t)ctx->sm_count * 4) blocks = (size_t)ctx->sm_count * 4;
      if (blocks < 1) blocks = 1;
      zero_copy_in_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>((const uint4 *)at.devicePointer, (uint4 *)stage.p, n16);
      ALINK_LAUNCHED(ctx);
      *out = stage.p;
      return ALINK_OK;
    }
    cudaGetLastError();  // pageable memory: not an error, take the copy-engine path
  }
  ALINK_CUDA(ctx, cudaMemcpyAsync(stage.p, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
  *out = stage.p;
  return ALINK_OK;
}

static inline int copy_out(alink_ctx *ctx, void *dst, const void *src_dev, size_t bytes, int space) {
  if (!dst || bytes == 0) return ALINK_OK;
  if (space == ALINK_DEVICE && bytes % 16 == 0 && bytes <= (1u << 20) && (((uintptr_t)dst | (uintptr_t)src_dev) & 15u) == 0) {
    // small device-to-device results (the B keys of a shard): a kernel, not a copy-engine job that could queue behind an upload
    const size_t n16 = bytes / 16;
    size_t blocks = (n16 + 255) / 256;
    if (blocks > 64) blocks = 64;
    zero_copy_in_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>((const uint4 *)src_dev, (uint4 *)dst, n16);
    ALINK_LAUNCHED(ctx);
    return ALINK_OK;
  }
  ALINK_CUDA(ctx, cudaMemcpyAsync(dst, src_dev, bytes,
                                  space == ALINK_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost,
                                  ctx->stream));
  return ALINK_OK;
}
