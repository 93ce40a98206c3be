// alink_b200.cu — C ABI (include/alink_b200.h) over the sm_100a kernels of the query-selection path.
// Build: alink_b200/csrc/Makefile (nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -lineinfo).
// -fmad=false: no implicit contraction anywhere; every fused multiply-add is an explicit __fmaf_rn.
#include <mutex>
#include "common.cuh"
#include "head_ffma.cuh"
#include "head_tc.cuh"
#include "head_tc_ko.cuh"
#include "utility.cuh"
#include "topk.cuh"
#include "disparity.cuh"
#include "head_tc_split.cuh"
#include "refine.cuh"
#include "roc.cuh"
#include "train.cuh"
#include "de.cuh"
#include "perturb.cuh"
#include "comm.cuh"

extern "C" {
static void push_teardown(alink_ctx *ctx);  // peer windows of the push exchange (below, with alink_comm_init)
static int noalloc_resolve(int device);     // device address of the static scratch of alink_topk_values_noalloc (set-up time)
}

#include <math.h>
#include <new>
#include <vector>
#include <algorithm>

extern "C" {

int alink_abi_version(void) { return ALINK_ABI_VERSION; }

int alink_create(alink_ctx **out, int device, unsigned flags) {
  if (!out) return ALINK_E_INVALID;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) return ALINK_E_CUDA;  // no CPU fallback
  if (device < 0 || device >= ndev) return ALINK_E_INVALID;
  alink_ctx *ctx = new (std::nothrow) alink_ctx();
  if (!ctx) return ALINK_E_NOMEM;
  ctx->device = device;
  ctx->flags = flags;
  if (cudaSetDevice(device) != cudaSuccess ||
      cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
    delete ctx;
    return ALINK_E_CUDA;
  }
  ctx->own_stream = ctx->stream;
  noalloc_resolve(device);  // once per device, with that device current; a failure is reported by the entry point that needs it
  cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
  for (int i = 0; i < 8; ++i) cudaEventCreate(&ctx->ev[i]);
  // watchdog record: written by a kernel whose mbarrier wait times out (a protocol deadlock), readable
  // from the host even after the launch failure has poisoned the CUDA context
  if (cudaHostAlloc((void **)&ctx->wd_host, 2048, cudaHostAllocMapped) == cudaSuccess) {
    memset(ctx->wd_host, 0, 2048);
    unsigned long long *dptr = nullptr;
    if (cudaHostGetDevicePointer((void **)&dptr, ctx->wd_host, 0) == cudaSuccess)
      cudaMemcpyToSymbol(ptx::g_alink_wd, &dptr, sizeof(dptr));
  } else {
    ctx->wd_host = nullptr;
  }
  *out = ctx;
  return ALINK_OK;
}

static void free_group(HeadGroup &g) {
  void *ptrs[] = {g.W1, g.b1, g.W2, g.b2, g.W3, g.b3, (void *)g.pW1, g.W1t, g.W2t, g.W1h, g.W2h, g.S1hi, g.S1lo, g.S2hi, g.S2lo};
  for (void *p : ptrs)
    if (p) cudaFree(p);
  for (int t = 0; t < 6; ++t) {
    if (g.opt_a[t]) cudaFree(g.opt_a[t]);
    if (g.opt_d[t]) cudaFree(g.opt_d[t]);
  }
  g = HeadGroup();
}

int alink_destroy(alink_ctx *ctx) {
  if (!ctx) return ALINK_OK;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  for (auto &g : ctx->groups) free_group(g);
  for (auto &t : ctx->tables) dev_release(t.own);
  DevBuf *bufs[] = {&ctx->idx_a, &ctx->idx_b, &ctx->h1ws, &ctx->probs, &ctx->cons, &ctx->util, &ctx->sel_state,
                    &ctx->sel_keys, &ctx->sel_idx, &ctx->sel_util, &ctx->sel_in, &ctx->sel_hist, &ctx->misc0, &ctx->misc1,
                    &ctx->misc2, &ctx->misc3, &ctx->cand, &ctx->cia, &ctx->cib, &ctx->util_c, &ctx->rstat, &ctx->xscratch,
                    &ctx->train_ws, &ctx->train_y, &ctx->train_out, &ctx->util2, &ctx->probs2, &ctx->cia2, &ctx->cib2, &ctx->map2, &ctx->src2};
  for (DevBuf *b : bufs) dev_release(*b);
  DevBuf *debufs[] = {&ctx->de.pop, &ctx->de.trial, &ctx->de.e_pop, &ctx->de.e_trial, &ctx->de.viol, &ctx->de.rows, &ctx->de.ia,
                      &ctx->de.ib, &ctx->de.consts, &ctx->de.aux};
  for (DevBuf *b : debufs) dev_release(*b);
  if (ctx->rs_host) cudaFreeHost(ctx->rs_host);
  if (ctx->comm_status) cudaFree(ctx->comm_status);
  if (ctx->comm_host) cudaFreeHost(ctx->comm_host);
  if (ctx->comm && nccl_api()) nccl_api()->CommDestroy((alink_nccl_comm)ctx->comm);
  push_teardown(ctx);
  dev_release(ctx->comm_keys);
  for (int i = 0; i < 8; ++i)
    if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
  if (ctx->copy_stream) {
    cudaStreamSynchronize(ctx->copy_stream);
    cudaStreamDestroy(ctx->copy_stream);
    cudaEventDestroy(ctx->copy_gate);
    cudaEventDestroy(ctx->copy_done);
  }
  for (auto &t : ctx->tables)
    if (t.ready) cudaEventDestroy(t.ready);
  cudaStreamDestroy(ctx->own_stream);
  delete ctx;
  return ALINK_OK;
}

const char *alink_last_error(const alink_ctx *ctx) { return ctx ? ctx->err : "null context"; }

int alink_sync(alink_ctx *ctx) {
  if (!ctx) return ALINK_E_INVALID;
  ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return ALINK_OK;
}

int alink_set_stream(alink_ctx *ctx, void *cuda_stream) {
  if (!ctx) return ALINK_E_INVALID;
  ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
  return ALINK_OK;
}

int64_t alink_launch_count(const alink_ctx *ctx) { return ctx ? ctx->launches : 0; }

int alink_set_reserved_sms(alink_ctx *ctx, int n) {
  if (!ctx) return ALINK_E_INVALID;
  if (n < 0 || n > ctx->sm_count - 2) return alink_fail(ctx, ALINK_E_INVALID, "reserved SMs outside [0, %d]", ctx->sm_count - 2);
  ctx->reserved_sms = n & ~1;  // the kernels run CTA pairs
  return ALINK_OK;
}

int alink_watchdog_record(const alink_ctx *ctx, uint64_t out[256]) {
  if (!ctx || !out) return ALINK_E_INVALID;
  for (int i = 0; i < 256; ++i) out[i] = ctx->wd_host ? (uint64_t)((volatile unsigned long long *)ctx->wd_host)[i] : 0;
  return ALINK_OK;
}

int alink_host_alloc(void **out, uint64_t bytes) {
  if (!out) return ALINK_E_INVALID;
  return cudaMallocHost(out, bytes ? bytes : 1) == cudaSuccess ? ALINK_OK : ALINK_E_NOMEM;
}
int alink_host_free(void *p) { return (!p || cudaFreeHost(p) == cudaSuccess) ? ALINK_OK : ALINK_E_CUDA; }

// ------------------------------------------------------------------------------------------ heads
static int upload_stack(alink_ctx *ctx, const float *const *src, int K, size_t each, float **dst) {
  ALINK_CUDA(ctx, cudaMalloc((void **)dst, sizeof(float) * each * K));
  for (int k = 0; k < K; ++k)
    ALINK_CUDA(ctx, cudaMemcpyAsync(*dst + each * k, src[k], sizeof(float) * each, cudaMemcpyHostToDevice,
                                    ctx->stream));
  return ALINK_OK;
}

int alink_set_heads(alink_ctx *ctx, int group, int K, int D_in, int H1, int H2, int C, int out_act,
                    int feat_mode, const float *const *W1, const float *const *b1, const float *const *W2,
                    const float *const *b2, const float *const *W3, const float *const *b3, int compute) {
  if (!ctx) return ALINK_E_INVALID;
  if (group < 0 || group >= ALINK_MAX_GROUPS) return alink_fail(ctx, ALINK_E_INVALID, "bad group %d", group);
  if (K < 1 || K > ALINK_MAX_K) return alink_fail(ctx, ALINK_E_INVALID, "K=%d outside [1,%d]", K, ALINK_MAX_K);
  if (!((C == 2 && out_act == ALINK_ACT_SOFTMAX) || (C == 1 && out_act == ALINK_ACT_SIGMOID)))
    return alink_fail(ctx, ALINK_E_INVALID, "head must be softmax/C=2 (siamese.py) or sigmoid/C=1 (siamese3.py)");
  if (feat_mode != ALINK_FEAT_ABSDIFF && feat_mode != ALINK_FEAT_ABSDIFF_HADAMARD)
    return alink_fail(ctx, ALINK_E_INVALID, "bad feat_mode %d", feat_mode);
  if (compute != ALINK_COMPUTE_FP32_EXACT && compute != ALINK_COMPUTE_BF16_TC && compute != ALINK_COMPUTE_FP32_REFINE)
    return alink_fail(ctx, ALINK_E_INVALID, "bad compute mode %d", compute);
  if (D_in < 8 || D_in % 8 || H1 < 8 || H1 % 8 || (feat_mode && (D_in % 16)))
    return alink_fail(ctx, ALINK_E_UNSUPPORTED, "D_in and H1 must be multiples of 8 (got %d, %d)", D_in, H1);
  if (H2 != 32 && H2 != 64) return alink_fail(ctx, ALINK_E_UNSUPPORTED, "H2 must be 32 or 64 (got %d)", H2);
  if (!W1 || !b1 || !W2 || !b2 || !W3 || !b3) return alink_fail(ctx, ALINK_E_INVALID, "null weight array");
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  HeadGroup &g = ctx->groups[group];
  free_group(g);
  g.K = K; g.D_in = D_in; g.H1 = H1; g.H2 = H2; g.C = C; g.act = out_act; g.feat_mode = feat_mode;
  g.compute = compute;
  ctx->refine_eps_hint[group] = 0.0;
  ALINK_TRY(upload_stack(ctx, W1, K, (size_t)D_in * H1, &g.W1));
  ALINK_TRY(upload_stack(ctx, b1, K, (size_t)H1, &g.b1));
  ALINK_TRY(upload_stack(ctx, W2, K, (size_t)H1 * H2, &g.W2));
  ALINK_TRY(upload_stack(ctx, b2, K, (size_t)H2, &g.b2));
  ALINK_TRY(upload_stack(ctx, W3, K, (size_t)H2 * C, &g.W3));
  ALINK_TRY(upload_stack(ctx, b3, K, (size_t)C, &g.b3));
  // pointer tables: 6 arrays of K device pointers in one allocation
  const float *host_ptrs[6 * ALINK_MAX_K];
  const float *bases[6] = {g.W1, g.b1, g.W2, g.b2, g.W3, g.b3};
  const size_t each[6] = {(size_t)D_in * H1, (size_t)H1, (size_t)H1 * H2, (size_t)H2, (size_t)H2 * C, (size_t)C};
  for (int a = 0; a < 6; ++a)
    for (int k = 0; k < K; ++k) host_ptrs[a * ALINK_MAX_K + k] = bases[a] + each[a] * k;
  const float **dptr = nullptr;
  ALINK_CUDA(ctx, cudaMalloc((void **)&dptr, sizeof(host_ptrs)));
  ALINK_CUDA(ctx, cudaMemcpyAsync((void *)dptr, host_ptrs, sizeof(host_ptrs), cudaMemcpyHostToDevice, ctx->stream));
  g.pW1 = dptr;
  g.pb1 = dptr + 1 * ALINK_MAX_K;
  g.pW2 = dptr + 2 * ALINK_MAX_K;
  g.pb2 = dptr + 3 * ALINK_MAX_K;
  g.pW3 = dptr + 4 * ALINK_MAX_K;
  g.pb3 = dptr + 5 * ALINK_MAX_K;
  if (compute == ALINK_COMPUTE_BF16_TC) {
    ALINK_TRY(tc_prepare_heads(ctx, g));
    g.tc_ok = true;
  } else if (compute == ALINK_COMPUTE_FP32_REFINE) {
    // bf16 and half operand images (the query picks half when it cannot overflow, refine.cuh);
    // shapes the tensor-core head cannot take run fully exact
    g.tc_ok = tc_prepare_heads(ctx, g, 3) == ALINK_OK;
    if (!g.tc_ok) ctx->err[0] = 0;
    // weight statistics: the overflow guard of the half-precision filter, the power-of-two scales of the split
    // stage and the a-priori error model of the filter (refine_apriori_eps)
    std::vector<double> col((size_t)H1), col2((size_t)H1), v((size_t)H1), t((size_t)D_in);
    for (int k = 0; k < K; ++k) {
      std::fill(col.begin(), col.end(), 0.0);
      std::fill(col2.begin(), col2.end(), 0.0);
      for (int i = 0; i < D_in; ++i) {
        const float *row = W1[k] + (size_t)i * H1;
        for (int j = 0; j < H1; ++j) {
          const double w = fabs((double)row[j]);
          col[j] += w;
          col2[j] += w * w;
          if (!(w <= g.w1_absmax)) g.w1_absmax = w;
        }
      }
      double b1ss = 0.0;
      for (int j = 0; j < H1; ++j) {
        if (!(col[j] <= g.w1_col_l1)) g.w1_col_l1 = col[j];  // NaN-propagating max
        const double c2 = sqrt(col2[j]);
        if (!(c2 <= g.w1_col_l2)) g.w1_col_l2 = c2;
        const double b = fabs((double)b1[k][j]);
        if (!(b <= g.b1_absmax)) g.b1_absmax = b;
        b1ss += b * b;
      }
      if (!(sqrt(b1ss) <= g.b1_l2)) g.b1_l2 = sqrt(b1ss);
      std::vector<double> c22((size_t)H2, 0.0);
      for (int i = 0; i < H1; ++i)
        for (int j = 0; j < H2; ++j) {
          const double w = fabs((double)W2[k][(size_t)i * H2 + j]);
          if (!(w <= g.w_absmax)) g.w_absmax = w;
          c22[j] += w * w;
        }
      for (int j = 0; j < H2; ++j)
        if (!(sqrt(c22[j]) <= g.w2_col_l2)) g.w2_col_l2 = sqrt(c22[j]);
      double w3ss = 0.0;  // the logit the probability depends on: z1 - z0 (softmax-2) or z (sigmoid-1)
      for (int j = 0; j < H2; ++j) {
        const double d = C == 2 ? (double)W3[k][(size_t)j * 2 + 1] - (double)W3[k][(size_t)j * 2] : (double)W3[k][j];
        w3ss += d * d;
      }
      if (!(sqrt(w3ss) <= g.w3_gain)) g.w3_gain = sqrt(w3ss);
      g.w3_gain_ms += w3ss / (double)K;
      if (!(b1ss / (double)H1 <= g.b1_ms)) g.b1_ms = b1ss / (double)H1;
      // spectral norm of W1 (D_in x H1) by power iteration on W1^T W1 (it approaches from below: + 10 %)
      for (int j = 0; j < H1; ++j) v[j] = 1.0 / sqrt((double)H1);
      double sigma = 0.0;
      for (int it = 0; it < 12; ++it) {
        for (int i = 0; i < D_in; ++i) {
          const float *row = W1[k] + (size_t)i * H1;
          double acc = 0.0;
          for (int j = 0; j < H1; ++j) acc += (double)row[j] * v[j];
          t[i] = acc;
        }
        std::fill(v.begin(), v.end(), 0.0);
        for (int i = 0; i < D_in; ++i) {
          const float *row = W1[k] + (size_t)i * H1;
          const double ti = t[i];
          for (int j = 0; j < H1; ++j) v[j] += (double)row[j] * ti;
        }
        double nv = 0.0;
        for (int j = 0; j < H1; ++j) nv += v[j] * v[j];
        nv = sqrt(nv);
        sigma = sqrt(nv);  // ||W^T W v|| -> sigma^2 for unit v
        if (!(nv > 0.0)) break;
        for (int j = 0; j < H1; ++j) v[j] /= nv;
      }
      sigma *= 1.10;
      if (!(sigma <= g.w1_spec)) g.w1_spec = sigma;
    }
    g.split_ok = false;
    if (g.tc_ok && split_supported(ctx, g) && g.w1_absmax < 1e30 && g.w_absmax < 1e30) {
      g.split_ok = split_prepare_heads(ctx, g) == ALINK_OK;
      if (!g.split_ok) ctx->err[0] = 0;
    }
  }
  ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // host_ptrs / caller buffers may go away
  g.set = true;
  return ALINK_OK;
}

// ------------------------------------------------------------------------------------------ tables
int alink_set_table(alink_ctx *ctx, int slot, const void *emb, int64_t M, int D, int dtype, int space) {
  if (!ctx) return ALINK_E_INVALID;
  if (slot < 0 || slot >= ALINK_MAX_SLOTS) return alink_fail(ctx, ALINK_E_INVALID, "bad slot %d", slot);
  if (!emb || M < 1 || D < 8 || D % 8) return alink_fail(ctx, ALINK_E_INVALID, "bad table (M=%lld, D=%d)", (long long)M, D);
  if (dtype != ALINK_F32 && dtype != ALINK_BF16) return alink_fail(ctx, ALINK_E_INVALID, "bad dtype %d", dtype);
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  Table &t = ctx->tables[slot];
  if (t.pending) {  // an asynchronous upload into this slot is still in flight: let it land first
    ALINK_CUDA(ctx, cudaEventSynchronize(t.ready));
    t.pending = false;
  }
  const size_t bytes = (size_t)M * D * (dtype == ALINK_F32 ? 4 : 2);
  if (space == ALINK_DEVICE) {
    t.p = emb;
  } else {
    ALINK_TRY(dev_reserve(ctx, t.own, bytes));
    ALINK_CUDA(ctx, cudaMemcpyAsync(t.own.p, emb, bytes, cudaMemcpyHostToDevice, ctx->stream));
    ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    t.p = t.own.p;
  }
  t.M = M; t.D = D; t.dtype = dtype;
  t.absmax = -1.0;
  return ALINK_OK;
}

// Upload a HOST table into a slot on the context's copy stream and return at once; the next call that reads the
// slot makes the compute stream wait for the upload. With two slots the upload of the next pool overlaps the query
// on the current one (bench.py e2e). The host buffer must stay valid (and should be pinned) until that next call.
int alink_set_table_async(alink_ctx *ctx, int slot, const void *emb, int64_t M, int D, int dtype) {
  if (!ctx) return ALINK_E_INVALID;
  if (slot < 0 || slot >= ALINK_MAX_SLOTS) return alink_fail(ctx, ALINK_E_INVALID, "bad slot %d", slot);
  if (!emb || M < 1 || D < 8 || D % 8) return alink_fail(ctx, ALINK_E_INVALID, "bad table (M=%lld, D=%d)", (long long)M, D);
  if (dtype != ALINK_F32 && dtype != ALINK_BF16) return alink_fail(ctx, ALINK_E_INVALID, "bad dtype %d", dtype);
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  if (!ctx->copy_stream) {
    ALINK_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    ALINK_CUDA(ctx, cudaEventCreateWithFlags(&ctx->copy_gate, cudaEventDisableTiming));
    ALINK_CUDA(ctx, cudaEventCreateWithFlags(&ctx->copy_done, cudaEventDisableTiming));
  }
  Table &t = ctx->tables[slot];
  if (!t.ready) ALINK_CUDA(ctx, cudaEventCreateWithFlags(&t.ready, cudaEventDisableTiming));
  const size_t bytes = (size_t)M * D * (dtype == ALINK_F32 ? 4 : 2);
  ALINK_TRY(dev_reserve(ctx, t.own, bytes));  // (a reallocation synchronises the compute stream)
  // kernels already queued on the compute stream may still read this slot: the copy starts behind them
  ALINK_CUDA(ctx, cudaEventRecord(ctx->copy_gate, ctx->stream));
  ALINK_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->copy_gate, 0));
  ALINK_CUDA(ctx, cudaMemcpyAsync(t.own.p, emb, bytes, cudaMemcpyHostToDevice, ctx->copy_stream));
  ALINK_CUDA(ctx, cudaEventRecord(t.ready, ctx->copy_stream));
  ALINK_CUDA(ctx, cudaEventRecord(ctx->copy_done, ctx->copy_stream));
  ctx->copy_busy = true;
  t.p = t.own.p;
  t.M = M; t.D = D; t.dtype = dtype;
  t.absmax = -1.0;
  t.pending = true;
  return ALINK_OK;
}

// the compute stream waits for an upload in flight into this slot (no host synchronisation)
static int table_wait(alink_ctx *ctx, int slot) {
  if (slot < 0 || slot >= ALINK_MAX_SLOTS) return ALINK_OK;  // the caller reports the bad slot
  Table &t = ctx->tables[slot];
  if (t.pending) {
    ALINK_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, t.ready, 0));
    t.pending = false;
  }
  return ALINK_OK;
}

// ------------------------------------------------------------------------------------------ scoring
// ia/ib are device pointers (or null = identity). Fills ctx->probs [N][K][C].
// fuse_measure >= 0: a compute mode that can reduce the K learners in its epilogue also fills ctx->util.
static int score_impl(alink_ctx *ctx, int group, const Table &A, const Table &Bt, const int32_t *ia,
                      const int32_t *ib, int64_t N, int fuse_measure = -1) {
  if (group < 0 || group >= ALINK_MAX_GROUPS || !ctx->groups[group].set)
    return alink_fail(ctx, ALINK_E_STATE, "head group %d not set (alink_set_heads)", group);
  const HeadGroup &g = ctx->groups[group];
  if (!A.p || !Bt.p) return alink_fail(ctx, ALINK_E_STATE, "embedding table slot not set (alink_set_table)");
  if (A.D != Bt.D || A.dtype != Bt.dtype) return alink_fail(ctx, ALINK_E_INVALID, "table shape/dtype mismatch");
  const int want = g.feat_mode ? 2 * A.D : A.D;
  if (want != g.D_in)
    return alink_fail(ctx, ALINK_E_INVALID, "table D=%d gives fan-in %d but heads expect %d", A.D, want, g.D_in);
  if (!ia && !ctx->ap_G && (N > A.M || N > Bt.M)) return alink_fail(ctx, ALINK_E_INVALID, "identity pairs exceed table rows");
  if (N < 0 || N > 0xFFFFFFFEll) return alink_fail(ctx, ALINK_E_INVALID, "N out of range");
  ctx->N = N; ctx->sK = g.K; ctx->sC = g.C; ctx->util_valid = false; ctx->probs_full = true;
  // an all-pairs query whose K-learner reduction is fused needs no [N][K][C] probabilities (100 GB at 50k x 50k)
  const bool skip_probs = ctx->ap_G && fuse_measure >= 0 && g.compute == ALINK_COMPUTE_BF16_TC;
  if (!skip_probs) ALINK_TRY(dev_reserve(ctx, ctx->probs, sizeof(float) * (size_t)(N ? N : 1) * g.K * g.C));
  if (N == 0) return ALINK_OK;
  if (g.compute == ALINK_COMPUTE_BF16_TC) {
    float *util = nullptr;
    ctx->probs_full = !skip_probs;
    if (fuse_measure >= 0) {
      ALINK_TRY(dev_reserve(ctx, ctx->util, sizeof(float) * (size_t)N));
      util = (float *)ctx->util.p;
    }
    ALINK_TRY(tc_score(ctx, g, A, Bt, ia, ib, N, skip_probs ? nullptr : (float *)ctx->probs.p, util,
                       fuse_measure >= 0 ? fuse_measure : 0));
    if (util) {
      ctx->util_valid = true;
      ctx->util_measure = fuse_measure;
    }
    return ALINK_OK;
  }
  if (A.dtype != ALINK_F32)
    return alink_fail(ctx, ALINK_E_UNSUPPORTED, "FP32_EXACT mode needs fp32 tables");
  const int64_t chunk = 65536;
  const int Mc_max = (int)(N < chunk ? N : chunk);
  const int64_t lstride = (int64_t)Mc_max * g.H1;
  ALINK_TRY(dev_reserve(ctx, ctx->h1ws, sizeof(float) * (size_t)lstride * g.K));
  for (int64_t row0 = 0; row0 < N; row0 += chunk) {
    const int Mc = (int)((N - row0) < chunk ? (N - row0) : chunk);
    ALINK_TRY(ffma_head_chunk(ctx, g, (const float *)A.p, (const float *)Bt.p, A.D, ia, ib, row0, Mc,
                              (float *)ctx->h1ws.p, lstride, (float *)ctx->probs.p));
  }
  return ALINK_OK;
}

static int emit_scores(alink_ctx *ctx, float *per_learner, float *consensus, int out_space) {
  const int64_t N = ctx->N;
  if (per_learner)
    ALINK_TRY(copy_out(ctx, per_learner, ctx->probs.p, sizeof(float) * (size_t)N * ctx->sK * ctx->sC, out_space));
  if (consensus) {
    float *dst = consensus;
    if (out_space == ALINK_HOST) {
      ALINK_TRY(dev_reserve(ctx, ctx->cons, sizeof(float) * (size_t)(N ? N : 1) * ctx->sC));
      dst = (float *)ctx->cons.p;
    }
    ALINK_TRY(launch_utility(ctx, (const float *)ctx->probs.p, N, ctx->sK, ctx->sC, 0, nullptr, dst));
    if (out_space == ALINK_HOST)
      ALINK_TRY(copy_out(ctx, consensus, dst, sizeof(float) * (size_t)N * ctx->sC, ALINK_HOST));
  }
  if (out_space == ALINK_HOST && (per_learner || consensus)) ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return ALINK_OK;
}

// ALINK_FLAG_CHECK_INDICES: largest index as an unsigned number (a negative one becomes huge) per side
__global__ void __launch_bounds__(256) index_range_kernel(const int32_t *__restrict__ ia, const int32_t *__restrict__ ib,
                                                          int64_t N, unsigned int *__restrict__ out) {
  unsigned int ma = 0, mb = 0;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < N; i += (int64_t)gridDim.x * 256) {
    ma = max(ma, (unsigned int)__ldg(ia + i));
    mb = max(mb, (unsigned int)__ldg(ib + i));
  }
  for (int o = 16; o; o >>= 1) {
    ma = max(ma, __shfl_xor_sync(0xffffffffu, ma, o));
    mb = max(mb, __shfl_xor_sync(0xffffffffu, mb, o));
  }
  if ((threadIdx.x & 31) == 0) {
    if (ma) atomicMax(out, ma);
    if (mb) atomicMax(out + 1, mb);
  }
}

static int check_indices(alink_ctx *ctx, const int32_t *dia, const int32_t *dib, int64_t N, int slot_a, int slot_b) {
  if (!(ctx->flags & ALINK_FLAG_CHECK_INDICES) || !dia || N <= 0) return ALINK_OK;
  ALINK_TRY(dev_reserve(ctx, ctx->rstat, 256));
  unsigned int *d = (unsigned int *)ctx->rstat.p + 6;
  ALINK_CUDA(ctx, cudaMemsetAsync(d, 0, 8, ctx->stream));
  int64_t blocks = (N + 2047) / 2048;
  if (blocks > (int64_t)ctx->sm_count * 8) blocks = (int64_t)ctx->sm_count * 8;
  index_range_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(dia, dib, N, d);
  ALINK_LAUNCHED(ctx);
  unsigned int h[2] = {0, 0};
  ALINK_CUDA(ctx, cudaMemcpyAsync(h, d, 8, cudaMemcpyDeviceToHost, ctx->stream));
  ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if ((int64_t)h[0] >= ctx->tables[slot_a].M)
    return alink_fail(ctx, ALINK_E_INVALID, "ia holds index %d, table slot %d has %lld rows", (int)h[0], slot_a,
                      (long long)ctx->tables[slot_a].M);
  if ((int64_t)h[1] >= ctx->tables[slot_b].M)
    return alink_fail(ctx, ALINK_E_INVALID, "ib holds index %d, table slot %d has %lld rows", (int)h[1], slot_b,
                      (long long)ctx->tables[slot_b].M);
  return ALINK_OK;
}

int alink_score_indexed(alink_ctx *ctx, int group, int slot_a, int slot_b, const int32_t *ia, const int32_t *ib,
                        int64_t N, int in_space, float *per_learner, float *consensus, int out_space) {
  if (!ctx) return ALINK_E_INVALID;
  if (slot_a < 0 || slot_a >= ALINK_MAX_SLOTS || slot_b < 0 || slot_b >= ALINK_MAX_SLOTS)
    return alink_fail(ctx, ALINK_E_INVALID, "bad slot");
  if ((ia == nullptr) != (ib == nullptr)) return alink_fail(ctx, ALINK_E_INVALID, "ia and ib must both be given or both NULL");
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  const void *dia, *dib;
  ALINK_TRY(stage_in(ctx, ia, sizeof(int32_t) * (size_t)N, in_space, ctx->idx_a, &dia));
  ALINK_TRY(stage_in(ctx, ib, sizeof(int32_t) * (size_t)N, in_space, ctx->idx_b, &dib));
  ALINK_TRY(check_indices(ctx, (const int32_t *)dia, (const int32_t *)dib, N, slot_a, slot_b));
  ALINK_TRY(table_wait(ctx, slot_a));
  ALINK_TRY(table_wait(ctx, slot_b));
  ALINK_TRY(score_impl(ctx, group, ctx->tables[slot_a], ctx->tables[slot_b], (const int32_t *)dia,
                       (const int32_t *)dib, N));
  return emit_scores(ctx, per_learner, consensus, out_space);
}

int alink_score_dense(alink_ctx *ctx, int group, const float *XL, const float *XR, int64_t N, int D, int in_space,
                      float *per_learner, float *consensus, int out_space) {
  if (!ctx) return ALINK_E_INVALID;
  if (N < 0 || D < 8 || D % 8) return alink_fail(ctx, ALINK_E_INVALID, "bad dense shape (N=%lld, D=%d)", (long long)N, D);
  if (N == 0) {
    ctx->N = 0;
    return ALINK_OK;
  }
  if (!XL || !XR) return alink_fail(ctx, ALINK_E_INVALID, "null XL/XR");
  ALINK_TRY(alink_set_table(ctx, ALINK_MAX_SLOTS - 2, XL, N, D, ALINK_F32, in_space));
  ALINK_TRY(alink_set_table(ctx, ALINK_MAX_SLOTS - 1, XR, N, D, ALINK_F32, in_space));
  ALINK_TRY(score_impl(ctx, group, ctx->tables[ALINK_MAX_SLOTS - 2], ctx->tables[ALINK_MAX_SLOTS - 1], nullptr,
                       nullptr, N));
  return emit_scores(ctx, per_learner, consensus, out_space);
}

// ------------------------------------------------------------------------------------------ utility
static int utility_impl(alink_ctx *ctx, int measure) {
  if (measure < 0 || measure >= ALINK_UTIL_COUNT) return alink_fail(ctx, ALINK_E_INVALID, "bad measure %d", measure);
  if (!ctx->probs.p || !ctx->probs_full)
    return alink_fail(ctx, ALINK_E_STATE, "no scores: call alink_score_* first");
  ALINK_TRY(dev_reserve(ctx, ctx->util, sizeof(float) * (size_t)(ctx->N ? ctx->N : 1)));
  ALINK_TRY(launch_utility(ctx, (const float *)ctx->probs.p, ctx->N, ctx->sK, ctx->sC, measure,
                           (float *)ctx->util.p, nullptr));
  ctx->util_valid = true;
  ctx->util_measure = measure;
  return ALINK_OK;
}

int alink_utility(alink_ctx *ctx, int measure, float *util, int out_space) {
  if (!ctx) return ALINK_E_INVALID;
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  ALINK_TRY(utility_impl(ctx, measure));
  ALINK_TRY(copy_out(ctx, util, ctx->util.p, sizeof(float) * (size_t)ctx->N, out_space));
  if (util && out_space == ALINK_HOST) ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return ALINK_OK;
}

int alink_utility_from_proba(alink_ctx *ctx, const float *per_learner, int64_t N, int K, int C, int measure,
                             int in_space, float *util, float *consensus, int out_space) {
  if (!ctx) return ALINK_E_INVALID;
  if (N < 0 || K < 1 || K > ALINK_MAX_K || (C != 1 && C != 2) || measure < 0 || measure >= ALINK_UTIL_COUNT)
    return alink_fail(ctx, ALINK_E_INVALID, "bad arguments (N=%lld K=%d C=%d measure=%d)", (long long)N, K, C, measure);
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t nbytes = sizeof(float) * (size_t)N * K * C;
  if (N > 0) {
    if (!per_learner) return alink_fail(ctx, ALINK_E_INVALID, "null per_learner");
    ALINK_TRY(dev_reserve(ctx, ctx->probs, nbytes));
    ALINK_CUDA(ctx, cudaMemcpyAsync(ctx->probs.p, per_learner, nbytes,
                                    in_space == ALINK_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice,
                                    ctx->stream));
  }
  ctx->N = N; ctx->sK = K; ctx->sC = C;
  ALINK_TRY(utility_impl(ctx, measure));
  ALINK_TRY(copy_out(ctx, util, ctx->util.p, sizeof(float) * (size_t)N, out_space));
  if (consensus) ALINK_TRY(emit_scores(ctx, nullptr, consensus, out_space));
  if (out_space == ALINK_HOST) ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return ALINK_OK;
}

// ------------------------------------------------------------------------------------------ top-B
static int emit_selection(alink_ctx *ctx, int64_t B, int64_t *idx_out, float *util_out, uint64_t *keys_out,
                          int out_space) {
  if (B == 0) return ALINK_OK;
  const uint64_t *keys = (const uint64_t *)ctx->sel_keys.p;
  int64_t *didx = idx_out;
  float *dutil = util_out;
  if (out_space == ALINK_HOST) {
    ALINK_TRY(dev_reserve(ctx, ctx->sel_idx, sizeof(int64_t) * (size_t)B));
    ALINK_TRY(dev_reserve(ctx, ctx->sel_util, sizeof(float) * (size_t)B));
    didx = idx_out ? (int64_t *)ctx->sel_idx.p : nullptr;
    dutil = util_out ? (float *)ctx->sel_util.p : nullptr;
  }
  if (didx || dutil) {
    int64_t blocks = (B + 255) / 256;
    if (blocks > 1024) blocks = 1024;
    unpack_keys_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(keys, B, didx, dutil);
    ALINK_LAUNCHED(ctx);
  }
  if (out_space == ALINK_HOST) {
    ALINK_TRY(copy_out(ctx, idx_out, didx, sizeof(int64_t) * (size_t)B, ALINK_HOST));
    ALINK_TRY(copy_out(ctx, util_out, dutil, sizeof(float) * (size_t)B, ALINK_HOST));
  }
  ALINK_TRY(copy_out(ctx, keys_out, keys, sizeof(uint64_t) * (size_t)B, out_space));
  if (out_space == ALINK_HOST) ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return ALINK_OK;
}

int alink_select_topk(alink_ctx *ctx, int64_t B, int64_t base, int64_t *idx_out, float *util_out,
                      uint64_t *keys_out, int out_space) {
  if (!ctx) return ALINK_E_INVALID;
  if (!ctx->util_valid) return alink_fail(ctx, ALINK_E_STATE, "no utilities: call alink_utility first");
  if (B < 0 || B > ctx->N) return alink_fail(ctx, ALINK_E_INVALID, "B=%lld outside [0, N=%lld]", (long long)B, (long long)ctx->N);
  if (base < 0 || base + ctx->N > 0xFFFFFFFFll) return alink_fail(ctx, ALINK_E_INVALID, "base + N exceeds 2^32 - 1");
  if (B == 0) return ALINK_OK;
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  KeySrc src{(const float *)ctx->util.p, nullptr, (uint64_t)base};
  ALINK_TRY(topk_device(ctx, src, ctx->N, B));
  return emit_selection(ctx, B, idx_out, util_out, keys_out, out_space);
}

int alink_topk_values(alink_ctx *ctx, const float *util, int64_t N, int64_t base, int64_t B, int in_space,
                      int64_t *idx_out, float *util_out, uint64_t *keys_out, int out_space) {
  if (!ctx) return ALINK_E_INVALID;
  if (N < 0 || B < 0 || B > N) return alink_fail(ctx, ALINK_E_INVALID, "B=%lld outside [0, N=%lld]", (long long)B, (long long)N);
  if (base < 0 || base + N > 0xFFFFFFFFll) return alink_fail(ctx, ALINK_E_INVALID, "base + N exceeds 2^32 - 1");
  if (B == 0) return ALINK_OK;
  if (!util) return alink_fail(ctx, ALINK_E_INVALID, "null util");
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  const void *du;
  ALINK_TRY(stage_in(ctx, util, sizeof(float) * (size_t)N, in_space, ctx->sel_in, &du));
  KeySrc src{(const float *)du, nullptr, (uint64_t)base};
  ALINK_TRY(topk_device(ctx, src, N, B));
  return emit_selection(ctx, B, idx_out, util_out, keys_out, out_space);
}

int alink_topk_keys(alink_ctx *ctx, const uint64_t *keys, int64_t n, int64_t B, int in_space, uint64_t *keys_out,
                    int out_space) {
  if (!ctx) return ALINK_E_INVALID;
  if (n < 0 || B < 0 || B > n) return alink_fail(ctx, ALINK_E_INVALID, "B=%lld outside [0, n=%lld]", (long long)B, (long long)n);
  if (B == 0) return ALINK_OK;
  if (!keys) return alink_fail(ctx, ALINK_E_INVALID, "null keys");
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  const void *dk;
  ALINK_TRY(stage_in(ctx, keys, sizeof(uint64_t) * (size_t)n, in_space, ctx->sel_in, &dk));
  KeySrc src{nullptr, (const uint64_t *)dk, 0};
  ALINK_TRY(topk_device(ctx, src, n, B));
  return emit_selection(ctx, B, nullptr, nullptr, keys_out, out_space);
}

// ------------------------------------------------------------------------------- no-allocation top-B
// alink_topk_values_noalloc: the multi-launch top-B (6 radix passes -> compaction -> one-CTA bitonic sort) with its
// scratch — the SelectState (threshold, ticket, 2048-bin histogram) and the B unsorted keys — in a file-scope device
// array instead of the context's growable buffers, for callers that may not allocate (graph capture, serving loops).
// 48 KB per GPU that loads the library, held until exit. B <= ALINK_NOALLOC_MAX_B. The array is shared by every call
// and keeps the last call's contents: select_init_kernel clears the state each time; calls that overlap on the GPU race.
#define ALINK_NOALLOC_STATE_BYTES 16384
#define ALINK_NOALLOC_SCRATCH_BYTES (ALINK_NOALLOC_STATE_BYTES + 8 * ALINK_NOALLOC_MAX_B)
static_assert(sizeof(SelectState) <= ALINK_NOALLOC_STATE_BYTES, "SelectState outgrew its scratch share");
static __device__ __align__(256) unsigned char g_noalloc_scratch[ALINK_NOALLOC_SCRATCH_BYTES];
static void *g_noalloc_addr[64];  // per device: the array's DEVICE address (its name is only a host-side handle here)
static std::mutex g_noalloc_mu;

// Called from alink_create — set-up time, once per device, with that device current — so that the entry point itself makes
// no CUDA call that could allocate or initialise anything (it may run under stream capture).
static int noalloc_resolve(int device) {
  if (device < 0 || device >= 64) return ALINK_E_UNSUPPORTED;
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (g_noalloc_addr[device]) return ALINK_OK;
  void *p = nullptr;
  const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
  if (e != cudaSuccess || !p) {
    cudaGetLastError();
    return ALINK_E_CUDA;
  }
  g_noalloc_addr[device] = p;
  return ALINK_OK;
}

static int noalloc_scratch(alink_ctx *ctx, unsigned char **out) {
  if (ctx->device < 0 || ctx->device >= 64) return alink_fail(ctx, ALINK_E_UNSUPPORTED, "device ordinal %d >= 64", ctx->device);
  void *p;
  {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    p = g_noalloc_addr[ctx->device];
  }
  if (!p) return alink_fail(ctx, ALINK_E_CUDA, "the static scratch's device address could not be resolved at alink_create (cudaGetSymbolAddress failed)");
  *out = (unsigned char *)p;
  return ALINK_OK;
}

int alink_topk_values_noalloc(alink_ctx *ctx, const float *util_dev, int64_t N, int64_t base, int64_t B, int64_t *idx_out_dev,
                              float *util_out_dev, uint64_t *keys_out_dev) {
  if (!ctx) return ALINK_E_INVALID;
  if (B > ALINK_NOALLOC_MAX_B)
    return alink_fail(ctx, ALINK_E_UNSUPPORTED, "alink_topk_values_noalloc: B=%lld exceeds the static scratch (max %d); use alink_topk_values",
                      (long long)B, ALINK_NOALLOC_MAX_B);
  if (N < 0 || B < 0 || B > N) return alink_fail(ctx, ALINK_E_INVALID, "B=%lld outside [0, N=%lld]", (long long)B, (long long)N);
  if (base < 0 || base + N > 0xFFFFFFFFll) return alink_fail(ctx, ALINK_E_INVALID, "base + N exceeds 2^32 - 1");
  if (B == 0) return ALINK_OK;
  if (!util_dev) return alink_fail(ctx, ALINK_E_INVALID, "null util");
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  unsigned char *scr = nullptr;
  ALINK_TRY(noalloc_scratch(ctx, &scr));
  SelectState *st = (SelectState *)scr;
  uint64_t *keys = (uint64_t *)(scr + ALINK_NOALLOC_STATE_BYTES);
  KeySrc src{util_dev, nullptr, (uint64_t)base, nullptr};
  select_init_kernel<<<1, 256, 0, ctx->stream>>>(st, (unsigned long long)B);
  ALINK_LAUNCHED(ctx);
  int64_t blocks = (N + 256 * 8 - 1) / (256 * 8);
  const int64_t cap = (int64_t)ctx->sm_count * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  static const int shifts[6] = {53, 42, 32, 21, 10, 0};
  static const int nbits[6] = {11, 11, 10, 11, 11, 10};
  for (int p = 0; p < 6; ++p) {
    radix_pass_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(src, N, shifts[p], nbits[p], p == 0, st);
    ALINK_LAUNCHED(ctx);
  }
  compact_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(src, N, st, keys, (unsigned int)B);
  ALINK_LAUNCHED(ctx);
  const int64_t npad = next_pow2(B) < 2 ? 2 : next_pow2(B);
  const int threads = npad >= 2048 ? 1024 : (npad >= 64 ? (int)(npad / 2) : 32);
  bitonic_smem_kernel<<<1, threads, sizeof(uint64_t) * (size_t)npad, ctx->stream>>>(keys, (int)B, (int)npad);
  ALINK_LAUNCHED(ctx);
  if (idx_out_dev || util_out_dev) {
    unpack_keys_kernel<<<(unsigned)((B + 255) / 256), 256, 0, ctx->stream>>>(keys, B, idx_out_dev, util_out_dev);
    ALINK_LAUNCHED(ctx);
  }
  if (keys_out_dev)
    ALINK_CUDA(ctx, cudaMemcpyAsync(keys_out_dev, keys, sizeof(uint64_t) * (size_t)B, cudaMemcpyDeviceToDevice, ctx->stream));
  ALINK_CUDA(ctx, cudaGetLastError());
  return ALINK_OK;
}

// ------------------------------------------------------------------------------------------ fused step
// FP32_REFINE (refine.cuh): bf16 tensor-core filter over all N pairs, bit-exact re-score of the candidates.
// On return ctx->sel_keys holds the B selected keys (device, sorted). ev (6 events) may be null.
// max |element| of an fp32 table, computed once per alink_set_table (one streaming pass)
static int table_absmax(alink_ctx *ctx, Table &t, double *out, double *rownorm = nullptr) {
  if (t.absmax < 0.0) {
    ALINK_TRY(dev_reserve(ctx, ctx->rstat, 256));
    unsigned int *bits = (unsigned int *)ctx->rstat.p + 32;
    ALINK_CUDA(ctx, cudaMemsetAsync(bits, 0, 16, ctx->stream));
    table_stats_kernel<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>((const float *)t.p, t.M, t.D, bits);
    ALINK_LAUNCHED(ctx);
    unsigned int h[4] = {0, 0, 0, 0};
    ALINK_CUDA(ctx, cudaMemcpyAsync(h, bits, 16, cudaMemcpyDeviceToHost, ctx->stream));
    ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float f, n2;
    double tot;
    memcpy(&f, &h[0], 4);
    memcpy(&n2, &h[1], 4);
    memcpy(&tot, &h[2], 8);
    t.absmax = (double)f;  // +inf stays +inf (fails the bound); NaN inputs never reach a finite utility anyway
    t.rownorm = sqrt((double)n2);
    t.rowms = tot / (double)t.M;
  }
  *out = t.absmax;
  if (rownorm) *rownorm = t.rownorm;
  return ALINK_OK;
}

// per-measure factor between an error of a PROBABILITY and the error of the utility near the top
// (margin = p(1) - p(2) doubles it; KL-max compares a learner with the mean: measured 4.3x)
static const double REFINE_SCALE[ALINK_UTIL_COUNT] = {1.0, 2.0, 1.25, 1.0, 1.25, 5.0};
static const double REFINE_EPS_BF16 = 1.2e-2;  // measured max |dp| 9.0e-3 (profiles/r1_refine_filter_error.txt)
static const double REFINE_EPS_F16 = 1.5e-3;   // measured max |dp| 8.4e-4, rms 1.2e-4 (same file)

static int refine_query(alink_ctx *ctx, const HeadGroup &g, Table &A, Table &Bt, const int32_t *ia,
                        const int32_t *ib, int64_t N, int measure, int64_t B, int64_t base, cudaEvent_t *ev) {
  ctx->N = N; ctx->sK = g.K; ctx->sC = g.C; ctx->util_valid = false; ctx->probs_full = false;
  // operand format of the filter: half unless a feature or hidden activation could exceed its range
  static const int fmt_env = [] { const char *e = getenv("ALINK_REFINE_FMT"); return e ? atoi(e) : 0; }();  // 1 = force bf16
  bool f16 = g.W1h && g.W2h && fmt_env != 1 && !g.bounds_stale;
  if (f16) {
    double ma, mb;
    ALINK_TRY(table_absmax(ctx, A, &ma));
    ALINK_TRY(table_absmax(ctx, Bt, &mb));
    double xmax = ma + mb;
    if (g.feat_mode == ALINK_FEAT_ABSDIFF_HADAMARD && ma * mb > xmax) xmax = ma * mb;
    const double h1max = xmax * g.w1_col_l1 + g.b1_absmax;
    f16 = xmax < 6.0e4 && h1max < 6.0e4 && g.w_absmax < 6.0e4;  // false for NaN / inf
  }
  ALINK_TRY(dev_reserve(ctx, ctx->util, sizeof(float) * (size_t)N));
  ALINK_TRY(dev_reserve(ctx, ctx->sel_state, 2 * sizeof(SelectState)));
  ALINK_TRY(dev_reserve(ctx, ctx->cand, sizeof(uint32_t) * (size_t)N));
  ALINK_TRY(dev_reserve(ctx, ctx->rstat, 256));
  float *util_a = (float *)ctx->util.p;
  // second state slot: tau must survive the exact top-B of step (5), which runs in slot 0, for a widening round
  SelectState *st = (SelectState *)ctx->sel_state.p + 1;
  unsigned int *rstat = (unsigned int *)ctx->rstat.p;  // [0] candidate count, [1] max error bits, [2..3] sum sq
  // (1) approximate utilities of every pair, fused K-learner reduction in the kernel's epilogue
  ALINK_TRY(tc_score(ctx, g, A, Bt, ia, ib, N, nullptr, util_a, measure, f16));
  if (ev) ALINK_CUDA(ctx, cudaEventRecord(ev[2], ctx->stream));
  // (2) B-th largest approximate utility
  KeySrc asrc{util_a, nullptr, 0, nullptr};
  ALINK_TRY(select_threshold(ctx, asrc, N, B, st, 3));  // the utility bits only: tau, not a tie-broken key
  int64_t blocks = (N + 2047) / 2048;
  const int64_t cap = (int64_t)ctx->sm_count * 8;
  if (blocks > cap) blocks = cap;
  // eps bounds the error of a PROBABILITY; a measure moves by at most `scale` times that near the top
  // (margin = p(1) - p(2) doubles it; KL-max compares a learner with the mean: measured 4.3x)
  const double *scale = REFINE_SCALE;
  double eps = ctx->refine_eps > 0.0 ? ctx->refine_eps : (f16 ? REFINE_EPS_F16 : REFINE_EPS_BF16);
  ctx->refine_f16 = f16 ? 1 : 0;
  // marks the candidates (ctx->cand, count -> *M_out) and scores them exactly (ctx->probs, ctx->util_c)
  auto mark_and_rescore = [&](const unsigned int *pilot_min, float margin, int64_t *M_out, int64_t *cb_out) -> int {
    ALINK_CUDA(ctx, cudaMemsetAsync(rstat, 0, 16, ctx->stream));
    refine_mark_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(util_a, N, st, pilot_min, margin, (uint32_t *)ctx->cand.p,
                                                                  rstat, 0xFFFFFFFFu);
    ALINK_LAUNCHED(ctx);
    unsigned int hM = 0;
    ALINK_CUDA(ctx, cudaMemcpyAsync(&hM, rstat, 4, cudaMemcpyDeviceToHost, ctx->stream));
    ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const int64_t M = (int64_t)hM;
    *M_out = M;
    if (M < B) return alink_fail(ctx, ALINK_E_STATE, "refine: %lld candidates < B=%lld", (long long)M, (long long)B);
    // exact re-score of the candidates (oracle summation order, CUDA cores)
    ALINK_TRY(dev_reserve(ctx, ctx->cia, sizeof(int32_t) * (size_t)M));
    ALINK_TRY(dev_reserve(ctx, ctx->cib, sizeof(int32_t) * (size_t)M));
    ALINK_TRY(dev_reserve(ctx, ctx->util_c, sizeof(float) * (size_t)M));
    ALINK_TRY(dev_reserve(ctx, ctx->probs, sizeof(float) * (size_t)M * g.K * g.C));
    int64_t cb = (M + 255) / 256;
    if (cb > cap) cb = cap;
    *cb_out = cb;
    refine_compose_kernel<<<(unsigned)cb, 256, 0, ctx->stream>>>((const uint32_t *)ctx->cand.p, (unsigned)M, ia, ib,
                                                                 ia ? 0u : ctx->ap_G, ctx->ap_pair0,
                                                                 (int32_t *)ctx->cia.p, (int32_t *)ctx->cib.p);
    ALINK_LAUNCHED(ctx);
    const int64_t chunk = 65536;
    const int Mc_max = (int)(M < chunk ? M : chunk);
    const int64_t lstride = (int64_t)Mc_max * g.H1;
    ALINK_TRY(dev_reserve(ctx, ctx->h1ws, sizeof(float) * (size_t)lstride * g.K));
    for (int64_t row0 = 0; row0 < M; row0 += chunk) {
      const int Mc = (int)((M - row0) < chunk ? (M - row0) : chunk);
      ALINK_TRY(ffma_head_chunk(ctx, g, (const float *)A.p, (const float *)Bt.p, A.D, (const int32_t *)ctx->cia.p,
                                (const int32_t *)ctx->cib.p, row0, Mc, (float *)ctx->h1ws.p, lstride,
                                (float *)ctx->probs.p));
    }
    return launch_utility(ctx, (const float *)ctx->probs.p, M, g.K, g.C, measure, (float *)ctx->util_c.p, nullptr);
  };
  // (2b) pilot: the pairs with u~ >= tau, scored exactly; the smallest of their exact utilities bounds the exact
  // B-th largest value from below (refine.cuh). ALINK_REFINE_PILOT=0 keeps the plain 2 eps band.
  static const int pilot_env = [] { const char *e = getenv("ALINK_REFINE_PILOT"); return e ? atoi(e) : 1; }();
  unsigned int *pilot_min = nullptr;
  if (pilot_env) {
    int64_t M0 = 0, cb0 = 1;
    ALINK_TRY(mark_and_rescore(nullptr, 0.0f, &M0, &cb0));
    pilot_min = rstat + 4;
    ALINK_CUDA(ctx, cudaMemsetAsync(pilot_min, 0xFF, 4, ctx->stream));
    refine_min_kernel<<<(unsigned)cb0, 256, 0, ctx->stream>>>((const float *)ctx->util_c.p, (unsigned)M0, pilot_min);
    ALINK_LAUNCHED(ctx);
  }
  for (int round = 1;; ++round) {
    // (3) candidates: within eps of the pilot's bound, or within 2 eps of tau; (4) their exact re-score
    int64_t M = 0, cb = 1;
    ALINK_TRY(mark_and_rescore(pilot_min, (float)((pilot_min ? 1.0 : 2.0) * eps * scale[measure]), &M, &cb));
    refine_maxerr_kernel<<<(unsigned)cb, 256, 0, ctx->stream>>>((const float *)ctx->util_c.p, util_a,
                                                                (const uint32_t *)ctx->cand.p, (unsigned)M, rstat);
    ALINK_LAUNCHED(ctx);
    if (ev) ALINK_CUDA(ctx, cudaEventRecord(ev[3], ctx->stream));
    // (5) exact top-B among the candidates, keyed by the pairs' original indices
    KeySrc csrc{(const float *)ctx->util_c.p, nullptr, (uint64_t)base, (const uint32_t *)ctx->cand.p};
    ALINK_TRY(topk_device(ctx, csrc, M, B));
    unsigned int hs[4] = {0, 0, 0, 0};
    ALINK_CUDA(ctx, cudaMemcpyAsync(hs, rstat, 16, cudaMemcpyDeviceToHost, ctx->stream));
    ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    double sumsq;
    memcpy(&sumsq, hs + 2, 8);
    ctx->refine_survivors = M;
    memcpy(&ctx->refine_max_err, hs + 1, 4);
    ctx->refine_rms_err = (float)sqrt(sumsq / (double)M);
    ctx->refine_eps_used = eps;
    ctx->refine_rounds = round;
    // the bound must hold on every candidate we can check; if not, widen it and redo (3)-(5)
    const double seen = (double)ctx->refine_max_err / scale[measure];
    if (seen <= eps || M >= N || round >= 4) break;
    eps = 2.0 * (seen > eps ? seen : eps);
  }
  return ALINK_OK;
}

static const double REFINE_EPS2 = 2.0e-5;  // default stage-2 (split operand) bound; measured: profiles/r2_refine_errors.txt

// A-priori error of the half-precision filter for this group and these tables. Rounding errors of the
// operands are modelled as independent with relative rms delta = u / sqrt 6 (u = the format's unit roundoff; a
// significand uniform in [1, 2)) and propagated through the three layers:
//   layer 1, active unit:  s1^2 = 2 delta^2 |x|^2 c1^2 / D_in                    (c1 = column L2 norm of W1)
//   layer 2, active unit:  s2^2 = c2^2 (s1^2 / 2 + 2 delta^2 |h1|^2 / H1)        (relu passes half of the units)
//   logit z1 - z0:         sd^2 = s2^2 |w3d|^2 / 2 ;  probability: sd / 4 ;  mean of K independent learners: / sqrt K
// refine_model_eps  uses TYPICAL norms (|x|^2 = mean row energy of both tables, |h1|^2 from it) and returns z = 8
//                   standard deviations: checked against an emulation of the filter (1.4x above its measured rms) and
//                   used as the filter width together with the per-format default;
// refine_worst_eps  propagates the same model with WORST-CASE norms (max row norm, max |W|, spectral norm of W1) and no
//                   averaging: 30-70x the observed error — it would make a third of the pool candidates — so it is
//                   reported only.
// Neither is a proof: the per-query check (max over all candidates, the stratified sample of non-candidates, and 8 x the
// candidates' rms) is what detects a violation, and the width is then raised and the stages repeated.
static double refine_model_eps(const HeadGroup &g, const Table &A, const Table &B, bool f16, double z = 8.0) {
  const double u = f16 ? 1.0 / 2048.0 : 1.0 / 256.0, d2 = u * u / 6.0;
  double xn2 = A.rowms + B.rowms;
  if (g.feat_mode == ALINK_FEAT_ABSDIFF_HADAMARD) xn2 += A.rowms * B.absmax * B.absmax;
  const double pre = xn2 * g.w1_col_l2 * g.w1_col_l2 / (double)g.D_in;  // variance of a layer-1 pre-activation
  const double s1 = 2.0 * d2 * pre;
  const double h1n2 = (double)g.H1 * (0.5 * pre + g.b1_ms);
  const double s2 = g.w2_col_l2 * g.w2_col_l2 * (0.5 * s1 + 2.0 * d2 * h1n2 / (double)g.H1);
  const double sd = sqrt(0.5 * s2 * g.w3_gain_ms / (double)g.K);
  return z * sd * 0.25;
}
static double refine_worst_eps(const HeadGroup &g, double xmax, double xnorm, bool f16, double z = 6.0) {
  const double u = f16 ? 1.0 / 2048.0 : 1.0 / 256.0, delta = u / sqrt(3.0);
  const double a1 = xmax * g.w1_col_l2, b1 = xnorm * g.w1_absmax;
  const double s1 = delta * sqrt(2.0) * (a1 < b1 ? a1 : b1);
  const double h1max = xmax * g.w1_col_l1 + g.b1_absmax, h1norm = xnorm * g.w1_spec + g.b1_l2;
  const double a2 = h1max * g.w2_col_l2, b2 = h1norm * g.w_absmax;
  const double r2 = delta * sqrt(2.0) * (a2 < b2 ? a2 : b2);
  const double s2 = sqrt(s1 * g.w2_col_l2 * s1 * g.w2_col_l2 + r2 * r2);
  return z * s2 * g.w3_gain * 0.25;
}

// the report of a three-stage query from its statistics words (after the synchronisation that brought them)
static void refine_fill_report(alink_ctx *ctx, const unsigned int *h, double eps1, double eps2, int round) {
  float e1, e2, ea;
  double ss1, ss2;
  memcpy(&e1, h + RS_ERR1, 4);
  memcpy(&e2, h + RS_ERR2, 4);
  memcpy(&ea, h + RS_ERRA, 4);
  memcpy(&ss2, h + RS_SS2, 8);
  memcpy(&ss1, h + RS_SS1, 8);
  ctx->refine_survivors = h[RS_N2];
  ctx->refine_n1 = h[RS_N1];
  ctx->refine_nsample = (int64_t)h[RS_M1] - (int64_t)h[RS_N1];
  ctx->refine_nsample2 = (int64_t)h[RS_M2] - (int64_t)h[RS_N2];
  ctx->refine_err1 = e1;
  ctx->refine_err2 = e2;
  ctx->refine_max_err = ea > e1 ? ea : e1;                       // the stage-1 filter's largest observed error
  ctx->refine_rms_err = h[RS_N1] ? (float)sqrt(ss1 / (double)h[RS_N1]) : 0.0f;
  ctx->refine_rms2 = h[RS_M2] ? (float)sqrt(ss2 / (double)h[RS_M2]) : 0.0f;
  ctx->refine_eps_used = eps1;
  ctx->refine_eps2_used = eps2;
  ctx->refine_rounds = round;
}

// FP32_REFINE in three stages with device-resident counts (refine.cuh): ONE host synchronisation, at the end, which
// brings the result and 64 bytes of statistics. Emits the selection itself; falls back to refine_query (filter +
// FFMA, host-sized grids) for shapes the split kernel does not take, B > 4096, or when a candidate list overflows.
static int refine_query_v2(alink_ctx *ctx, const HeadGroup &g, Table &A, Table &Bt, const int32_t *ia, const int32_t *ib,
                           int64_t N, int measure, int64_t B, int64_t base, cudaEvent_t *ev, int64_t *idx_out,
                           float *util_out, uint64_t *keys_out, int out_space) {
  static const int ver_env = [] { const char *e = getenv("ALINK_REFINE_V"); return e ? atoi(e) : 2; }();
  static const int fmt_env = [] { const char *e = getenv("ALINK_REFINE_FMT"); return e ? atoi(e) : 0; }();  // 1 = force bf16
  double ma = 0, mb = 0, ra = 0, rb = 0;
  bool ok = ver_env != 1 && g.split_ok && !g.bounds_stale && g.W1h && g.W2h && fmt_env != 1 && B <= 4096 && B >= 1;
  double xmax = 0, xnorm = 0, h1max = 0;
  if (ok) {
    ALINK_TRY(table_absmax(ctx, A, &ma, &ra));
    ALINK_TRY(table_absmax(ctx, Bt, &mb, &rb));
    xmax = ma + mb;
    xnorm = ra + rb;
    if (g.feat_mode == ALINK_FEAT_ABSDIFF_HADAMARD) {
      if (ma * mb > xmax) xmax = ma * mb;
      const double hn = ra * mb < rb * ma ? ra * mb : rb * ma;
      xnorm = sqrt(xnorm * xnorm + hn * hn);
    }
    h1max = xmax * g.w1_col_l1 + g.b1_absmax;
    ok = xmax < 6.0e4 && h1max < 6.0e4 && g.w_absmax < 6.0e4 && xmax > 0.0;  // the half filter cannot overflow
  }
  ctx->refine_fallback = 0;
  if (!ok) {
    ctx->refine_version = 1;
    ctx->refine_fallback = 1;
    ALINK_TRY(refine_query(ctx, g, A, Bt, ia, ib, N, measure, B, base, ev));
    if (ev) ALINK_CUDA(ctx, cudaEventRecord(ev[4], ctx->stream));
    return emit_selection(ctx, B, idx_out, util_out, keys_out, out_space);
  }
  ctx->refine_version = 2;
  ctx->N = N; ctx->sK = g.K; ctx->sC = g.C; ctx->util_valid = false; ctx->probs_full = false;
  ctx->refine_f16 = 1;
  // rows one round of the split kernel covers: its CTA pairs each take one (tile pair, learner) unit per round
  const unsigned int wave = ((unsigned int)(ctx->sm_count / 2) << 16) | (unsigned int)g.K;  // (clusters, K) for the sample's rounding
  const unsigned int ns_min = N < 4096 ? (unsigned int)N : 4096u, ns_max = ns_min + (unsigned int)(ctx->sm_count / 2) * 256u, ns2_max = 256;
  // candidate capacities grow with the pool: at 2.5 G pairs (the 50k x 50k all-pairs matrix) the half filter's band alone
  // holds tens of millions of pairs, which is exactly what the split stage is for
  int64_t cap1 = N / 32 > (1 << 18) ? N / 32 : (1 << 18);
  if (cap1 > (1ll << 27)) cap1 = 1ll << 27;
  if (cap1 > N) cap1 = N;
  int64_t cap2l = N / 2048 > 8192 ? N / 2048 : 8192;
  if (cap2l > (1 << 21)) cap2l = 1 << 21;
  const unsigned int cap2 = (unsigned int)cap2l;
  const int64_t rows1 = cap1 + ns_max, rows2 = (int64_t)cap2 + ns2_max;
  ALINK_TRY(dev_reserve(ctx, ctx->util, sizeof(float) * (size_t)N));
  ALINK_TRY(dev_reserve(ctx, ctx->sel_state, 2 * sizeof(SelectState)));
  ALINK_TRY(dev_reserve(ctx, ctx->rstat, 32768));
  ALINK_TRY(dev_reserve(ctx, ctx->cand, sizeof(uint32_t) * (size_t)(N > rows1 ? N : rows1)));
  ALINK_TRY(dev_reserve(ctx, ctx->cia, sizeof(int32_t) * (size_t)rows1));
  ALINK_TRY(dev_reserve(ctx, ctx->cib, sizeof(int32_t) * (size_t)rows1));
  ALINK_TRY(dev_reserve(ctx, ctx->util2, sizeof(float) * (size_t)rows1));
  ALINK_TRY(dev_reserve(ctx, ctx->probs2, sizeof(float) * (size_t)rows1 * g.K * g.C));
  ALINK_TRY(dev_reserve(ctx, ctx->cia2, sizeof(int32_t) * (size_t)rows2));
  ALINK_TRY(dev_reserve(ctx, ctx->cib2, sizeof(int32_t) * (size_t)rows2));
  ALINK_TRY(dev_reserve(ctx, ctx->map2, sizeof(uint32_t) * (size_t)rows2));
  ALINK_TRY(dev_reserve(ctx, ctx->src2, sizeof(uint32_t) * (size_t)rows2));
  ALINK_TRY(dev_reserve(ctx, ctx->util_c, sizeof(float) * (size_t)rows2));
  ALINK_TRY(dev_reserve(ctx, ctx->probs, sizeof(float) * (size_t)rows2 * g.K * g.C));
  ALINK_TRY(dev_reserve(ctx, ctx->h1ws, sizeof(float) * (size_t)rows2 * g.H1 * g.K));
  const int64_t npad = next_pow2(B);
  ALINK_TRY(dev_reserve(ctx, ctx->sel_keys, sizeof(uint64_t) * (size_t)npad));
  if (!ctx->rs_host) ALINK_CUDA(ctx, cudaMallocHost((void **)&ctx->rs_host, RS_WORDS * sizeof(unsigned int)));
  float *util_a = (float *)ctx->util.p;
  SelectState *st = (SelectState *)ctx->sel_state.p + 1;
  unsigned int *rs = (unsigned int *)ctx->rstat.p + 224;    // 32 statistics words, then 3 x 2048 histogram counters
  unsigned int *hist = (unsigned int *)ctx->rstat.p + 256;
  static const int coop_env = [] { const char *e = getenv("ALINK_REFINE_COOP"); return e ? atoi(e) : 1; }();
  static int coop_blocks_per_sm = -1;
  if (coop_blocks_per_sm < 0) {
    int occ = 0, coop = 0;
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, ctx->device);
    if (coop && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, refine_stage1_coop_kernel, 256, 0) == cudaSuccess && occ > 0)
      coop_blocks_per_sm = occ < 4 ? occ : 4;
    else
      coop_blocks_per_sm = 0;
  }
  const bool coop = coop_env != 0 && coop_blocks_per_sm > 0;
  // the bounds: caller's value, else the larger of the per-format default and the a-priori statistical model
  const double apriori = refine_model_eps(g, A, Bt, true);
  ctx->refine_eps_apriori = apriori;
  ctx->refine_eps_worst = refine_worst_eps(g, xmax, xnorm, true);
  const int gi = (int)(&g - ctx->groups);
  double eps1 = ctx->refine_eps > 0.0 ? ctx->refine_eps : (REFINE_EPS_F16 > apriori ? REFINE_EPS_F16 : apriori);
  if (ctx->refine_eps <= 0.0 && ctx->refine_eps_hint[gi] > eps1) eps1 = ctx->refine_eps_hint[gi];
  double eps2 = ctx->refine_eps2 > 0.0 ? ctx->refine_eps2 : REFINE_EPS2;
  const double sc = REFINE_SCALE[measure];
  // the band of the consensus measures lives in u = min(pbar, 1 - pbar) (refine_threshold): eps bounds the probability, and
  // every learner's LOGIT is then off by at most ez = 4 sqrt(K) eps (eps is the 8-sigma bound of a mean of K learners at
  // slope 1/4). ALINK_REFINE_BAND=0 keeps the additive band in utility space for every measure.
  static const int band_env = [] { const char *e = getenv("ALINK_REFINE_BAND"); return e ? atoi(e) : 1; }();
  const int band_measure = band_env ? measure : -1;
  const bool band_on = band_measure == ALINK_UTIL_UNCERTAINTY || band_measure == ALINK_UTIL_NEG_MARGIN ||
                       band_measure == ALINK_UTIL_ENTROPY || band_measure == ALINK_UTIL_CONSENSUS_ENTROPY;
  const int band_stat = band_on ? band_measure : -1;  // error statistics in probability space (refine_finish_kernel)
  const double scv = band_on ? 1.0 : sc;              // ... where the bounds apply without a per-measure factor
  const double ezk = 4.0 * sqrt((double)g.K);
  // (1) approximate utilities of every pair (fused K-learner reduction in the kernel's epilogue)
  ALINK_TRY(tc_score(ctx, g, A, Bt, ia, ib, N, nullptr, util_a, measure, true));
  if (ev) ALINK_CUDA(ctx, cudaEventRecord(ev[2], ctx->stream));
  // (2) B-th largest approximate utility tau1 (inside the cooperative stage-1 kernel when available)
  KeySrc asrc{util_a, nullptr, 0, nullptr};
  if (!coop) ALINK_TRY(select_threshold(ctx, asrc, N, B, st, 3));
  int64_t blocks = (N + 2047) / 2048;
  const int64_t capb = (int64_t)ctx->sm_count * 8;
  if (blocks > capb) blocks = capb;
  for (int round = 1;; ++round) {
    if (eps2 * 4.0 > eps1) eps2 = eps1 / 4.0;
    // (3) C1 = {u~ >= tau1 - 2 eps1}; + the stratified sample; table rows of both
    if (coop) {
      ALINK_CUDA(ctx, cudaMemsetAsync(rs, 0, (RS_WORDS + 3 * 2048) * sizeof(unsigned int), ctx->stream));
      Stage1Args sa1;
      sa1.util = util_a; sa1.N = N; sa1.B = (unsigned long long)B; sa1.margin = (float)(2.0 * eps1 * sc);
      sa1.measure = band_measure; sa1.eps = eps1; sa1.ez = ezk * eps1;
      sa1.cap1 = (unsigned int)cap1; sa1.ns_min = ns_min; sa1.ns_max = ns_max; sa1.wave_rows = wave;
      sa1.hist = hist; sa1.rs = rs; sa1.cand1 = (uint32_t *)ctx->cand.p; sa1.ia = ia; sa1.ib = ib;
      sa1.G = ia ? 0u : ctx->ap_G; sa1.pair0 = ctx->ap_pair0; sa1.cia = (int32_t *)ctx->cia.p; sa1.cib = (int32_t *)ctx->cib.p;
      void *kargs[] = {(void *)&sa1};
      int64_t cb = (N + 2047) / 2048;
      const int64_t cmax = (int64_t)(ctx->sm_count - ctx->reserved_sms) * coop_blocks_per_sm;
      if (cb > cmax) cb = cmax;
      if (cb < 1) cb = 1;
      ALINK_CUDA(ctx, cudaLaunchCooperativeKernel((const void *)refine_stage1_coop_kernel, dim3((unsigned)cb), dim3(256), kargs, 0, ctx->stream));
      ALINK_LAUNCHED(ctx);
    } else {
      ALINK_CUDA(ctx, cudaMemsetAsync(rs, 0, RS_WORDS * sizeof(unsigned int), ctx->stream));
      refine_mark_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(util_a, N, st, nullptr, (float)(2.0 * eps1 * sc),
                                                                    (uint32_t *)ctx->cand.p, rs + RS_COUNT1, (unsigned int)cap1,
                                                                    band_measure, eps1, ezk * eps1);
      ALINK_LAUNCHED(ctx);
      refine_stage1_finish_kernel<<<(unsigned)(rows1 / 4096 > 64 ? (rows1 / 4096 > capb ? capb : rows1 / 4096) : 64), 256, 0, ctx->stream>>>(
          rs, (unsigned int)cap1, N, ns_min, ns_max, wave, (uint32_t *)ctx->cand.p, ia, ib, ia ? 0u : ctx->ap_G, ctx->ap_pair0,
          (int32_t *)ctx->cia.p, (int32_t *)ctx->cib.p);
      ALINK_LAUNCHED(ctx);
    }
    // (4) split-operand tensor-core utilities u2 of C1 u S
    {
      const uint32_t apG = ctx->ap_G;
      ctx->ap_G = 0;  // the rows are explicit now
      const int rc = tc_score_split(ctx, g, A, Bt, (const int32_t *)ctx->cia.p, (const int32_t *)ctx->cib.p, rows1, rs + RS_M1,
                                    (float *)ctx->probs2.p, nullptr, measure, xmax, h1max);
      ctx->ap_G = apG;
      ALINK_TRY(rc);
      // the kernel's unit of work is (tile pair, learner): the K-learner utility is this streaming pass
      ALINK_TRY(launch_utility(ctx, (const float *)ctx->probs2.p, rows1, g.K, g.C, measure, (float *)ctx->util2.p, nullptr, rs + RS_M1));
    }
    // (5) tau2, C2 = {u2 >= tau2 - 2 eps2} + every 16th sample row; their table rows
    refine_stage2_kernel<<<1, 1024, 0, ctx->stream>>>(rs, (const float *)ctx->util2.p, (unsigned long long)B, (float)(2.0 * eps2 * sc),
                                                      band_measure, eps2, ezk * eps2, cap2,
                                                      ns2_max, (const uint32_t *)ctx->cand.p, (const int32_t *)ctx->cia.p,
                                                      (const int32_t *)ctx->cib.p, (int32_t *)ctx->cia2.p, (int32_t *)ctx->cib2.p,
                                                      (uint32_t *)ctx->map2.p, (uint32_t *)ctx->src2.p);
    ALINK_LAUNCHED(ctx);
    // (6) bit-exact re-score of C2 u S2 (oracle summation order, CUDA cores)
    {
      const uint32_t apG = ctx->ap_G;
      ctx->ap_G = 0;
      const int rc = ffma_head_chunk(ctx, g, (const float *)A.p, (const float *)Bt.p, A.D, (const int32_t *)ctx->cia2.p,
                                     (const int32_t *)ctx->cib2.p, 0, (int)rows2, (float *)ctx->h1ws.p, rows2 * g.H1,
                                     (float *)ctx->probs.p, rs + RS_M2);
      ctx->ap_G = apG;
      ALINK_TRY(rc);
    }
    // (7) exact utilities + the error statistics of both filters
    {
      int64_t fbl = rows1 / 4096 > 48 ? rows1 / 4096 : 48;
      if (fbl > capb) fbl = capb;
      const unsigned fb = (unsigned)fbl;
#define ALINK_FINISH(KT)                                                                                                   \
  refine_finish_kernel<KT><<<fb, 256, 0, ctx->stream>>>(rs, (const float *)ctx->probs.p, g.K, g.C, measure, util_a,        \
                                                        (const float *)ctx->util2.p, (const uint32_t *)ctx->cand.p,        \
                                                        (const uint32_t *)ctx->map2.p, (const uint32_t *)ctx->src2.p,      \
                                                        (float *)ctx->util_c.p, (const float *)ctx->probs2.p, band_stat)
      switch (g.K) {
        case 1: ALINK_FINISH(1); break;
        case 3: ALINK_FINISH(3); break;
        case 5: ALINK_FINISH(5); break;
        default: ALINK_FINISH(0); break;
      }
#undef ALINK_FINISH
      ALINK_LAUNCHED(ctx);
    }
    if (ev && round == 1) ALINK_CUDA(ctx, cudaEventRecord(ev[3], ctx->stream));
    // (8) exact top-B of C2, keyed by the pairs' original indices
    KeySrc csrc{(const float *)ctx->util_c.p, nullptr, (uint64_t)base, (const uint32_t *)ctx->map2.p};
    topk_small_kernel<<<1, 1024, 0, ctx->stream>>>(csrc, (int)cap2, (int)B, (int)(npad < 2 ? 2 : npad), (uint64_t *)ctx->sel_keys.p,
                                                   rs + RS_N2);
    ALINK_LAUNCHED(ctx);
    if (ctx->defer_req) {
      // a sharded step does not stop here: the verdict is evaluated on the device and travels with the keys
      // (alink_query_sharded reads the statistics with the merged result and calls refine_fill_report)
      VerifyConsts vc{eps1, eps2, scv, band_on ? expm1(ezk * eps1) : -1.0};
      refine_verify_kernel<<<1, 32, 0, ctx->stream>>>(rs, vc);
      ALINK_LAUNCHED(ctx);
      ALINK_CUDA(ctx, cudaMemcpyAsync(ctx->rs_host, rs, RS_WORDS * sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->stream));
      if (ev) ALINK_CUDA(ctx, cudaEventRecord(ev[4], ctx->stream));
      ALINK_TRY(emit_selection(ctx, B, idx_out, util_out, keys_out, out_space));  // (device outputs: no synchronisation)
      ctx->deferred = true;
      ctx->defer_eps1 = eps1; ctx->defer_eps2 = eps2; ctx->defer_scv = scv;
      ctx->defer_ok_dev = rs + RS_OK;
      return ALINK_OK;
    }
    ALINK_CUDA(ctx, cudaMemcpyAsync(ctx->rs_host, rs, RS_WORDS * sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->stream));
    if (ev && round == 1) ALINK_CUDA(ctx, cudaEventRecord(ev[4], ctx->stream));
    ALINK_TRY(emit_selection(ctx, B, idx_out, util_out, keys_out, out_space));
    ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // the one host synchronisation of the step
    const unsigned int *h = ctx->rs_host;
    refine_fill_report(ctx, h, eps1, eps2, round);
    const float e1 = ctx->refine_err1, e2 = ctx->refine_err2;
    float ea;
    memcpy(&ea, h + RS_ERRA, 4);
    if (h[RS_FLAGS]) {  // a candidate list overflowed (plateaus of tied utilities): the host-sized pipeline takes any size
      ctx->refine_version = 1;
      ctx->refine_fallback = 2 + (int)(h[RS_FLAGS] & 3u);  // 3: stage-1 list, 4: stage-2 list, 5: both
      ALINK_TRY(refine_query(ctx, g, A, Bt, ia, ib, N, measure, B, base, nullptr));
      return emit_selection(ctx, B, idx_out, util_out, keys_out, out_space);
    }
    // Both bounds must hold on everything we can check: stage 2 against the exact values (C2 u S2), stage 1 against
    // stage 2 (C1 u S, to within eps2) and against the exact values. Otherwise widen past what was seen and redo.
    float rel1;
    memcpy(&rel1, h + RS_REL1, 4);
    const double seen2 = (double)e2 / scv;
    const double s1a = (double)e1 / scv + eps2, s1b = (double)ea / scv;  // |u~ - u| <= |u~ - u2| + eps2
    double seen1 = s1a > s1b ? s1a : s1b;
    const double tail1 = 8.0 * (double)ctx->refine_rms_err / scv;  // 8 sigma of the candidates' error must fit as well
    // the multiplicative band's premise: relative error of u below e^ez - 1 (checked where u >= 1e-3)
    const bool badrel = band_on && (double)rel1 > expm1(ezk * eps1);
    if (badrel && (double)rel1 / ezk * 1.05 > seen1) seen1 = log1p((double)rel1) / ezk * 1.05;
    const bool bad2 = seen2 > eps2, bad1 = seen1 > eps1 || tail1 > eps1 || badrel;
    if (tail1 > 2.0 * seen1) seen1 = 0.5 * tail1;
    if ((!bad1 && !bad2) || round >= 5) {
      if (bad1 || bad2) {  // could not be verified in 5 rounds: run it exactly
        ctx->refine_version = 1;
        ctx->refine_fallback = 6;
        ALINK_TRY(refine_query(ctx, g, A, Bt, ia, ib, N, measure, B, base, nullptr));
        return emit_selection(ctx, B, idx_out, util_out, keys_out, out_space);
      }
      break;
    }
    if (bad2) eps2 = 2.0 * seen2;
    if (bad1) {
      eps1 = 2.0 * seen1;
      if (ctx->refine_eps <= 0.0) ctx->refine_eps_hint[gi] = eps1;  // the group's next query starts from here
    }
  }
  return ALINK_OK;
}

// FP32_REFINE for vote entropy (refine.cuh, last section): the half filter's votes are certain wherever every learner's
// |p1 - p0| clears the per-learner bound; the rest is re-scored with split operands, and what is still within that
// stage's bound exactly. All N utilities are exact afterwards; the selection is the ordinary top-B over them.
// *handled = false: shape not taken, or a list overflowed / a bound could not be verified (the caller runs the exact
// pipeline; ctx->refine_fallback says why).
static int refine_vote_query(alink_ctx *ctx, const HeadGroup &g, Table &A, Table &Bt, const int32_t *ia, const int32_t *ib,
                             int64_t N, int64_t B, int64_t base, cudaEvent_t *ev, int64_t *idx_out, float *util_out,
                             uint64_t *keys_out, int out_space, bool *handled) {
  static const int ver_env = [] { const char *e = getenv("ALINK_REFINE_V"); return e ? atoi(e) : 2; }();
  static const int vote_env = [] { const char *e = getenv("ALINK_REFINE_VOTE"); return e ? atoi(e) : 1; }();
  static const int fmt_env = [] { const char *e = getenv("ALINK_REFINE_FMT"); return e ? atoi(e) : 0; }();
  *handled = false;
  ctx->refine_fallback = 1;
  ctx->refine_version = 1;
  double ma = 0, mb = 0, ra = 0, rb = 0;
  if (ver_env == 1 || !vote_env || fmt_env == 1 || !g.split_ok || g.bounds_stale || !g.W1h || !g.W2h || !tc_takes_vmin(g))
    return ALINK_OK;
  ALINK_TRY(table_absmax(ctx, A, &ma, &ra));
  ALINK_TRY(table_absmax(ctx, Bt, &mb, &rb));
  double xmax = ma + mb, xnorm = ra + rb;
  if (g.feat_mode == ALINK_FEAT_ABSDIFF_HADAMARD) {
    if (ma * mb > xmax) xmax = ma * mb;
    const double hn = ra * mb < rb * ma ? ra * mb : rb * ma;
    xnorm = sqrt(xnorm * xnorm + hn * hn);
  }
  const double h1max = xmax * g.w1_col_l1 + g.b1_absmax;
  if (!(xmax < 6.0e4 && h1max < 6.0e4 && g.w_absmax < 6.0e4 && xmax > 0.0)) return ALINK_OK;
  const int measure = ALINK_UTIL_VOTE_ENTROPY;
  ctx->refine_version = 2;
  ctx->refine_fallback = 0;
  ctx->N = N; ctx->sK = g.K; ctx->sC = g.C; ctx->util_valid = false; ctx->probs_full = false;
  ctx->refine_f16 = 1;
  const unsigned int wave = ((unsigned int)(ctx->sm_count / 2) << 16) | (unsigned int)g.K;
  const unsigned int ns_min = N < 4096 ? (unsigned int)N : 4096u, ns_max = ns_min + (unsigned int)(ctx->sm_count / 2) * 256u, ns2_max = 256;
  int64_t cap1 = N / 4 > (1 << 18) ? N / 4 : (1 << 18);  // a vote is uncertain far more often than a utility is near tau
  if (cap1 > (1ll << 27)) cap1 = 1ll << 27;
  if (cap1 > N) cap1 = N;
  int64_t cap2l = N / 256 > 8192 ? N / 256 : 8192;
  if (cap2l > (1 << 20)) cap2l = 1 << 20;
  const unsigned int cap2 = (unsigned int)cap2l;
  const int64_t rows1 = cap1 + ns_max, rows2 = (int64_t)cap2 + ns2_max;
  ALINK_TRY(dev_reserve(ctx, ctx->util, sizeof(float) * (size_t)N));
  ALINK_TRY(dev_reserve(ctx, ctx->util2, sizeof(float) * (size_t)N));  // vmin of the filter
  ALINK_TRY(dev_reserve(ctx, ctx->rstat, 32768));
  ALINK_TRY(dev_reserve(ctx, ctx->cand, sizeof(uint32_t) * (size_t)rows1));
  ALINK_TRY(dev_reserve(ctx, ctx->cia, sizeof(int32_t) * (size_t)rows1));
  ALINK_TRY(dev_reserve(ctx, ctx->cib, sizeof(int32_t) * (size_t)rows1));
  ALINK_TRY(dev_reserve(ctx, ctx->probs2, sizeof(float) * (size_t)rows1 * g.K * g.C));
  ALINK_TRY(dev_reserve(ctx, ctx->cia2, sizeof(int32_t) * (size_t)rows2));
  ALINK_TRY(dev_reserve(ctx, ctx->cib2, sizeof(int32_t) * (size_t)rows2));
  ALINK_TRY(dev_reserve(ctx, ctx->map2, sizeof(uint32_t) * (size_t)rows2));
  ALINK_TRY(dev_reserve(ctx, ctx->src2, sizeof(uint32_t) * (size_t)rows2));
  ALINK_TRY(dev_reserve(ctx, ctx->probs, sizeof(float) * (size_t)rows2 * g.K * g.C));
  ALINK_TRY(dev_reserve(ctx, ctx->h1ws, sizeof(float) * (size_t)rows2 * g.H1 * g.K));
  if (!ctx->rs_host) ALINK_CUDA(ctx, cudaMallocHost((void **)&ctx->rs_host, RS_WORDS * sizeof(unsigned int)));
  float *util_a = (float *)ctx->util.p, *vmin = (float *)ctx->util2.p;
  unsigned int *rs = (unsigned int *)ctx->rstat.p + 224;
  const double apriori = refine_model_eps(g, A, Bt, true);
  ctx->refine_eps_apriori = apriori;
  ctx->refine_eps_worst = refine_worst_eps(g, xmax, xnorm, true);
  const int gi = (int)(&g - ctx->groups);
  // eps1 / eps2 bound the error of a MEAN of K learners; one learner's probability is allowed sqrt(K) times that
  const double rk = sqrt((double)g.K);
  double eps1 = ctx->refine_eps > 0.0 ? ctx->refine_eps : (REFINE_EPS_F16 > apriori ? REFINE_EPS_F16 : apriori);
  if (ctx->refine_eps <= 0.0 && ctx->refine_eps_hint[gi] > eps1) eps1 = ctx->refine_eps_hint[gi];
  double eps2 = ctx->refine_eps2 > 0.0 ? ctx->refine_eps2 : REFINE_EPS2;
  // (1) half-precision votes and vote entropy of every pair + how far each pair is from a flipped vote
  ALINK_TRY(tc_score(ctx, g, A, Bt, ia, ib, N, nullptr, util_a, measure, true, vmin));
  if (ev) ALINK_CUDA(ctx, cudaEventRecord(ev[2], ctx->stream));
  int64_t blocks = (N + 2047) / 2048;
  const int64_t capb = (int64_t)ctx->sm_count * 8;
  if (blocks > capb) blocks = capb;
  for (int round = 1;; ++round) {
    if (eps2 * 4.0 > eps1) eps2 = eps1 / 4.0;
    // |p1 - p0| moves by twice a learner's probability error (<= sqrt(K) eps); x 1.5 for safety
    const float thr1 = (float)(3.0 * eps1 * rk), thr2 = (float)(3.0 * eps2 * rk);
    ALINK_CUDA(ctx, cudaMemsetAsync(rs, 0, RS_WORDS * sizeof(unsigned int), ctx->stream));
    // (2) C1 = {some learner within thr1 of a flip} + the stratified sample; their table rows
    vote_mark_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(vmin, N, thr1, (uint32_t *)ctx->cand.p, rs + RS_COUNT1, (unsigned int)cap1);
    ALINK_LAUNCHED(ctx);
    refine_stage1_finish_kernel<<<(unsigned)(rows1 / 4096 > 64 ? (rows1 / 4096 > capb ? capb : rows1 / 4096) : 64), 256, 0, ctx->stream>>>(
        rs, (unsigned int)cap1, N, ns_min, ns_max, wave, (uint32_t *)ctx->cand.p, ia, ib, ia ? 0u : ctx->ap_G, ctx->ap_pair0,
        (int32_t *)ctx->cia.p, (int32_t *)ctx->cib.p);
    ALINK_LAUNCHED(ctx);
    // (3) split-operand tensor-core probabilities of C1 u S
    {
      const uint32_t apG = ctx->ap_G;
      ctx->ap_G = 0;
      const int rc = tc_score_split(ctx, g, A, Bt, (const int32_t *)ctx->cia.p, (const int32_t *)ctx->cib.p, rows1, rs + RS_M1,
                                    (float *)ctx->probs2.p, nullptr, measure, xmax, h1max);
      ctx->ap_G = apG;
      ALINK_TRY(rc);
    }
    // (4) their votes replace the filter's; C2 = {still within thr2 of a flip} + 256 sample rows
    {
      int64_t fbl = rows1 / 4096 > 48 ? rows1 / 4096 : 48;
      if (fbl > capb) fbl = capb;
      const unsigned fb = (unsigned)fbl;
#define ALINK_VOTE2(KT)                                                                                                        \
  vote_stage2_kernel<KT><<<fb, 256, 0, ctx->stream>>>(rs, (const float *)ctx->probs2.p, g.K, g.C, util_a, vmin,                 \
                                                      (const uint32_t *)ctx->cand.p, thr1, thr2, cap2,                          \
                                                      (const int32_t *)ctx->cia.p, (const int32_t *)ctx->cib.p,                 \
                                                      (int32_t *)ctx->cia2.p, (int32_t *)ctx->cib2.p, (uint32_t *)ctx->map2.p,  \
                                                      (uint32_t *)ctx->src2.p)
      switch (g.K) {
        case 1: ALINK_VOTE2(1); break;
        case 3: ALINK_VOTE2(3); break;
        case 5: ALINK_VOTE2(5); break;
        default: ALINK_VOTE2(0); break;
      }
#undef ALINK_VOTE2
      ALINK_LAUNCHED(ctx);
      vote_stage2_finish_kernel<<<1, 256, 0, ctx->stream>>>(rs, cap2, ns2_max, (const uint32_t *)ctx->cand.p, (const int32_t *)ctx->cia.p,
                                                            (const int32_t *)ctx->cib.p, (int32_t *)ctx->cia2.p, (int32_t *)ctx->cib2.p,
                                                            (uint32_t *)ctx->map2.p, (uint32_t *)ctx->src2.p);
      ALINK_LAUNCHED(ctx);
    }
    // (5) bit-exact re-score of C2 u S2, exact votes
    {
      const uint32_t apG = ctx->ap_G;
      ctx->ap_G = 0;
      const int rc = ffma_head_chunk(ctx, g, (const float *)A.p, (const float *)Bt.p, A.D, (const int32_t *)ctx->cia2.p,
                                     (const int32_t *)ctx->cib2.p, 0, (int)rows2, (float *)ctx->h1ws.p, rows2 * g.H1,
                                     (float *)ctx->probs.p, rs + RS_M2);
      ctx->ap_G = apG;
      ALINK_TRY(rc);
      const unsigned fb = (unsigned)(rows2 / 2048 > 8 ? rows2 / 2048 : 8);
#define ALINK_VOTEF(KT)                                                                                                        \
  vote_finish_kernel<KT><<<fb, 256, 0, ctx->stream>>>(rs, (const float *)ctx->probs.p, (const float *)ctx->probs2.p, g.K, g.C,  \
                                                      util_a, (const uint32_t *)ctx->map2.p, (const uint32_t *)ctx->src2.p, thr2)
      switch (g.K) {
        case 1: ALINK_VOTEF(1); break;
        case 3: ALINK_VOTEF(3); break;
        case 5: ALINK_VOTEF(5); break;
        default: ALINK_VOTEF(0); break;
      }
#undef ALINK_VOTEF
      ALINK_LAUNCHED(ctx);
    }
    if (ev && round == 1) ALINK_CUDA(ctx, cudaEventRecord(ev[3], ctx->stream));
    // (6) top-B of the N exact utilities (ties by lowest pair index, as everywhere)
    KeySrc src{util_a, nullptr, (uint64_t)base};
    ALINK_TRY(topk_device(ctx, src, N, B));
    ALINK_CUDA(ctx, cudaMemcpyAsync(ctx->rs_host, rs, RS_WORDS * sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->stream));
    if (ev && round == 1) ALINK_CUDA(ctx, cudaEventRecord(ev[4], ctx->stream));
    ALINK_TRY(emit_selection(ctx, B, idx_out, util_out, keys_out, out_space));
    ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const unsigned int *h = ctx->rs_host;
    float e1, e2;
    double ss1, ss2;
    memcpy(&e1, h + RS_ERR1, 4);
    memcpy(&e2, h + RS_ERR2, 4);
    memcpy(&ss1, h + RS_SS1, 8);
    memcpy(&ss2, h + RS_SS2, 8);
    // reported at the scale of eps (a mean of K learners): the per-learner figures over sqrt(K)
    ctx->refine_survivors = h[RS_N2];
    ctx->refine_n1 = h[RS_N1];
    ctx->refine_nsample = (int64_t)h[RS_M1] - (int64_t)h[RS_N1];
    ctx->refine_nsample2 = (int64_t)h[RS_M2] - (int64_t)h[RS_N2];
    ctx->refine_err1 = (float)(e1 / rk);
    ctx->refine_err2 = (float)(e2 / rk);
    ctx->refine_max_err = ctx->refine_err1;
    ctx->refine_rms_err = h[RS_N1] ? (float)(sqrt(ss1 / (double)h[RS_N1]) / rk) : 0.0f;
    ctx->refine_rms2 = h[RS_M2] ? (float)(sqrt(ss2 / (double)h[RS_M2]) / rk) : 0.0f;
    ctx->refine_eps_used = eps1;
    ctx->refine_eps2_used = eps2;
    ctx->refine_rounds = round;
    if (h[RS_FLAGS] & 3u) {  // a candidate list overflowed: the exact pipeline takes any pool
      ctx->refine_version = 1;
      ctx->refine_fallback = 2 + (int)(h[RS_FLAGS] & 3u);
      return ALINK_OK;  // *handled stays false
    }
    // stage 2 against the exact values (C2 u S2), stage 1 against stage 2 (C1 u S, to within eps2); a sampled pair the
    // filter (flag 4) or the split stage (flag 8) called certain and whose votes changed is a violation whatever the errors
    const bool wit1 = (h[RS_FLAGS] & 4u) != 0, wit2 = (h[RS_FLAGS] & 8u) != 0;
    const double seen2 = (double)e2 / rk;
    double seen1 = (double)e1 / rk + eps2;
    const double tail1 = 8.0 * (double)ctx->refine_rms_err;
    const bool bad2 = seen2 > eps2 || wit2, bad1 = seen1 > eps1 || tail1 > eps1 || wit1;
    if (tail1 > 2.0 * seen1) seen1 = 0.5 * tail1;
    if ((!bad1 && !bad2) || round >= 5) {
      if (bad1 || bad2) {
        ctx->refine_version = 1;
        ctx->refine_fallback = 6;
        return ALINK_OK;
      }
      break;
    }
    if (bad2) eps2 = 2.0 * (seen2 > eps2 ? seen2 : eps2);
    if (bad1) {
      eps1 = 2.0 * (seen1 > eps1 ? seen1 : eps1);
      if (ctx->refine_eps <= 0.0) ctx->refine_eps_hint[gi] = eps1;
    }
  }
  ctx->util_valid = true;
  ctx->util_measure = measure;
  *handled = true;
  return ALINK_OK;
}

int alink_set_refine_eps(alink_ctx *ctx, double eps) {
  if (!ctx) return ALINK_E_INVALID;
  if (!(eps >= 0.0 && eps < 1.0)) return alink_fail(ctx, ALINK_E_INVALID, "eps outside [0, 1) (0 = per-format default)");
  ctx->refine_eps = eps;
  return ALINK_OK;
}

int alink_set_refine_eps2(alink_ctx *ctx, double eps2) {
  if (!ctx) return ALINK_E_INVALID;
  if (!(eps2 >= 0.0 && eps2 < 1.0)) return alink_fail(ctx, ALINK_E_INVALID, "eps2 outside [0, 1) (0 = default)");
  ctx->refine_eps2 = eps2;
  return ALINK_OK;
}

int alink_refine_report_get(const alink_ctx *ctx, alink_refine_report *out) {
  if (!ctx || !out) return ALINK_E_INVALID;
  memset(out, 0, sizeof(*out));
  out->version = ctx->refine_version;
  out->filter_f16 = ctx->refine_f16;
  out->rounds = ctx->refine_rounds;
  out->fallback = ctx->refine_fallback;
  out->n_exact = ctx->refine_survivors;
  out->eps1_used = ctx->refine_eps_used;
  out->err1_max = ctx->refine_max_err;
  out->err1_rms = ctx->refine_rms_err;
  if (ctx->refine_version == 2) {
    out->n_stage2 = ctx->refine_n1;
    out->n_sample = ctx->refine_nsample;
    out->n_sample_exact = ctx->refine_nsample2;
    out->eps1_apriori = ctx->refine_eps_apriori;
    out->eps1_worst = ctx->refine_eps_worst;
    out->eps2_used = ctx->refine_eps2_used;
    out->err2_max = ctx->refine_err2;
    out->err2_rms = ctx->refine_rms2;
  }
  return ALINK_OK;
}

int alink_refine_stats(const alink_ctx *ctx, int64_t *survivors, float *max_err) {
  if (!ctx) return ALINK_E_INVALID;
  if (survivors) *survivors = ctx->refine_survivors;
  if (max_err) *max_err = ctx->refine_max_err;
  return ALINK_OK;
}

int alink_refine_detail(const alink_ctx *ctx, int *filter_f16, double *eps_used, float *rms_err, int *rounds) {
  if (!ctx) return ALINK_E_INVALID;
  if (filter_f16) *filter_f16 = ctx->refine_f16;
  if (eps_used) *eps_used = ctx->refine_eps_used;
  if (rms_err) *rms_err = ctx->refine_rms_err;
  if (rounds) *rounds = ctx->refine_rounds;
  return ALINK_OK;
}

int alink_query(alink_ctx *ctx, int group, int slot_a, int slot_b, const int32_t *ia, const int32_t *ib, int64_t N,
                int in_space, int measure, int64_t B, int64_t base, int64_t *idx_out, float *util_out,
                uint64_t *keys_out, int out_space, alink_timing *timing) {
  if (!ctx) return ALINK_E_INVALID;
  if (slot_a < 0 || slot_a >= ALINK_MAX_SLOTS || slot_b < 0 || slot_b >= ALINK_MAX_SLOTS)
    return alink_fail(ctx, ALINK_E_INVALID, "bad slot");
  if ((ia == nullptr) != (ib == nullptr)) return alink_fail(ctx, ALINK_E_INVALID, "ia and ib must both be given or both NULL");
  if (B < 0 || B > N) return alink_fail(ctx, ALINK_E_INVALID, "B=%lld outside [0, N=%lld]", (long long)B, (long long)N);
  if (base < 0 || base + N > 0xFFFFFFFFll) return alink_fail(ctx, ALINK_E_INVALID, "base + N exceeds 2^32 - 1");
  if (measure < 0 || measure >= ALINK_UTIL_COUNT) return alink_fail(ctx, ALINK_E_INVALID, "bad measure %d", measure);
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  const int64_t l0 = ctx->launches;
  cudaEvent_t *e = ctx->ev;
  if (timing) ALINK_CUDA(ctx, cudaEventRecord(e[0], ctx->stream));
  const void *dia, *dib;
  if (ctx->copy_busy && cudaEventQuery(ctx->copy_done) == cudaSuccess) ctx->copy_busy = false;
  ALINK_TRY(stage_in(ctx, ia, sizeof(int32_t) * (size_t)N, in_space, ctx->idx_a, &dia));
  ALINK_TRY(stage_in(ctx, ib, sizeof(int32_t) * (size_t)N, in_space, ctx->idx_b, &dib));
  if (timing) ALINK_CUDA(ctx, cudaEventRecord(e[1], ctx->stream));
  ALINK_TRY(check_indices(ctx, (const int32_t *)dia, (const int32_t *)dib, N, slot_a, slot_b));
  ALINK_TRY(table_wait(ctx, slot_a));
  ALINK_TRY(table_wait(ctx, slot_b));
  ctx->refine_survivors = 0;
  ctx->refine_max_err = 0.0f;
  ctx->refine_version = 0;
  if (group >= 0 && group < ALINK_MAX_GROUPS && ctx->groups[group].set &&
      ctx->groups[group].compute == ALINK_COMPUTE_FP32_REFINE && ctx->groups[group].tc_ok && B > 0 && N >= 8192) {
    const HeadGroup &g = ctx->groups[group];
    Table &TA = ctx->tables[slot_a], &TB = ctx->tables[slot_b];
    if (!TA.p || !TB.p) return alink_fail(ctx, ALINK_E_STATE, "embedding table slot not set (alink_set_table)");
    if (TA.D != TB.D || TA.dtype != ALINK_F32 || TB.dtype != ALINK_F32 || (g.feat_mode ? 2 * TA.D : TA.D) != g.D_in)
      return alink_fail(ctx, ALINK_E_INVALID, "FP32_REFINE needs fp32 tables matching the heads' fan-in");
    if (!dia && !ctx->ap_G && (N > TA.M || N > TB.M))
      return alink_fail(ctx, ALINK_E_INVALID, "identity pairs exceed table rows");
    bool handled = true;
    if (measure == ALINK_UTIL_VOTE_ENTROPY)
      ALINK_TRY(refine_vote_query(ctx, g, TA, TB, (const int32_t *)dia, (const int32_t *)dib, N, B, base, timing ? e : nullptr,
                                  idx_out, util_out, keys_out, out_space, &handled));
    else
      ALINK_TRY(refine_query_v2(ctx, g, TA, TB, (const int32_t *)dia, (const int32_t *)dib, N, measure, B, base,
                                timing ? e : nullptr, idx_out, util_out, keys_out, out_space));
    if (handled) {
    if (ctx->deferred) {  // alink_query_sharded: no synchronisation here, the times are read with the merged result
      if (timing) ALINK_CUDA(ctx, cudaEventRecord(e[5], ctx->stream));
      return ALINK_OK;
    }
    if (timing) {
      ALINK_CUDA(ctx, cudaEventRecord(e[5], ctx->stream));
      ALINK_CUDA(ctx, cudaEventSynchronize(e[5]));
      cudaEventElapsedTime(&timing->ms_h2d, e[0], e[1]);
      cudaEventElapsedTime(&timing->ms_score, e[1], e[2]);    // tensor-core filter
      cudaEventElapsedTime(&timing->ms_utility, e[2], e[3]);  // candidate marking + exact re-score
      cudaEventElapsedTime(&timing->ms_select, e[3], e[4]);
      cudaEventElapsedTime(&timing->ms_d2h, e[4], e[5]);
      cudaEventElapsedTime(&timing->ms_total, e[0], e[5]);
      timing->launches = ctx->launches - l0;
    }
    return ALINK_OK;
    }  // (else: vote entropy on a shape / pool the refine path does not take — the exact pipeline below)
  }
  ALINK_TRY(score_impl(ctx, group, ctx->tables[slot_a], ctx->tables[slot_b], (const int32_t *)dia,
                       (const int32_t *)dib, N, measure));
  if (timing) ALINK_CUDA(ctx, cudaEventRecord(e[2], ctx->stream));
  if (!(ctx->util_valid && ctx->util_measure == measure)) ALINK_TRY(utility_impl(ctx, measure));
  if (timing) ALINK_CUDA(ctx, cudaEventRecord(e[3], ctx->stream));
  if (B > 0) {
    KeySrc src{(const float *)ctx->util.p, nullptr, (uint64_t)base};
    ALINK_TRY(topk_device(ctx, src, N, B));
  }
  if (timing) ALINK_CUDA(ctx, cudaEventRecord(e[4], ctx->stream));
  ALINK_TRY(emit_selection(ctx, B, idx_out, util_out, keys_out, out_space));
  if (timing) {
    ALINK_CUDA(ctx, cudaEventRecord(e[5], ctx->stream));
    ALINK_CUDA(ctx, cudaEventSynchronize(e[5]));
    cudaEventElapsedTime(&timing->ms_h2d, e[0], e[1]);
    cudaEventElapsedTime(&timing->ms_score, e[1], e[2]);
    cudaEventElapsedTime(&timing->ms_utility, e[2], e[3]);
    cudaEventElapsedTime(&timing->ms_select, e[3], e[4]);
    cudaEventElapsedTime(&timing->ms_d2h, e[4], e[5]);
    cudaEventElapsedTime(&timing->ms_total, e[0], e[5]);
    timing->launches = ctx->launches - l0;
  }
  return ALINK_OK;
}

// ------------------------------------------------------------------------------------------ all pairs
// Probe rows [row0, row0 + nrows) of slot_probe against EVERY row of slot_gallery, pair index = i * G + j:
// the row-at-a-time loops of utilities/generateMatrixDFW.py:28-35 and code/ALINK_MTP.py:279-288.
static int allpairs_begin(alink_ctx *ctx, int slot_probe, int slot_gallery, int64_t row0, int64_t nrows, int64_t *N,
                          int64_t *base) {
  if (slot_probe < 0 || slot_probe >= ALINK_MAX_SLOTS || slot_gallery < 0 || slot_gallery >= ALINK_MAX_SLOTS)
    return alink_fail(ctx, ALINK_E_INVALID, "bad slot");
  ALINK_TRY(table_wait(ctx, slot_probe));
  ALINK_TRY(table_wait(ctx, slot_gallery));
  const Table &P = ctx->tables[slot_probe], &G = ctx->tables[slot_gallery];
  if (!P.p || !G.p) return alink_fail(ctx, ALINK_E_STATE, "embedding table slot not set (alink_set_table)");
  if (row0 < 0 || nrows < 0 || row0 + nrows > P.M) return alink_fail(ctx, ALINK_E_INVALID, "probe rows outside the table");
  if ((double)(row0 + nrows) * (double)G.M > 4294967294.0)
    return alink_fail(ctx, ALINK_E_INVALID, "pair index (row0 + nrows) * G exceeds 2^32 - 2");
  *N = nrows * G.M;
  *base = row0 * G.M;
  ctx->ap_G = (uint32_t)G.M;
  ctx->ap_pair0 = *base;
  return ALINK_OK;
}

int alink_score_allpairs(alink_ctx *ctx, int group, int slot_probe, int slot_gallery, int64_t row0, int64_t nrows,
                         float *per_learner, float *consensus, int out_space) {
  if (!ctx) return ALINK_E_INVALID;
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  int64_t N, base;
  ALINK_TRY(allpairs_begin(ctx, slot_probe, slot_gallery, row0, nrows, &N, &base));
  int rc = score_impl(ctx, group, ctx->tables[slot_probe], ctx->tables[slot_gallery], nullptr, nullptr, N);
  ctx->ap_G = 0;
  ctx->ap_pair0 = 0;
  if (rc != ALINK_OK) return rc;
  return emit_scores(ctx, per_learner, consensus, out_space);
}

int alink_query_allpairs(alink_ctx *ctx, int group, int slot_probe, int slot_gallery, int64_t row0, int64_t nrows,
                         int measure, int64_t B, int64_t *idx_out, float *util_out, uint64_t *keys_out, int out_space,
                         alink_timing *timing) {
  if (!ctx) return ALINK_E_INVALID;
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  int64_t N, base;
  ALINK_TRY(allpairs_begin(ctx, slot_probe, slot_gallery, row0, nrows, &N, &base));
  const int rc = alink_query(ctx, group, slot_probe, slot_gallery, nullptr, nullptr, N, ALINK_DEVICE, measure, B, base,
                             idx_out, util_out, keys_out, out_space, timing);
  ctx->ap_G = 0;
  ctx->ap_pair0 = 0;
  return rc;
}

// ------------------------------------------------------------------------------------------ finetune step
static void group_tensor_sizes(const HeadGroup &g, size_t each[6]) {
  each[0] = (size_t)g.D_in * g.H1; each[1] = (size_t)g.H1; each[2] = (size_t)g.H1 * g.H2;
  each[3] = (size_t)g.H2; each[4] = (size_t)g.H2 * g.C; each[5] = (size_t)g.C;
}

static int ensure_opt_state(alink_ctx *ctx, HeadGroup &g) {
  size_t each[6];
  group_tensor_sizes(g, each);
  for (int t = 0; t < 6; ++t) {
    if (g.opt_a[t]) continue;
    ALINK_CUDA(ctx, cudaMalloc((void **)&g.opt_a[t], sizeof(float) * each[t] * g.K));
    ALINK_CUDA(ctx, cudaMalloc((void **)&g.opt_d[t], sizeof(float) * each[t] * g.K));
    ALINK_CUDA(ctx, cudaMemsetAsync(g.opt_a[t], 0, sizeof(float) * each[t] * g.K, ctx->stream));
    ALINK_CUDA(ctx, cudaMemsetAsync(g.opt_d[t], 0, sizeof(float) * each[t] * g.K, ctx->stream));
  }
  return ALINK_OK;
}

// argument checks shared by the step and the epoch call
static int train_check(alink_ctx *ctx, int group, int learner, int slot_a, int slot_b, const int32_t *ia, const int32_t *ib,
                       const float *y, int64_t n) {
  if (group < 0 || group >= ALINK_MAX_GROUPS || !ctx->groups[group].set)
    return alink_fail(ctx, ALINK_E_STATE, "head group %d not set (alink_set_heads)", group);
  HeadGroup &g = ctx->groups[group];
  if (learner < 0 || learner >= g.K) return alink_fail(ctx, ALINK_E_INVALID, "learner %d outside [0, %d)", learner, g.K);
  if (slot_a < 0 || slot_a >= ALINK_MAX_SLOTS || slot_b < 0 || slot_b >= ALINK_MAX_SLOTS) return alink_fail(ctx, ALINK_E_INVALID, "bad slot");
  if (n < 1 || !y) return alink_fail(ctx, ALINK_E_INVALID, "bad batch (n=%lld)", (long long)n);
  if ((ia == nullptr) != (ib == nullptr)) return alink_fail(ctx, ALINK_E_INVALID, "ia and ib must both be given or both NULL");
  ALINK_TRY(table_wait(ctx, slot_a));
  ALINK_TRY(table_wait(ctx, slot_b));
  const Table &A = ctx->tables[slot_a], &Bt = ctx->tables[slot_b];
  if (!A.p || !Bt.p) return alink_fail(ctx, ALINK_E_STATE, "embedding table slot not set (alink_set_table)");
  if (A.dtype != ALINK_F32 || Bt.dtype != ALINK_F32 || A.D != Bt.D || (g.feat_mode ? 2 * A.D : A.D) != g.D_in)
    return alink_fail(ctx, ALINK_E_INVALID, "the finetune step needs fp32 tables matching the heads' fan-in");
  if (!ia && (n > A.M || n > Bt.M)) return alink_fail(ctx, ALINK_E_INVALID, "identity pairs exceed table rows");
  return ALINK_OK;
}

// One minibatch on the stream: forward, loss / accuracy means -> out2[0..1], and (apply) backward + Adadelta on the fp32
// masters. dia/dib (or null = identity rows from row0) and dy are DEVICE pointers to this batch's nb entries.
// The workspace (ctx->train_ws) must hold nb_max rows; successive batches reuse it in stream order.
static int train_enqueue(alink_ctx *ctx, HeadGroup &g, int learner, const Table &A, const Table &Bt, const int32_t *dia,
                         const int32_t *dib, const float *dy, int nb, double lr, double rho, double eps, int apply, float *out2,
                         float **probs_dev) {
  const size_t nX = (size_t)nb * g.D_in, n1 = (size_t)nb * g.H1, n2 = (size_t)nb * g.H2, n3 = (size_t)nb * 2;
  // workspace: X | h1 | h2 | dz1 | dz2 | dz3 | loss | acc | probs
  float *X = (float *)ctx->train_ws.p, *h1 = X + nX, *h2 = h1 + n1, *dz1 = h2 + n2, *dz2 = dz1 + n1, *dz3 = dz2 + n2;
  float *lossv = dz3 + n3, *accv = lossv + nb, *probs = accv + nb;
  if (probs_dev) *probs_dev = probs;
  size_t each[6];
  group_tensor_sizes(g, each);
  float *W1 = g.W1 + each[0] * learner, *b1 = g.b1 + each[1] * learner, *W2 = g.W2 + each[2] * learner;
  float *b2 = g.b2 + each[3] * learner, *W3 = g.W3 + each[4] * learner, *b3 = g.b3 + each[5] * learner;
  auto grid = [&](size_t n) { size_t b = (n + 255) / 256; return (unsigned)(b < 1 ? 1 : (b > 4096 ? 4096 : b)); };
  cudaStream_t st = ctx->stream;
  train_feat_kernel<<<grid(nX), 256, 0, st>>>((const float *)A.p, (const float *)Bt.p, dia, dib, A.D,
                                              g.feat_mode == ALINK_FEAT_ABSDIFF_HADAMARD, nb, X);
  ALINK_LAUNCHED(ctx);
  auto tiles = [&](int J) {
    const long long t = (long long)((J + TD_COLS - 1) / TD_COLS) * ((nb + TD_ROWS - 1) / TD_ROWS);
    return (unsigned)(t > 4096 ? 4096 : t);
  };
  train_dense_fwd_kernel<<<tiles(g.H1), 256, 0, st>>>(X, g.D_in, W1, b1, g.H1, nb, h1);
  ALINK_LAUNCHED(ctx);
  train_dense_fwd_kernel<<<tiles(g.H2), 256, 0, st>>>(h1, g.H1, W2, b2, g.H2, nb, h2);
  ALINK_LAUNCHED(ctx);
  train_head_kernel<<<grid(nb), 256, 0, st>>>(h2, W3, b3, g.H2, g.C, dy, nb, lossv, accv, dz3, probs, nb <= 256 ? out2 : nullptr);
  ALINK_LAUNCHED(ctx);
  if (nb > 256) {  // (one CTA computes the means in the head kernel's tail)
    train_mean_kernel<<<1, 32, 0, st>>>(lossv, accv, nb, out2);
    ALINK_LAUNCHED(ctx);
  }
  if (apply) {
    const float flr = (float)lr, frho = (float)rho, feps = (float)eps;
    train_grad_in_kernel<<<grid(n2), 256, 0, st>>>(dz3, g.C, W3, g.H2, h2, nb, dz2);
    ALINK_LAUNCHED(ctx);
    train_grad_in_kernel<<<grid(n1), 256, 0, st>>>(dz2, g.H2, W2, g.H1, h1, nb, dz1);
    ALINK_LAUNCHED(ctx);
    float *oa[6], *od[6];
    for (int t = 0; t < 6; ++t) {
      oa[t] = g.opt_a[t] + each[t] * learner;
      od[t] = g.opt_d[t] + each[t] * learner;
    }
    // the six gradient + Adadelta updates in one launch (train_update_all_kernel: same arithmetic per element)
    TrainUpdArgs ua;
    memset(&ua, 0, sizeof(ua));
    ua.A[0] = X; ua.dZ[0] = dz1; ua.Kd[0] = g.D_in; ua.J[0] = g.H1; ua.W[0] = W1; ua.b[0] = b1;
    ua.A[1] = h1; ua.dZ[1] = dz2; ua.Kd[1] = g.H1; ua.J[1] = g.H2; ua.W[1] = W2; ua.b[1] = b2;
    ua.A[2] = h2; ua.dZ[2] = dz3; ua.Kd[2] = g.H2; ua.J[2] = g.C; ua.W[2] = W3; ua.b[2] = b3;
    for (int l = 0; l < 3; ++l) {
      ua.ag_w[l] = oa[2 * l]; ua.ad_w[l] = od[2 * l];
      ua.ag_b[l] = oa[2 * l + 1]; ua.ad_b[l] = od[2 * l + 1];
    }
    ua.nb = nb; ua.lr = flr; ua.rho = frho; ua.eps = feps;
    train_update_all_kernel<<<grid(each[0] + each[1] + each[2] + each[3] + each[4] + each[5]), 256, 0, st>>>(ua);
    ALINK_LAUNCHED(ctx);
  }
  return ALINK_OK;
}

// the tensor-core operand images follow the fp32 masters; the half-overflow bound is recomputed by the next
// alink_set_heads (until then FP32_REFINE filters with bf16)
static int train_refresh_images(alink_ctx *ctx, HeadGroup &g) {
  cudaStream_t st = ctx->stream;
  if (g.W1t) {
    tc::prep_w1_kernel<false><<<ctx->sm_count * 4, 256, 0, st>>>(g.W1, (uint8_t *)g.W1t, g.K, g.D_in, g.H1);
    ALINK_LAUNCHED(ctx);
    tc::prep_w2_kernel<false><<<ctx->sm_count, 256, 0, st>>>(g.W2, (uint8_t *)g.W2t, g.K, g.H1, g.H2);
    ALINK_LAUNCHED(ctx);
    tc::prep_b1_kernel<false><<<ctx->sm_count, 256, 0, st>>>(g.b1, (uint8_t *)g.W2t + (size_t)g.K * g.H1 * g.H2 * 2, g.K, g.H1);
    ALINK_LAUNCHED(ctx);
  }
  g.bounds_stale = true;
  return ALINK_OK;
}

static size_t train_ws_floats(const HeadGroup &g, size_t nb) {
  return nb * g.D_in + 2 * nb * g.H1 + 2 * nb * g.H2 + 2 * nb * 2 + 2 * nb;
}

int alink_train_step(alink_ctx *ctx, int group, int learner, int slot_a, int slot_b, const int32_t *ia, const int32_t *ib,
                     const float *y, int nb, int in_space, double lr, double rho, double eps, int apply, float *loss_out,
                     float *acc_out, float *probs_out) {
  if (!ctx) return ALINK_E_INVALID;
  if (nb > 65536) return alink_fail(ctx, ALINK_E_INVALID, "bad batch (nb=%d)", nb);
  ALINK_TRY(train_check(ctx, group, learner, slot_a, slot_b, ia, ib, y, nb));
  HeadGroup &g = ctx->groups[group];
  const Table &A = ctx->tables[slot_a], &Bt = ctx->tables[slot_b];
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  ALINK_TRY(ensure_opt_state(ctx, g));
  const void *dia, *dib, *dy;
  ALINK_TRY(stage_in(ctx, ia, sizeof(int32_t) * (size_t)nb, in_space, ctx->idx_a, &dia));
  ALINK_TRY(stage_in(ctx, ib, sizeof(int32_t) * (size_t)nb, in_space, ctx->idx_b, &dib));
  ALINK_TRY(stage_in(ctx, y, sizeof(float) * (size_t)nb, in_space, ctx->train_y, &dy));
  ALINK_TRY(dev_reserve(ctx, ctx->train_ws, sizeof(float) * train_ws_floats(g, (size_t)nb)));
  ALINK_TRY(dev_reserve(ctx, ctx->train_out, 16));
  float *probs = nullptr;
  ALINK_TRY(train_enqueue(ctx, g, learner, A, Bt, (const int32_t *)dia, (const int32_t *)dib, (const float *)dy, nb, lr, rho, eps,
                          apply, (float *)ctx->train_out.p, &probs));
  if (apply) ALINK_TRY(train_refresh_images(ctx, g));
  cudaStream_t st = ctx->stream;
  float hout[2] = {0, 0};
  ALINK_CUDA(ctx, cudaMemcpyAsync(hout, ctx->train_out.p, 8, cudaMemcpyDeviceToHost, st));
  if (probs_out) ALINK_CUDA(ctx, cudaMemcpyAsync(probs_out, probs, sizeof(float) * (size_t)nb * g.C, cudaMemcpyDeviceToHost, st));
  ALINK_CUDA(ctx, cudaStreamSynchronize(st));
  if (loss_out) *loss_out = hout[0];
  if (acc_out) *acc_out = hout[1];
  return ALINK_OK;
}

// A whole pass over n pairs in minibatches of `batch` (the last one ragged), WITHOUT a host round trip per batch: the
// indices and labels cross PCIe once, the ceil(n / batch) steps are enqueued back to back — the same kernels in the same
// order as alink_train_step, so parameters, accumulators and every batch's loss are bit-identical to the step loop —
// the tensor-core images are refreshed once, and the per-batch (loss, accuracy) pairs come back in one copy.
int alink_train_epoch(alink_ctx *ctx, int group, int learner, int slot_a, int slot_b, const int32_t *ia, const int32_t *ib,
                      const float *y, int64_t n, int batch, int in_space, double lr, double rho, double eps, int apply,
                      float *loss_acc_out) {
  if (!ctx) return ALINK_E_INVALID;
  if (batch < 1 || batch > 65536) return alink_fail(ctx, ALINK_E_INVALID, "bad batch size %d", batch);
  if (n > (1ll << 30)) return alink_fail(ctx, ALINK_E_INVALID, "n too large");
  ALINK_TRY(train_check(ctx, group, learner, slot_a, slot_b, ia, ib, y, n));
  HeadGroup &g = ctx->groups[group];
  const Table &A = ctx->tables[slot_a], &Bt = ctx->tables[slot_b];
  if (!ia) return alink_fail(ctx, ALINK_E_INVALID, "alink_train_epoch needs explicit pair indices");
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  ALINK_TRY(ensure_opt_state(ctx, g));
  const void *dia, *dib, *dy;
  ALINK_TRY(stage_in(ctx, ia, sizeof(int32_t) * (size_t)n, in_space, ctx->idx_a, &dia));
  ALINK_TRY(stage_in(ctx, ib, sizeof(int32_t) * (size_t)n, in_space, ctx->idx_b, &dib));
  ALINK_TRY(stage_in(ctx, y, sizeof(float) * (size_t)n, in_space, ctx->train_y, &dy));
  const int64_t nbatch = (n + batch - 1) / batch;
  ALINK_TRY(dev_reserve(ctx, ctx->train_ws, sizeof(float) * train_ws_floats(g, (size_t)(batch < n ? batch : n))));
  ALINK_TRY(dev_reserve(ctx, ctx->train_out, sizeof(float) * 2 * (size_t)nbatch));
  float *out = (float *)ctx->train_out.p;
  for (int64_t b = 0; b < nbatch; ++b) {
    const int64_t lo = b * batch;
    const int nb = (int)(n - lo < batch ? n - lo : batch);
    ALINK_TRY(train_enqueue(ctx, g, learner, A, Bt, (const int32_t *)dia + lo, (const int32_t *)dib + lo, (const float *)dy + lo, nb,
                            lr, rho, eps, apply, out + 2 * b, nullptr));
  }
  if (apply) ALINK_TRY(train_refresh_images(ctx, g));
  if (loss_acc_out) {
    ALINK_CUDA(ctx, cudaMemcpyAsync(loss_acc_out, out, sizeof(float) * 2 * (size_t)nbatch, cudaMemcpyDeviceToHost, ctx->stream));
  }
  ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return ALINK_OK;
}

// which: 0 = parameters, 1 = Adadelta E[g^2], 2 = Adadelta E[dx^2]; six host arrays in Keras order (W1,b1,W2,b2,W3,b3)
static int train_state_xfer(alink_ctx *ctx, int group, int learner, int which, float *const *host, bool to_host) {
  if (!ctx) return ALINK_E_INVALID;
  if (group < 0 || group >= ALINK_MAX_GROUPS || !ctx->groups[group].set)
    return alink_fail(ctx, ALINK_E_STATE, "head group %d not set (alink_set_heads)", group);
  HeadGroup &g = ctx->groups[group];
  if (learner < 0 || learner >= g.K || which < 0 || which > 2 || !host) return alink_fail(ctx, ALINK_E_INVALID, "bad arguments");
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  if (which) ALINK_TRY(ensure_opt_state(ctx, g));
  size_t each[6];
  group_tensor_sizes(g, each);
  float *base[6] = {g.W1, g.b1, g.W2, g.b2, g.W3, g.b3};
  for (int t = 0; t < 6; ++t) {
    if (!host[t]) continue;
    float *dev = (which == 0 ? base[t] : (which == 1 ? g.opt_a[t] : g.opt_d[t])) + each[t] * learner;
    if (to_host) ALINK_CUDA(ctx, cudaMemcpyAsync(host[t], dev, sizeof(float) * each[t], cudaMemcpyDeviceToHost, ctx->stream));
    else ALINK_CUDA(ctx, cudaMemcpyAsync(dev, host[t], sizeof(float) * each[t], cudaMemcpyHostToDevice, ctx->stream));
  }
  ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return ALINK_OK;
}

int alink_get_train_state(alink_ctx *ctx, int group, int learner, int which, float *const *host_out) {
  return train_state_xfer(ctx, group, learner, which, host_out, true);
}

int alink_set_train_state(alink_ctx *ctx, int group, int learner, int which, const float *const *host_in) {
  if (which == 0) return ctx ? alink_fail(ctx, ALINK_E_INVALID, "parameters are set with alink_set_heads") : ALINK_E_INVALID;
  return train_state_xfer(ctx, group, learner, which, (float *const *)host_in, false);
}


// ------------------------------------------------------------------------------------------ multi-GPU exchange
int alink_comm_unique_id(void *id128) {
  NcclApi *api = nccl_api();
  if (!api || !id128) return ALINK_E_UNSUPPORTED;
  alink_nccl_uid u;
  if (api->GetUniqueId(&u) != 0) return ALINK_E_CUDA;
  memcpy(id128, &u, sizeof(u));
  return ALINK_OK;
}

// The peer windows of the push exchange (comm.cuh). Every rank allocates one, the IPC handles travel through one
// ncclAllGather, every rank maps its peers' windows; the path is used only if EVERY rank mapped every window (one
// ncclAllReduce of the success bits, which also orders the windows' zero-fill before any peer's first store).
static void push_teardown(alink_ctx *ctx) {
  for (int r = 0; r < PUSH_MAX_W; ++r) {
    if (ctx->push_win[r] && ctx->push_win[r] != ctx->push_own) cudaIpcCloseMemHandle(ctx->push_win[r]);
    ctx->push_win[r] = nullptr;
  }
  if (ctx->push_own) cudaFree(ctx->push_own);
  if (ctx->push_status) cudaFree(ctx->push_status);
  if (ctx->push_host) cudaFreeHost(ctx->push_host);
  ctx->push_own = nullptr;
  ctx->push_status = nullptr;
  ctx->push_host = nullptr;
  ctx->push_ok = false;
  ctx->push_epoch = 0;
  cudaGetLastError();
}

static int push_setup(alink_ctx *ctx, NcclApi *api) {
  const int W = ctx->comm_world, rank = ctx->comm_rank;
  ALINK_CUDA(ctx, cudaMalloc((void **)&ctx->push_own, PUSH_WIN_BYTES));
  ALINK_CUDA(ctx, cudaMalloc((void **)&ctx->push_status, 64 + 64 * PUSH_MAX_W + 16));
  ALINK_CUDA(ctx, cudaMallocHost((void **)&ctx->push_host, 64));
  ALINK_CUDA(ctx, cudaMemsetAsync(ctx->push_own, 0, PUSH_WIN_BYTES, ctx->stream));
  ALINK_CUDA(ctx, cudaMemsetAsync(ctx->push_status, 0, 64 + 64 * PUSH_MAX_W + 16, ctx->stream));
  unsigned char *hbuf = (unsigned char *)ctx->push_status + 64;       // W handles of 64 B
  unsigned int *okw = (unsigned int *)(hbuf + 64 * PUSH_MAX_W);       // this rank's success bit, then the sum
  cudaIpcMemHandle_t mine;
  bool ok = cudaIpcGetMemHandle(&mine, ctx->push_own) == cudaSuccess;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  ALINK_CUDA(ctx, cudaMemcpyAsync(hbuf + 64 * rank, &mine, 64, cudaMemcpyHostToDevice, ctx->stream));
  ALINK_NCCL(ctx, api, api->AllGather(hbuf + 64 * rank, hbuf, 8, ALINK_NCCL_UINT64, (alink_nccl_comm)ctx->comm, ctx->stream));
  cudaIpcMemHandle_t all[PUSH_MAX_W];
  ALINK_CUDA(ctx, cudaMemcpyAsync(all, hbuf, 64 * (size_t)W, cudaMemcpyDeviceToHost, ctx->stream));
  ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  for (int r = 0; r < W && ok; ++r) {
    if (r == rank) {
      ctx->push_win[r] = ctx->push_own;
      continue;
    }
    void *p = nullptr;
    if (cudaIpcOpenMemHandle(&p, all[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess || !p) {
      ok = false;
      cudaGetLastError();
    } else {
      ctx->push_win[r] = (uint8_t *)p;
    }
  }
  const unsigned int okbit = ok ? 1u : 0u;
  ALINK_CUDA(ctx, cudaMemcpyAsync(okw, &okbit, 4, cudaMemcpyHostToDevice, ctx->stream));
  ALINK_NCCL(ctx, api, api->AllReduce(okw, okw + 1, 1, ALINK_NCCL_UINT32, ALINK_NCCL_SUM, (alink_nccl_comm)ctx->comm, ctx->stream));
  unsigned int sum = 0;
  ALINK_CUDA(ctx, cudaMemcpyAsync(&sum, okw + 1, 4, cudaMemcpyDeviceToHost, ctx->stream));
  ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (sum != (unsigned int)W) {  // some rank could not map a peer (no P2P / IPC in this container): NCCL serves the exchange
    push_teardown(ctx);
    return ALINK_OK;
  }
  ctx->push_ok = true;
  return ALINK_OK;
}

int alink_comm_init(alink_ctx *ctx, const void *id128, int rank, int world) {
  if (!ctx) return ALINK_E_INVALID;
  if (!id128 || world < 1 || rank < 0 || rank >= world) return alink_fail(ctx, ALINK_E_INVALID, "bad rank %d of %d", rank, world);
  NcclApi *api = nccl_api();
  if (!api) return alink_fail(ctx, ALINK_E_UNSUPPORTED, "libnccl.so.2 could not be loaded: %s", dlerror() ? dlerror() : "missing symbols");
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  if (ctx->comm) {
    api->CommDestroy((alink_nccl_comm)ctx->comm);
    ctx->comm = nullptr;
  }
  alink_nccl_uid u;
  memcpy(&u, id128, sizeof(u));
  alink_nccl_comm c = nullptr;
  ALINK_NCCL(ctx, api, api->CommInitRank(&c, world, u, rank));
  ctx->comm = c;
  ctx->comm_rank = rank;
  ctx->comm_world = world;
  push_teardown(ctx);
  // off by default: measured on 2 B200s the NCCL all-gather + merge adds 25-35 us to the slowest rank's local query, the
  // push kernel 130-200 us (tools/sharded_probe.py); ALINK_COMM_PUSH=1 selects it
  static const int push_env = [] { const char *e = getenv("ALINK_COMM_PUSH"); return e ? atoi(e) : 0; }();
  if (world > 1 && world <= PUSH_MAX_W && push_env) ALINK_TRY(push_setup(ctx, api));
  return ALINK_OK;
}

int alink_comm_destroy(alink_ctx *ctx) {
  if (!ctx) return ALINK_E_INVALID;
  if (ctx->comm && nccl_api()) {
    cudaStreamSynchronize(ctx->stream);
    nccl_api()->CommDestroy((alink_nccl_comm)ctx->comm);
  }
  push_teardown(ctx);
  ctx->comm = nullptr;
  ctx->comm_world = 1;
  ctx->comm_rank = 0;
  return ALINK_OK;
}

// The whole sharded step behind one call: alink_query on this rank's shard (keys stay on the device, indices carry `base`),
// ncclAllGather of B keys per rank on the context's stream, the same deterministic merge on every rank, results out.
int alink_query_sharded(alink_ctx *ctx, int group, int slot_a, int slot_b, const int32_t *ia, const int32_t *ib, int64_t N,
                        int in_space, int measure, int64_t B, int64_t base, int64_t *idx_out, float *util_out, uint64_t *keys_out,
                        int out_space, alink_timing *timing) {
  if (!ctx) return ALINK_E_INVALID;
  if (!ctx->comm || ctx->comm_world == 1)
    return alink_query(ctx, group, slot_a, slot_b, ia, ib, N, in_space, measure, B < N ? B : N, base, idx_out, util_out, keys_out,
                       out_space, timing);
  if (B < 1) return alink_fail(ctx, ALINK_E_INVALID, "B must be >= 1");
  NcclApi *api = nccl_api();
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  const int W = ctx->comm_world;
  const int64_t bl = B < N ? B : N;  // a shard smaller than B contributes all its pairs; the rest of its slot stays 0
  // payload per rank: B keys + the verdict word (comm.cuh); slot W holds this rank's own
  const int64_t S = B + 1;
  ALINK_TRY(dev_reserve(ctx, ctx->comm_keys, sizeof(uint64_t) * (size_t)S * (W + 1)));
  uint64_t *mine = (uint64_t *)ctx->comm_keys.p, *all = mine + S;
  if (!ctx->comm_status) {
    ALINK_CUDA(ctx, cudaMalloc((void **)&ctx->comm_status, 64));
    ALINK_CUDA(ctx, cudaMallocHost((void **)&ctx->comm_host, 64));
  }
  static const int defer_env = [] { const char *e = getenv("ALINK_COMM_DEFER"); return e ? atoi(e) : 1; }();
  ALINK_CUDA(ctx, cudaMemsetAsync(mine, 0, sizeof(uint64_t) * (size_t)B, ctx->stream));  // key 0 sorts below every real key
  const int64_t l0 = ctx->launches;
  ctx->deferred = false;
  if (bl > 0) {
    // FP32_REFINE: enqueue the three stages WITHOUT their host synchronisation; the verdict is left on the device
    ctx->defer_req = defer_env != 0 && !(ctx->push_ok && B <= PUSH_MAX_B);
    const int rc = alink_query(ctx, group, slot_a, slot_b, ia, ib, N, in_space, measure, bl, base, nullptr, nullptr, mine, ALINK_DEVICE, timing);
    ctx->defer_req = false;
    ALINK_TRY(rc);
  }
  const bool deferred = ctx->deferred;
  ctx->deferred = false;
  ctx->push_used = 0;
  if (ctx->push_ok && B <= PUSH_MAX_B) {
    // exchange + merge in ONE launch over the peers' mapped windows (comm.cuh)
    const int64_t npad = next_pow2(B);
    ALINK_TRY(dev_reserve(ctx, ctx->sel_keys, sizeof(uint64_t) * (size_t)npad));
    PushArgs pa;
    memset(&pa, 0, sizeof(pa));
    for (int r = 0; r < W; ++r) pa.win[r] = ctx->push_win[r];
    pa.rank = ctx->comm_rank; pa.W = W; pa.B = (int)B; pa.npad = (int)(npad < 2 ? 2 : npad);
    pa.epoch = ++ctx->push_epoch;
    pa.mine = mine; pa.gathered = all; pa.out = (uint64_t *)ctx->sel_keys.p;
    pa.status = ctx->push_status; pa.ok_dev = nullptr;
    push_merge_kernel<<<1, 1024, 0, ctx->stream>>>(pa);
    ALINK_LAUNCHED(ctx);
    ctx->push_used = 1;
    ALINK_CUDA(ctx, cudaMemcpyAsync(ctx->push_host, ctx->push_status, 8, cudaMemcpyDeviceToHost, ctx->stream));
    ALINK_TRY(emit_selection(ctx, B, idx_out, util_out, keys_out, out_space));
    if (out_space == ALINK_HOST && ctx->push_host[0])  // (device outputs: the caller's next synchronising call reports it)
      return alink_fail(ctx, ALINK_E_CUDA, "peer push: a rank did not arrive within 2 s (epoch %llu)", ctx->push_epoch);
    return ALINK_OK;
  }
  // keys + verdict -> ncclAllGather -> every rank reduces the W verdicts to the same all_ok and merges
  auto exchange = [&](const unsigned int *ok_dev) -> int {
    comm_flag_kernel<<<1, 32, 0, ctx->stream>>>(mine + B, ok_dev);
    ALINK_LAUNCHED(ctx);
    ALINK_NCCL(ctx, api, api->AllGather(mine, all, (size_t)S, ALINK_NCCL_UINT64, (alink_nccl_comm)ctx->comm, ctx->stream));
    ctx->launches++;  // the collective's kernel
    comm_flags_reduce_kernel<<<1, 32, 0, ctx->stream>>>(all, W, (int)S, (int)B, ctx->comm_status);
    ALINK_LAUNCHED(ctx);
    KeySrc src{nullptr, all, 0, nullptr};
    ALINK_TRY(topk_device(ctx, src, S * W, B));
    ALINK_CUDA(ctx, cudaMemcpyAsync(ctx->comm_host, ctx->comm_status, 8, cudaMemcpyDeviceToHost, ctx->stream));
    return ALINK_OK;
  };
  ALINK_TRY(exchange(deferred ? ctx->defer_ok_dev : nullptr));
  // ONE synchronisation: results, statistics and every rank's verdict arrive together. Every rank reads all_ok, deferred or
  // not (a short shard runs exact and always reports ok) — the ranks must agree on whether the exchange is repeated.
  cudaEvent_t *e = ctx->ev;
  if (timing && deferred) ALINK_CUDA(ctx, cudaEventRecord(e[6], ctx->stream));
  ALINK_TRY(emit_selection(ctx, B, idx_out, util_out, keys_out, out_space));
  ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (deferred) refine_fill_report(ctx, ctx->rs_host, ctx->defer_eps1, ctx->defer_eps2, 1);
  if (timing && deferred) {
    cudaEventElapsedTime(&timing->ms_h2d, e[0], e[1]);
    cudaEventElapsedTime(&timing->ms_score, e[1], e[2]);
    cudaEventElapsedTime(&timing->ms_utility, e[2], e[3]);
    cudaEventElapsedTime(&timing->ms_select, e[3], e[4]);
    cudaEventElapsedTime(&timing->ms_d2h, e[4], e[6]);  // (here: verdict + exchange + merge)
    cudaEventElapsedTime(&timing->ms_total, e[0], e[6]);
    timing->launches = ctx->launches - l0;
  }
  if (ctx->comm_host[1]) return ALINK_OK;
  // Some rank's verification failed (every rank sees the same W verdicts, so every rank is here): that rank repeats its
  // query the ordinary way — widened bounds or the exact pipeline, synchronising as often as it needs — the others keep
  // their keys, and the exchange is repeated.
  if (deferred && !ctx->rs_host[RS_OK] && bl > 0) {
    ALINK_CUDA(ctx, cudaMemsetAsync(mine, 0, sizeof(uint64_t) * (size_t)B, ctx->stream));
    ALINK_TRY(alink_query(ctx, group, slot_a, slot_b, ia, ib, N, in_space, measure, bl, base, nullptr, nullptr, mine, ALINK_DEVICE, nullptr));
  }
  ALINK_TRY(exchange(nullptr));
  return emit_selection(ctx, B, idx_out, util_out, keys_out, out_space);
}

int alink_comm_push_state(const alink_ctx *ctx, int *enabled, int *used_last) {
  if (!ctx) return ALINK_E_INVALID;
  if (enabled) *enabled = ctx->push_ok ? 1 : 0;
  if (used_last) *used_last = ctx->push_used;
  return ALINK_OK;
}

// ------------------------------------------------------------------------------------------ A2-LINK pool
int alink_perturb_rows(alink_ctx *ctx, int slot_src, const int32_t *idx, int64_t N, int in_space, int side, int t, uint64_t seed,
                       int64_t pair_base, int ncoord, double vmax, int normalize, int slot_dst) {
  if (!ctx) return ALINK_E_INVALID;
  if (slot_src < 0 || slot_src >= ALINK_MAX_SLOTS || slot_dst < 0 || slot_dst >= ALINK_MAX_SLOTS || slot_src == slot_dst)
    return alink_fail(ctx, ALINK_E_INVALID, "bad slots");
  if (N < 1 || side < 0 || side > 1 || t < 0 || ncoord < 0 || pair_base < 0) return alink_fail(ctx, ALINK_E_INVALID, "bad arguments");
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  ALINK_TRY(table_wait(ctx, slot_src));
  const Table &S = ctx->tables[slot_src];
  if (!S.p || S.dtype != ALINK_F32) return alink_fail(ctx, ALINK_E_STATE, "the source slot needs an fp32 table");
  if (!idx && N > S.M) return alink_fail(ctx, ALINK_E_INVALID, "identity rows exceed the table");
  const void *didx;
  ALINK_TRY(stage_in(ctx, idx, sizeof(int32_t) * (size_t)N, in_space, side ? ctx->idx_b : ctx->idx_a, &didx));
  Table &T = ctx->tables[slot_dst];
  if (T.pending) {
    ALINK_CUDA(ctx, cudaEventSynchronize(T.ready));
    T.pending = false;
  }
  ALINK_TRY(dev_reserve(ctx, T.own, sizeof(float) * (size_t)N * S.D));
  T.p = T.own.p; T.M = N; T.D = S.D; T.dtype = ALINK_F32; T.absmax = -1.0;
  const size_t smem = sizeof(float) * 8 * (size_t)S.D;
  if (smem > 48 * 1024) ALINK_CUDA(ctx, cudaFuncSetAttribute(perturb_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t blocks = (N + 7) / 8;
  if (blocks > (int64_t)ctx->sm_count * 16) blocks = (int64_t)ctx->sm_count * 16;
  perturb_rows_kernel<<<(unsigned)blocks, 256, smem, ctx->stream>>>((const float *)S.p, (const int32_t *)didx, N, S.D, side, t, seed,
                                                                    (uint64_t)pair_base, ncoord, (float)vmax, normalize, (float *)T.own.p);
  ALINK_LAUNCHED(ctx);
  return ALINK_OK;
}

// rows [row0, row0 + nrows) of an fp32 table slot -> caller's buffer (the perturbed rows of alink_perturb_rows for a host
// driver that wants them, code/ALINK.py:219-224; everything else keeps tables on the device)
int alink_get_table_rows(alink_ctx *ctx, int slot, int64_t row0, int64_t nrows, float *out, int out_space) {
  if (!ctx) return ALINK_E_INVALID;
  if (slot < 0 || slot >= ALINK_MAX_SLOTS) return alink_fail(ctx, ALINK_E_INVALID, "bad slot");
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  ALINK_TRY(table_wait(ctx, slot));
  const Table &T = ctx->tables[slot];
  if (!T.p || T.dtype != ALINK_F32) return alink_fail(ctx, ALINK_E_STATE, "slot %d holds no fp32 table", slot);
  if (row0 < 0 || nrows < 0 || row0 + nrows > T.M || !out) return alink_fail(ctx, ALINK_E_INVALID, "rows outside the table");
  if (nrows == 0) return ALINK_OK;
  ALINK_CUDA(ctx, cudaMemcpyAsync(out, (const float *)T.p + (size_t)row0 * T.D, sizeof(float) * (size_t)nrows * T.D,
                                  out_space == ALINK_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, ctx->stream));
  if (out_space == ALINK_HOST) ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return ALINK_OK;
}

int alink_score_column(alink_ctx *ctx, int learner, int col, float *out, int out_space) {
  if (!ctx) return ALINK_E_INVALID;
  if (!ctx->probs.p || !ctx->probs_full) return alink_fail(ctx, ALINK_E_STATE, "no scores: call alink_score_* first");
  if (learner < -1 || learner >= ctx->sK || col < 0 || col > 1 || !out) return alink_fail(ctx, ALINK_E_INVALID, "bad learner / column");
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  const int64_t N = ctx->N;
  if (N == 0) return ALINK_OK;
  const float *src = (const float *)ctx->probs.p;
  int K = ctx->sK, lrn = learner;
  if (learner < 0) {  // the Bagging mean (code/committee.py:18) first, then its column
    ALINK_TRY(dev_reserve(ctx, ctx->cons, sizeof(float) * (size_t)N * ctx->sC));
    ALINK_TRY(launch_utility(ctx, src, N, ctx->sK, ctx->sC, 0, nullptr, (float *)ctx->cons.p));
    src = (const float *)ctx->cons.p;
    K = 1;
    lrn = 0;
  }
  float *dst = out;
  if (out_space == ALINK_HOST) {
    ALINK_TRY(dev_reserve(ctx, ctx->misc3, sizeof(float) * (size_t)N));
    dst = (float *)ctx->misc3.p;
  }
  int64_t blocks = (N + 255) / 256;
  if (blocks > (int64_t)ctx->sm_count * 8) blocks = (int64_t)ctx->sm_count * 8;
  extract_col_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(src, N, K, ctx->sC, lrn, col, dst);
  ALINK_LAUNCHED(ctx);
  if (out_space == ALINK_HOST) {
    ALINK_TRY(copy_out(ctx, out, dst, sizeof(float) * (size_t)N, ALINK_HOST));
    ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  return ALINK_OK;
}

int alink_min_disparity(alink_ctx *ctx, const float *p2, const float *p1, int64_t N, int T, int in_space, float *util, int out_space) {
  if (!ctx) return ALINK_E_INVALID;
  if (N < 0 || T < 1 || (N && (!p2 || !p1))) return alink_fail(ctx, ALINK_E_INVALID, "bad arguments");
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  const void *dp2, *dp1;
  ALINK_TRY(stage_in(ctx, p2, sizeof(float) * (size_t)N * T, in_space, ctx->misc0, &dp2));
  ALINK_TRY(stage_in(ctx, p1, sizeof(float) * (size_t)N, in_space, ctx->misc1, &dp1));
  ALINK_TRY(dev_reserve(ctx, ctx->util, sizeof(float) * (size_t)(N ? N : 1)));
  if (N > 0) {
    int64_t blocks = (N + 255) / 256;
    if (blocks > (int64_t)ctx->sm_count * 8) blocks = (int64_t)ctx->sm_count * 8;
    min_disparity_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>((const float *)dp2, (const float *)dp1, N, T, (float *)ctx->util.p);
    ALINK_LAUNCHED(ctx);
  }
  ctx->N = N;
  ctx->util_valid = true;
  ctx->util_measure = -1;  // not one of the probability measures: alink_select_topk ranks it as it is
  ALINK_TRY(copy_out(ctx, util, ctx->util.p, sizeof(float) * (size_t)N, out_space));
  if (util && out_space == ALINK_HOST) ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return ALINK_OK;
}

// ------------------------------------------------------------------------------------------ adversarial pool (DE)
static size_t de_al(size_t x) { return (x + 255) & ~(size_t)255; }

int alink_de_init(alink_ctx *ctx, int P, int n, int G, const double *pop, int slot_left, int slot_right, const int32_t *il,
                  const int32_t *ir, const int32_t *targets, const double *lo, const double *hi) {
  if (!ctx) return ALINK_E_INVALID;
  if (P < 1 || n < 5 || G < 2 || (G & 1) || !pop || !il || !ir || !targets || !lo || !hi)
    return alink_fail(ctx, ALINK_E_INVALID, "bad DE shape (P=%d problems, n=%d >= 5 members, G=%d genes: (index, value) pairs)", P, n, G);
  if ((int64_t)P * n > 0x7FFFFFFFll / 8) return alink_fail(ctx, ALINK_E_INVALID, "P * n too large");
  if (slot_left < 0 || slot_left >= ALINK_MAX_SLOTS || slot_right < 0 || slot_right >= ALINK_MAX_SLOTS)
    return alink_fail(ctx, ALINK_E_INVALID, "bad slot");
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  DeState &d = ctx->de;
  d.P = P; d.n = n; d.G = G; d.slot_left = slot_left; d.slot_right = slot_right; d.evaluated = false;
  const size_t npg = (size_t)P * n * G, np_ = (size_t)P * n;
  ALINK_TRY(dev_reserve(ctx, d.pop, sizeof(double) * npg));
  ALINK_TRY(dev_reserve(ctx, d.trial, sizeof(double) * npg));
  ALINK_TRY(dev_reserve(ctx, d.e_pop, sizeof(float) * np_));
  ALINK_TRY(dev_reserve(ctx, d.e_trial, sizeof(float) * np_));
  ALINK_TRY(dev_reserve(ctx, d.viol, sizeof(uint32_t) * (npg + P)));
  ALINK_TRY(dev_reserve(ctx, d.consts, de_al(12 * (size_t)P) + de_al(16 * (size_t)G) + de_al((size_t)P)));
  char *c = (char *)d.consts.p;
  ALINK_CUDA(ctx, cudaMemcpyAsync(d.pop.p, pop, sizeof(double) * npg, cudaMemcpyHostToDevice, ctx->stream));
  ALINK_CUDA(ctx, cudaMemcpyAsync(c, il, 4 * (size_t)P, cudaMemcpyHostToDevice, ctx->stream));
  ALINK_CUDA(ctx, cudaMemcpyAsync(c + 4 * (size_t)P, ir, 4 * (size_t)P, cudaMemcpyHostToDevice, ctx->stream));
  ALINK_CUDA(ctx, cudaMemcpyAsync(c + 8 * (size_t)P, targets, 4 * (size_t)P, cudaMemcpyHostToDevice, ctx->stream));
  char *c2 = c + de_al(12 * (size_t)P);
  ALINK_CUDA(ctx, cudaMemcpyAsync(c2, lo, 8 * (size_t)G, cudaMemcpyHostToDevice, ctx->stream));
  ALINK_CUDA(ctx, cudaMemcpyAsync(c2 + 8 * (size_t)G, hi, 8 * (size_t)G, cudaMemcpyHostToDevice, ctx->stream));
  ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return ALINK_OK;
}

int alink_de_mutate(alink_ctx *ctx, const double *scale, const int32_t *r0, const int32_t *r1, const uint8_t *cross, uint64_t seed,
                    uint64_t gen, double dither_lo, double dither_hi, double cr, uint32_t *nviol_out) {
  if (!ctx) return ALINK_E_INVALID;
  DeState &d = ctx->de;
  if (d.P < 1) return alink_fail(ctx, ALINK_E_STATE, "alink_de_init first");
  if ((r0 == nullptr) != (r1 == nullptr) || (r0 && !scale)) return alink_fail(ctx, ALINK_E_INVALID, "scale, r0 and r1 come together");
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t npg = (size_t)d.P * d.n * d.G, np_ = (size_t)d.P * d.n;
  const double *dscale = nullptr;
  const int32_t *dr0 = nullptr, *dr1 = nullptr;
  const uint8_t *dcross = nullptr;
  uint32_t *viol = (uint32_t *)d.viol.p, *nviol = viol + npg;
  if (r0) {  // the reference's random stream, drawn on the host
    const size_t o1 = de_al(8 * (size_t)d.P), o2 = o1 + de_al(4 * np_), o3 = o2 + de_al(4 * np_);
    ALINK_TRY(dev_reserve(ctx, d.aux, o3 + (cross ? npg : 0) + 256));
    char *a = (char *)d.aux.p;
    ALINK_CUDA(ctx, cudaMemcpyAsync(a, scale, 8 * (size_t)d.P, cudaMemcpyHostToDevice, ctx->stream));
    ALINK_CUDA(ctx, cudaMemcpyAsync(a + o1, r0, 4 * np_, cudaMemcpyHostToDevice, ctx->stream));
    ALINK_CUDA(ctx, cudaMemcpyAsync(a + o2, r1, 4 * np_, cudaMemcpyHostToDevice, ctx->stream));
    if (cross) ALINK_CUDA(ctx, cudaMemcpyAsync(a + o3, cross, npg, cudaMemcpyHostToDevice, ctx->stream));
    dscale = (const double *)a;
    dr0 = (const int32_t *)(a + o1);
    dr1 = (const int32_t *)(a + o2);
    dcross = cross ? (const uint8_t *)(a + o3) : nullptr;
    ALINK_CUDA(ctx, cudaMemsetAsync(nviol, 0, 4 * (size_t)d.P, ctx->stream));
  }
  size_t blocks = (npg + 255) / 256;
  if (blocks > (size_t)ctx->sm_count * 16) blocks = (size_t)ctx->sm_count * 16;
  de_mutate_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>((const double *)d.pop.p, d.P, d.n, d.G, dscale, dr0, dr1, dcross, seed, gen,
                                                              dither_lo, dither_hi, cr, (double *)d.trial.p, r0 ? viol : nullptr, nviol);
  ALINK_LAUNCHED(ctx);
  if (r0 && nviol_out) {
    ALINK_CUDA(ctx, cudaMemcpyAsync(nviol_out, nviol, 4 * (size_t)d.P, cudaMemcpyDeviceToHost, ctx->stream));
    ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  return ALINK_OK;
}

int alink_de_repair(alink_ctx *ctx, const double *rnd, const int64_t *off) {
  if (!ctx) return ALINK_E_INVALID;
  DeState &d = ctx->de;
  if (d.P < 1 || !off) return alink_fail(ctx, ALINK_E_STATE, "alink_de_init / alink_de_mutate first");
  const int64_t total = off[d.P];
  if (total <= 0) return ALINK_OK;
  if (!rnd) return alink_fail(ctx, ALINK_E_INVALID, "null repair values");
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t o1 = de_al(8 * (size_t)total);
  ALINK_TRY(dev_reserve(ctx, d.aux, o1 + 8 * (size_t)(d.P + 1)));
  char *a = (char *)d.aux.p;
  ALINK_CUDA(ctx, cudaMemcpyAsync(a, rnd, 8 * (size_t)total, cudaMemcpyHostToDevice, ctx->stream));
  ALINK_CUDA(ctx, cudaMemcpyAsync(a + o1, off, 8 * (size_t)(d.P + 1), cudaMemcpyHostToDevice, ctx->stream));
  de_repair_kernel<<<d.P, 256, 0, ctx->stream>>>((double *)d.trial.p, (const uint32_t *)d.viol.p, d.n, d.G, (const double *)a,
                                                 (const int64_t *)(a + o1));
  ALINK_LAUNCHED(ctx);
  ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // rnd / off are the caller's
  return ALINK_OK;
}

// which = 0: the population (first evaluation; the best member swaps into slot 0); 1: the trials; 2: the best members only
// (probs_out: host [P][K][C] per-learner probabilities of the perturbed pairs; rows_out: host [P][D] perturbed left rows)
int alink_de_evaluate(alink_ctx *ctx, int group, int which, float *probs_out, float *rows_out) {
  if (!ctx) return ALINK_E_INVALID;
  DeState &d = ctx->de;
  if (d.P < 1) return alink_fail(ctx, ALINK_E_STATE, "alink_de_init first");
  if (which < 0 || which > 2) return alink_fail(ctx, ALINK_E_INVALID, "which outside 0..2");
  if (group < 0 || group >= ALINK_MAX_GROUPS || !ctx->groups[group].set) return alink_fail(ctx, ALINK_E_STATE, "head group %d not set", group);
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  ALINK_TRY(table_wait(ctx, d.slot_left));
  ALINK_TRY(table_wait(ctx, d.slot_right));
  const Table &L = ctx->tables[d.slot_left], &R = ctx->tables[d.slot_right];
  if (!L.p || !R.p || L.dtype != ALINK_F32 || R.dtype != ALINK_F32 || L.D != R.D)
    return alink_fail(ctx, ALINK_E_STATE, "the DE objective needs fp32 tables of one width in both slots");
  const HeadGroup &g = ctx->groups[group];
  const int D = L.D, nn = which == 2 ? 1 : d.n;
  const int64_t N = (int64_t)d.P * nn;
  ALINK_TRY(dev_reserve(ctx, d.rows, sizeof(float) * (size_t)N * D));
  ALINK_TRY(dev_reserve(ctx, d.ia, sizeof(int32_t) * (size_t)N));
  ALINK_TRY(dev_reserve(ctx, d.ib, sizeof(int32_t) * (size_t)N));
  const char *c = (const char *)d.consts.p;
  const int32_t *il = (const int32_t *)c, *ir = il + d.P, *tg = ir + d.P;
  const double *lo = (const double *)(c + de_al(12 * (size_t)d.P)), *hi = lo + d.G;
  // which == 2 reads member 0 of every problem: a population of stride n * G with n = 1 member taken
  const double *genes = which == 1 ? (const double *)d.trial.p : (const double *)d.pop.p;
  int64_t blocks = (N + 7) / 8;
  if (blocks > (int64_t)ctx->sm_count * 16) blocks = (int64_t)ctx->sm_count * 16;
  if (which == 2) {
    de_perturb_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(genes, d.P, 1, d.G, (int64_t)d.G * d.n, (const float *)L.p, il, ir, D, lo, hi,
                                                                 (float *)d.rows.p, (int32_t *)d.ia.p, (int32_t *)d.ib.p);
  } else {
    de_perturb_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(genes, d.P, d.n, d.G, (int64_t)d.G, (const float *)L.p, il, ir, D, lo, hi,
                                                                 (float *)d.rows.p, (int32_t *)d.ia.p, (int32_t *)d.ib.p);
  }
  ALINK_LAUNCHED(ctx);
  Table A;
  A.p = d.rows.p; A.M = N; A.D = D; A.dtype = ALINK_F32;
  const uint32_t apG = ctx->ap_G;
  ctx->ap_G = 0;
  const int rc = score_impl(ctx, group, A, R, (const int32_t *)d.ia.p, (const int32_t *)d.ib.p, N);
  ctx->ap_G = apG;
  ALINK_TRY(rc);
  if (which == 2) {
    if (probs_out) ALINK_TRY(copy_out(ctx, probs_out, ctx->probs.p, sizeof(float) * (size_t)N * g.K * g.C, ALINK_HOST));
    if (rows_out) ALINK_TRY(copy_out(ctx, rows_out, d.rows.p, sizeof(float) * (size_t)N * D, ALINK_HOST));
    ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ALINK_OK;
  }
  float *e = which == 1 ? (float *)d.e_trial.p : (float *)d.e_pop.p;
  int64_t eb = (N + 255) / 256;
  if (eb > (int64_t)ctx->sm_count * 8) eb = (int64_t)ctx->sm_count * 8;
  de_energy_kernel<<<(unsigned)eb, 256, 0, ctx->stream>>>((const float *)ctx->probs.p, N, d.n, g.K, g.C, tg, e);
  ALINK_LAUNCHED(ctx);
  if (which == 0) {
    de_best_first_kernel<<<d.P, 256, 0, ctx->stream>>>((double *)d.pop.p, (float *)d.e_pop.p, d.n, d.G);
    ALINK_LAUNCHED(ctx);
    d.evaluated = true;
  }
  return ALINK_OK;
}

int alink_de_select(alink_ctx *ctx, const uint8_t *active) {
  if (!ctx) return ALINK_E_INVALID;
  DeState &d = ctx->de;
  if (d.P < 1 || !d.evaluated) return alink_fail(ctx, ALINK_E_STATE, "evaluate the population (which = 0) and the trials (which = 1) first");
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  uint8_t *dact = nullptr;
  if (active) {
    dact = (uint8_t *)d.consts.p + de_al(12 * (size_t)d.P) + de_al(16 * (size_t)d.G);
    ALINK_CUDA(ctx, cudaMemcpyAsync(dact, active, (size_t)d.P, cudaMemcpyHostToDevice, ctx->stream));
  }
  de_select_kernel<<<d.P, 256, 0, ctx->stream>>>((double *)d.pop.p, (float *)d.e_pop.p, (const double *)d.trial.p, (const float *)d.e_trial.p,
                                                 d.n, d.G, dact);
  ALINK_LAUNCHED(ctx);
  if (active) ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return ALINK_OK;
}

int alink_de_get(alink_ctx *ctx, double *pop_out, float *e_out) {
  if (!ctx) return ALINK_E_INVALID;
  DeState &d = ctx->de;
  if (d.P < 1) return alink_fail(ctx, ALINK_E_STATE, "alink_de_init first");
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  if (pop_out) ALINK_TRY(copy_out(ctx, pop_out, d.pop.p, sizeof(double) * (size_t)d.P * d.n * d.G, ALINK_HOST));
  if (e_out) ALINK_TRY(copy_out(ctx, e_out, d.e_pop.p, sizeof(float) * (size_t)d.P * d.n, ALINK_HOST));
  ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return ALINK_OK;
}

// ------------------------------------------------------------------------------------------ ROC counts
int alink_roc_counts(alink_ctx *ctx, const float *scores, const int32_t *mask, int64_t P, int64_t G, int64_t row0,
                     int roc_case, const double *thresholds, int nthr, int in_space, uint64_t *tp, uint64_t *fp,
                     uint64_t *n_genuine, uint64_t *n_impostor) {
  if (!ctx) return ALINK_E_INVALID;
  if (P < 0 || G < 0 || row0 < 0 || nthr < 0 || nthr > ROC_MAX_THR || roc_case < 1 || roc_case > 3 || (nthr && (!thresholds || !tp || !fp)))
    return alink_fail(ctx, ALINK_E_INVALID, "bad arguments (P=%lld G=%lld row0=%lld nthr=%d case=%d)", (long long)P,
                      (long long)G, (long long)row0, nthr, roc_case);
  if ((double)P * (double)G > 9.0e18) return alink_fail(ctx, ALINK_E_INVALID, "matrix too large");
  // sorted copy of the thresholds + where each one came from
  std::vector<int> order((size_t)nthr);
  for (int i = 0; i < nthr; ++i) {
    if (thresholds[i] != thresholds[i]) return alink_fail(ctx, ALINK_E_INVALID, "threshold %d is NaN", i);
    order[i] = i;
  }
  std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return thresholds[x] < thresholds[y]; });
  std::vector<double> sorted((size_t)nthr);
  for (int i = 0; i < nthr; ++i) sorted[i] = thresholds[order[i]];
  const int nb = nthr + 1;
  std::vector<unsigned long long> hist((size_t)2 * nb, 0ull);
  if (P > 0 && G > 0) {
    if (!scores || !mask) return alink_fail(ctx, ALINK_E_INVALID, "null scores / mask");
    ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
    const void *ds, *dm;
    ALINK_TRY(stage_in(ctx, scores, sizeof(float) * (size_t)P * G, in_space, ctx->misc0, &ds));
    ALINK_TRY(stage_in(ctx, mask, sizeof(int32_t) * (size_t)P * G, in_space, ctx->misc1, &dm));
    const size_t thr_bytes = sizeof(double) * (size_t)(nthr ? nthr : 1), hist_bytes = sizeof(unsigned long long) * 2 * nb;
    ALINK_TRY(dev_reserve(ctx, ctx->misc2, thr_bytes));
    ALINK_TRY(dev_reserve(ctx, ctx->misc3, hist_bytes));
    if (nthr) ALINK_CUDA(ctx, cudaMemcpyAsync(ctx->misc2.p, sorted.data(), sizeof(double) * nthr, cudaMemcpyHostToDevice, ctx->stream));
    ALINK_CUDA(ctx, cudaMemsetAsync(ctx->misc3.p, 0, hist_bytes, ctx->stream));
    int64_t blocks = (P * G + 256 * 16 - 1) / (256 * 16);
    const int64_t cap = (int64_t)ctx->sm_count * 8;
    if (blocks > cap) blocks = cap;
    const size_t smem = thr_bytes + sizeof(unsigned int) * 2 * nb;
    roc_hist_kernel<<<(unsigned)blocks, 256, smem, ctx->stream>>>((const float *)ds, (const int32_t *)dm, P, G, row0, roc_case,
                                                                 (const double *)ctx->misc2.p, nthr,
                                                                 (unsigned long long *)ctx->misc3.p);
    ALINK_LAUNCHED(ctx);
    ALINK_CUDA(ctx, cudaMemcpyAsync(hist.data(), ctx->misc3.p, hist_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  // bin k = #thresholds <= score; score >= sorted[t]  <=>  k >= t + 1
  for (int cls = 0; cls < 2; ++cls) {
    unsigned long long run = 0, all = hist[(size_t)cls * nb];
    uint64_t *out = cls == 0 ? tp : fp;
    for (int t = nthr - 1; t >= 0; --t) {
      run += hist[(size_t)cls * nb + t + 1];
      out[order[t]] = (uint64_t)run;
    }
    all += run;
    if (cls == 0 && n_genuine) *n_genuine = (uint64_t)all;
    if (cls == 1 && n_impostor) *n_impostor = (uint64_t)all;
  }
  return ALINK_OK;
}

// ------------------------------------------------------------------------------------------ A-LINK rule
// thr != null: sharded pool — thr[t] is the GLOBAL threshold key of noise t (0 = nothing passes), keys and
// the returned indices carry base + j; thr == null: single shard, the threshold is selected here.
static int disparity_impl(alink_ctx *ctx, const float *p2, const float *p1, int64_t N, int T, double ratio,
                          double eps, int blind, const float *y, int in_space, int64_t *idx_out, int out_space,
                          int64_t *n_out, int64_t *active_count, int64_t base, const uint64_t *thr,
                          int64_t n_total_comm = -1) {
  if (!ctx) return ALINK_E_INVALID;
  if (N < 0 || T < 1 || !n_out) return alink_fail(ctx, ALINK_E_INVALID, "bad arguments (N=%lld, T=%d)", (long long)N, T);
  if (N > 0xFFFFFFFEll || base < 0 || base + N > 0xFFFFFFFFll) return alink_fail(ctx, ALINK_E_INVALID, "N / base out of range");
  if (!(ratio >= 0.0 && ratio <= 1.0)) return alink_fail(ctx, ALINK_E_INVALID, "ratio outside [0,1]");
  *n_out = 0;
  if (active_count) *active_count = 0;
  if (N == 0) return ALINK_OK;
  if (!p2 || !p1) return alink_fail(ctx, ALINK_E_INVALID, "null scores");
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  const void *dp2, *dp1, *dy;
  ALINK_TRY(stage_in(ctx, p2, sizeof(float) * (size_t)N * T, in_space, ctx->misc0, &dp2));
  ALINK_TRY(stage_in(ctx, p1, sizeof(float) * (size_t)N, in_space, ctx->misc1, &dp1));
  ALINK_TRY(stage_in(ctx, y, sizeof(float) * (size_t)N, in_space, ctx->misc2, &dy));
  // scratch: cnt[N] int | d[N] float | blk counts | 2 x u64 counters
  const int nblk = (int)((N + 256 * DISP_ITEMS - 1) / (256 * DISP_ITEMS));
  const size_t off_d = sizeof(int) * (size_t)N;
  const size_t off_blk = off_d + sizeof(float) * (size_t)N;
  const size_t off_cnt = (off_blk + sizeof(unsigned int) * (size_t)nblk + 15) & ~(size_t)15;
  ALINK_TRY(dev_reserve(ctx, ctx->misc3, off_cnt + 16));
  ALINK_TRY(dev_reserve(ctx, ctx->sel_state, sizeof(SelectState)));
  char *scr = (char *)ctx->misc3.p;
  int *cnt = (int *)scr;
  float *d = (float *)(scr + off_d);
  unsigned int *blk = (unsigned int *)(scr + off_blk);
  unsigned long long *counters = (unsigned long long *)(scr + off_cnt);
  ALINK_CUDA(ctx, cudaMemsetAsync(cnt, 0, sizeof(int) * (size_t)N, ctx->stream));
  ALINK_CUDA(ctx, cudaMemsetAsync(counters, 0, 16, ctx->stream));
  SelectState *st = (SelectState *)ctx->sel_state.p;
  int64_t sb = (N + 2047) / 2048;
  const int64_t cap = (int64_t)ctx->sm_count * 8;
  if (sb > cap) sb = cap;
  const int64_t m = (int64_t)((double)(n_total_comm >= 0 ? n_total_comm : N) * ratio);  // int(len * ratio), code/ALINK.py:183
  NcclApi *napi = n_total_comm >= 0 ? nccl_api() : nullptr;
  static const int shifts6[6] = {53, 42, 32, 21, 10, 0};
  static const int nbits6[6] = {11, 11, 10, 11, 11, 10};
  for (int t = 0; t < T; ++t) {
    const float *q = (const float *)dp2 + (size_t)t * N;
    if (blind) {
      disparity_blind_kernel<<<(unsigned)sb, 256, 0, ctx->stream>>>(q, (const float *)dp1, N, cnt);
      ALINK_LAUNCHED(ctx);
    } else if (napi) {
      // sharded pool, everything on the device: six histogram passes over this shard's keys, each summed across the
      // ranks by ncclAllReduce on the context's stream, the bin picked by one CTA; then the marks against that key
      if (m > 0) {
        select_init_kernel<<<1, 256, 0, ctx->stream>>>(st, (unsigned long long)m);
        ALINK_LAUNCHED(ctx);
        for (int p = 0; p < 6; ++p) {
          disparity_hist_kernel<<<(unsigned)sb, 256, 0, ctx->stream>>>(q, (const float *)dp1, N, (uint64_t)base, 0ull, shifts6[p], nbits6[p],
                                                                        p == 0, st->hist, st);
          ALINK_LAUNCHED(ctx);
          ALINK_NCCL(ctx, napi, napi->AllReduce(st->hist, st->hist, 2048, ALINK_NCCL_UINT32, ALINK_NCCL_SUM, (alink_nccl_comm)ctx->comm, ctx->stream));
          ctx->launches++;
          select_pick_kernel<<<1, 256, 0, ctx->stream>>>(st->hist, nbits6[p], p == 0, st);
          ALINK_LAUNCHED(ctx);
        }
        disparity_mark_thr_kernel<<<(unsigned)sb, 256, 0, ctx->stream>>>(q, (const float *)dp1, N, (uint64_t)base, 1ull, cnt, &st->prefix);
        ALINK_LAUNCHED(ctx);
      }
    } else if (thr) {
      if (thr[t]) {
        disparity_mark_thr_kernel<<<(unsigned)sb, 256, 0, ctx->stream>>>(q, (const float *)dp1, N, (uint64_t)base,
                                                                          (unsigned long long)thr[t], cnt);
        ALINK_LAUNCHED(ctx);
      }
    } else if (m > 0) {
      absdiff_vec_kernel<<<(unsigned)sb, 256, 0, ctx->stream>>>(q, (const float *)dp1, N, d);
      ALINK_LAUNCHED(ctx);
      KeySrc src{d, nullptr, 0};
      ALINK_TRY(select_threshold(ctx, src, N, m, st));
      disparity_mark_kernel<<<(unsigned)sb, 256, 0, ctx->stream>>>(d, N, st, 1, cnt);
      ALINK_LAUNCHED(ctx);
    }
  }
  // python compares the fp32 score with float64 bounds 0.5 -/+ eps (code/ALINK.py:195): fold the
  // bounds onto the nearest fp32 values that give the same decisions.
  const double lo_d = 0.5 - eps, hi_d = 0.5 + eps;
  float lo = (float)lo_d, hi = (float)hi_d;
  if ((double)lo > lo_d) lo = nextafterf(lo, -INFINITY);
  if ((double)hi < hi_d) hi = nextafterf(hi, INFINITY);
  DispFilter f{cnt, (const float *)dp1, (const float *)dy, T, lo, hi};
  disparity_count_kernel<<<nblk, 256, 0, ctx->stream>>>(f, N, blk, counters + 0);
  ALINK_LAUNCHED(ctx);
  disparity_scan_kernel<<<1, 1024, 0, ctx->stream>>>(blk, nblk, counters + 1);
  ALINK_LAUNCHED(ctx);
  int64_t *didx = idx_out;
  if (idx_out && out_space == ALINK_HOST) {
    ALINK_TRY(dev_reserve(ctx, ctx->sel_idx, sizeof(int64_t) * (size_t)N));
    didx = (int64_t *)ctx->sel_idx.p;
  }
  if (didx) {
    disparity_scatter_kernel<<<nblk, 256, 0, ctx->stream>>>(f, N, blk, didx);
    ALINK_LAUNCHED(ctx);
    if (base) {
      offset_i64_kernel<<<ctx->sm_count, 256, 0, ctx->stream>>>(didx, N, counters + 1, base);
      ALINK_LAUNCHED(ctx);
    }
  }
  unsigned long long hc[2];
  ALINK_CUDA(ctx, cudaMemcpyAsync(hc + 1, counters + 1, 8, cudaMemcpyDeviceToHost, ctx->stream));
  if (napi)  // ACTIVE_COUNT is a count over the whole pool
    ALINK_NCCL(ctx, napi, napi->AllReduce(counters, counters, 1, ALINK_NCCL_UINT64, ALINK_NCCL_SUM, (alink_nccl_comm)ctx->comm, ctx->stream));
  ALINK_CUDA(ctx, cudaMemcpyAsync(hc, counters, 8, cudaMemcpyDeviceToHost, ctx->stream));
  ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *n_out = (int64_t)hc[1];
  if (active_count) *active_count = (int64_t)hc[0];
  if (idx_out && out_space == ALINK_HOST && hc[1] > 0) {
    ALINK_CUDA(ctx, cudaMemcpyAsync(idx_out, didx, sizeof(int64_t) * (size_t)hc[1], cudaMemcpyDeviceToHost, ctx->stream));
    ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  return ALINK_OK;
}

int alink_disparity_select(alink_ctx *ctx, const float *p2, const float *p1, int64_t N, int T, double ratio,
                           double eps, int blind, const float *y, int in_space, int64_t *idx_out, int out_space,
                           int64_t *n_out, int64_t *active_count) {
  return disparity_impl(ctx, p2, p1, N, T, ratio, eps, blind, y, in_space, idx_out, out_space, n_out, active_count, 0,
                        nullptr);
}

int alink_disparity_select_sharded(alink_ctx *ctx, const float *p2, const float *p1, int64_t N, int T, int64_t base,
                                   const uint64_t *thr_keys, double eps, int blind, const float *y, int in_space,
                                   int64_t *idx_out, int out_space, int64_t *n_out, int64_t *active_count) {
  if (!ctx) return ALINK_E_INVALID;
  if (!blind && !thr_keys) return alink_fail(ctx, ALINK_E_INVALID, "null thr_keys");
  static const uint64_t none[1] = {0};
  return disparity_impl(ctx, p2, p1, N, T, 1.0, eps, blind, y, in_space, idx_out, out_space, n_out, active_count, base,
                        blind ? none : thr_keys);
}

int alink_disparity_select_comm(alink_ctx *ctx, const float *p2, const float *p1, int64_t N, int T, int64_t base, int64_t n_total,
                                double ratio, double eps, int blind, const float *y, int in_space, int64_t *idx_out, int out_space,
                                int64_t *n_out, int64_t *active_count) {
  if (!ctx) return ALINK_E_INVALID;
  if (!ctx->comm || !nccl_api()) return alink_fail(ctx, ALINK_E_STATE, "alink_comm_init first");
  if (n_total < N) return alink_fail(ctx, ALINK_E_INVALID, "n_total (the whole pool) is smaller than this shard");
  return disparity_impl(ctx, p2, p1, N, T, ratio, eps, blind, y, in_space, idx_out, out_space, n_out, active_count, base, nullptr,
                        n_total);
}

int alink_disparity_hist(alink_ctx *ctx, const float *q, const float *p1, int64_t N, int64_t base, uint64_t prefix,
                         int pass, int in_space, uint32_t *hist_out) {
  if (!ctx) return ALINK_E_INVALID;
  if (N < 0 || base < 0 || base + N > 0xFFFFFFFFll || pass < 0 || pass > 5 || !hist_out)
    return alink_fail(ctx, ALINK_E_INVALID, "bad arguments (N=%lld, base=%lld, pass=%d)", (long long)N, (long long)base, pass);
  static const int shifts[6] = {53, 42, 32, 21, 10, 0};
  static const int nbits[6] = {11, 11, 10, 11, 11, 10};
  ALINK_CUDA(ctx, cudaSetDevice(ctx->device));
  const void *dq, *dp1;
  ALINK_TRY(stage_in(ctx, N ? q : nullptr, sizeof(float) * (size_t)N, in_space, ctx->misc0, &dq));
  ALINK_TRY(stage_in(ctx, N ? p1 : nullptr, sizeof(float) * (size_t)N, in_space, ctx->misc1, &dp1));
  ALINK_TRY(dev_reserve(ctx, ctx->sel_state, sizeof(SelectState)));
  unsigned int *hist = ((SelectState *)ctx->sel_state.p)->hist;
  ALINK_CUDA(ctx, cudaMemsetAsync(hist, 0, sizeof(unsigned int) * 2048, ctx->stream));
  if (N > 0) {
    if (!dq || !dp1) return alink_fail(ctx, ALINK_E_INVALID, "null scores");
    int64_t sb = (N + 2047) / 2048;
    const int64_t cap = (int64_t)ctx->sm_count * 8;
    if (sb > cap) sb = cap;
    disparity_hist_kernel<<<(unsigned)sb, 256, 0, ctx->stream>>>((const float *)dq, (const float *)dp1, N, (uint64_t)base,
                                                                  (unsigned long long)prefix, shifts[pass], nbits[pass],
                                                                  pass == 0, hist);
    ALINK_LAUNCHED(ctx);
  }
  ALINK_CUDA(ctx, cudaMemcpyAsync(hist_out, hist, sizeof(unsigned int) * 2048, cudaMemcpyDeviceToHost, ctx->stream));
  ALINK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return ALINK_OK;
}

}  // extern "C"
