/* alink_b200.h — C ABI of the B200-native A-LINK query-selection path (libalink_b200.so).
 *
 * The reference (iamgroot42/A-LINK, under /root/reference) is pure Python; its "FFI" for this path is the
 * set of Python call sites below. Each entry point names the reference interface it replaces
 * (paths relative to /root/reference). The Python shims in alink_b200/*.py keep the reference's
 * class / function names and bind these symbols with ctypes.
 *
 * Conventions
 *   - extern "C"; plain pointers and sizes; no exceptions or aborts cross the ABI.
 *   - every call returns ALINK_OK (0) or a negative ALINK_E_*; alink_last_error(ctx) gives the text.
 *   - `in_space` / `out_space` say where the caller's buffers live: ALINK_HOST (the library copies
 *     over PCIe on the context's stream) or ALINK_DEVICE (device pointers on the context's GPU, used
 *     as they are: no copy). The caller owns every buffer it passes.
 *   - a context owns one GPU, one stream and all scratch memory; it is not thread-safe; distinct
 *     contexts are independent. Calls are synchronous on return when any output is ALINK_HOST,
 *     and stream-ordered (use alink_sync) when all outputs are ALINK_DEVICE.
 *   - there is NO CPU fallback: without a CUDA device alink_create fails with ALINK_E_CUDA.
 *   - probabilities: per_learner is [N][K][C] fp32 (C = 2 softmax head, C = 1 sigmoid head),
 *     consensus is [N][C]; utilities are computed on rows widened to 2 columns ([1-p, p] for C = 1,
 *     code/keras_wrapper.py:263-270).
 *   - selections are ordered by (utility descending, pair index ascending): modAL's multi_argmax
 *     (code/uncertainty.py:155) with the north_star tie rule. Reported index = base + position,
 *     base + N <= 2^32 - 1.
 */
#ifndef ALINK_B200_H
#define ALINK_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ALINK_ABI_VERSION 1

enum { ALINK_OK = 0, ALINK_E_INVALID = -1, ALINK_E_CUDA = -2, ALINK_E_STATE = -3, ALINK_E_NOMEM = -4,
       ALINK_E_UNSUPPORTED = -5 };
enum { ALINK_HOST = 0, ALINK_DEVICE = 1 };
enum { ALINK_F32 = 0, ALINK_BF16 = 1 };
enum { ALINK_FEAT_ABSDIFF = 0, ALINK_FEAT_ABSDIFF_HADAMARD = 1 };
/* compute modes: FP32_EXACT = CUDA-core FFMA chains in the oracle's summation order (bit-exact with
 * oracle/); BF16_TC = tcgen05 bf16 tiles with fp32 accumulation in TMEM (|dp| <= 2e-2). */
enum { ALINK_COMPUTE_FP32_EXACT = 0, ALINK_COMPUTE_BF16_TC = 1, ALINK_COMPUTE_FP32_REFINE = 2 };
/* FP32_REFINE: alink_score_* behave as FP32_EXACT; alink_query filters with the tensor-core head (IEEE
 * half operands when the features and hidden activations provably stay below 6e4, else bf16) and
 * re-scores only the pairs within 2*eps of the B-th best approximate utility with the bit-exact head, so
 * the selection is identical to FP32_EXACT whenever |filter utility - fp32 utility| <= eps. The bound is
 * verified on every candidate of every query and widened (re-marking, re-scoring) when it fails
 * (alink_set_refine_eps, alink_refine_stats, alink_refine_detail). VOTE_ENTROPY (discrete plateaus)
 * runs exactly. */
/* activation and utility ids are shared with the oracle: see include/alink_math.h
 * (ALINK_ACT_SOFTMAX/SIGMOID, ALINK_UTIL_*). */

#define ALINK_MAX_GROUPS 4  /* head groups per context: 0 = M1 committee, 1 = M2 (code/ALINK.py:96-108) */
#define ALINK_MAX_SLOTS 8   /* embedding-table slots */

typedef struct alink_ctx alink_ctx;

/* device-side milliseconds (CUDA events on the context stream) of the last alink_query */
typedef struct alink_timing {
  float ms_h2d;      /* host->device copies of this call's inputs (0 when inputs are ALINK_DEVICE) */
  float ms_score;    /* gather + head kernels */
  float ms_utility;  /* K-learner reduction / measure */
  float ms_select;   /* top-B */
  float ms_d2h;      /* device->host copy of the results */
  float ms_total;
  int64_t launches;  /* kernels launched by this call */
} alink_timing;

int alink_abi_version(void);

/* Context on CUDA device `device`. Replaces the reference's global TF session
 * (code/ALINK.py:22-25: ConfigProto(allow_growth) + K.set_session).
 * flags: ALINK_FLAG_CHECK_INDICES — every indexed call (alink_score_indexed, alink_query) first range-checks
 * ia / ib against the row counts of the two table slots on the device (one 8 B/pair streaming pass + one
 * 4-byte read-back) and returns ALINK_E_INVALID instead of letting a caller's bad index fault the context.
 * The reference has no index arrays (it materialises rows, code/readDFW.py:222-244), hence no such failure;
 * off by default because the hot path is otherwise free of host synchronisation before the head kernel. */
#define ALINK_FLAG_CHECK_INDICES 1u
int alink_create(alink_ctx **out, int device, unsigned flags);
int alink_destroy(alink_ctx *ctx);
const char *alink_last_error(const alink_ctx *ctx);
int alink_sync(alink_ctx *ctx);
/* Launch on the caller's CUDA stream (a cudaStream_t, e.g. torch.cuda.current_stream().cuda_stream)
 * instead of the context's own; NULL restores the private stream. Lets the host framework order its
 * own work and CUDA events with the library's kernels. */
int alink_set_stream(alink_ctx *ctx, void *cuda_stream);
/* The persistent tensor-core kernels take every SM (one CTA per SM, all of its registers and shared memory), so kernels of
 * OTHER streams — the barriers of a peer-to-peer table feed, a collective — cannot start until they end. n > 0 (rounded down
 * to an even number) leaves n SMs free: the head kernels then run on sm_count - n SMs (1.4 % slower at n = 2 of 148). */
int alink_set_reserved_sms(alink_ctx *ctx, int n);
/* total kernels launched by this context so far (bench.py's gpu_launches) */
int64_t alink_launch_count(const alink_ctx *ctx);
/* Failure detection: a device-side wait that exceeds ~2 s (a protocol deadlock) records itself in host-mapped
 * memory and traps, instead of hanging the GPU. out[0] = number of waiters that timed out; for the first 63 of
 * them, out[1 + 2 i] = blockIdx << 32 | threadIdx, out[2 + 2 i] = shared-memory barrier address << 32 |
 * parity << 8 | cluster rank. Readable after the launch failure. */
int alink_watchdog_record(const alink_ctx *ctx, uint64_t out[256]);  /* words 128.. = debug progress counters (ALINK_TC_DBG_DELAY bit 6) */
/* pinned host memory for ALINK_HOST buffers that should move at full PCIe rate */
int alink_host_alloc(void **out, uint64_t bytes);
int alink_host_free(void *p);

/* Install the K heads of one group. Replaces SiameseNetwork.__init__ + maybeLoadFromMemory
 * (code/siamese.py:20-35,114-119; code/siamese3.py:15-28) for each learner of Bagging(models, ...)
 * (code/committee.py:7, code/ALINK.py:90-97). Weights are Keras Dense layout: W (in,out) row-major
 * fp32, b (out,); W1[k] is (D_in,H1), W2[k] (H1,H2), W3[k] (H2,C). D_in = D (ABSDIFF) or 2*D
 * (ABSDIFF_HADAMARD). Host pointers; copied. */
int alink_set_heads(alink_ctx *ctx, int group, int K, int D_in, int H1, int H2, int C, int out_act,
                    int feat_mode, const float *const *W1, const float *const *b1,
                    const float *const *W2, const float *const *b2, const float *const *W3,
                    const float *const *b3, int compute);

/* Install an embedding table (M rows x D) in a slot. Replaces the materialised feature arrays the
 * reference builds with Python loops (code/readDFW.py:222-244, utilities/generateMatrixDFW.py:25-33).
 * space = ALINK_HOST: copied to the GPU; ALINK_DEVICE: borrowed (caller keeps it alive). */
int alink_set_table(alink_ctx *ctx, int slot, const void *emb, int64_t M, int D, int dtype, int space);

/* Same for a HOST table, without waiting: the upload runs on the context's copy stream behind whatever the
 * compute stream has queued, and the next call that reads the slot waits for it on the device. With two slots the
 * table of the next pool crosses PCIe while the current pool is scored (the per-iteration feature arrays of
 * code/ALINK.py:139-160 are host arrays in the reference). `emb` must stay valid — and should be page-locked
 * (alink_host_alloc) — until that next call returns. */
int alink_set_table_async(alink_ctx *ctx, int slot, const void *emb, int64_t M, int D, int dtype);

/* Score N indexed pairs (row ia[n] of slot_a vs row ib[n] of slot_b) with every learner of `group`.
 * Replaces SiameseNetwork.predict (code/siamese.py:130-131) called per learner by Bagging.predict
 * (code/committee.py:13-20) / Committee.vote_proba (code/learners.py:383-416).
 * ia/ib may be NULL = identity (pair n = row n of both tables). Outputs may be NULL; the scores also
 * stay on the device inside the context for alink_utility / alink_select_topk. */
int alink_score_indexed(alink_ctx *ctx, int group, int slot_a, int slot_b, const int32_t *ia,
                        const int32_t *ib, int64_t N, int in_space, float *per_learner,
                        float *consensus, int out_space);

/* Same, on materialised [X_left, X_right] (N x D fp32 each): the reference's own calling convention
 * (code/ALINK.py:160,170; code/existing_al.py:108). Uses slots ALINK_MAX_SLOTS-2/-1 as staging. */
int alink_score_dense(alink_ctx *ctx, int group, const float *XL, const float *XR, int64_t N, int D,
                      int in_space, float *per_learner, float *consensus, int out_space);

/* Utility of every scored pair. Replaces classifier_uncertainty / classifier_margin /
 * classifier_entropy (code/uncertainty.py:63-130) and modAL's vote_entropy / consensus_entropy /
 * KL_max_disagreement (imported at code/learners.py:12). util (N) may be NULL (kept on device). */
int alink_utility(alink_ctx *ctx, int measure, float *util, int out_space);

/* Utility from caller-held probabilities ([N][K][C]); same measures. consensus (N*C) may be NULL. */
int alink_utility_from_proba(alink_ctx *ctx, const float *per_learner, int64_t N, int K, int C,
                             int measure, int in_space, float *util, float *consensus, int out_space);

/* FP32_REFINE knobs: eps = assumed bound on the filter's probability error (0 = the per-format default:
 * 1e-3 half, 6e-3 bf16; see DESIGN.md for the measured errors); stats of the last alink_query: candidates
 * re-scored exactly and the largest |approximate - exact| utility seen among them (survivors = 0 when the
 * query ran fully exact). eps bounds a probability error; the candidate margin is 2 * eps * scale(measure),
 * scale = 1 (uncertainty), 2 (margin), 1.25 (entropy, consensus entropy), 5 (KL-max).
 * alink_refine_detail: operand format of the last filter (1 = half, 0 = bf16), the bound it ended with,
 * the rms error over the candidates, and the number of mark + re-score rounds (1 unless widened). */
int alink_set_refine_eps(alink_ctx *ctx, double eps);
int alink_refine_stats(const alink_ctx *ctx, int64_t *survivors, float *max_err);
int alink_refine_detail(const alink_ctx *ctx, int *filter_f16, double *eps_used, float *rms_err, int *rounds);

/* FP32_REFINE since round 2 runs in THREE stages without host round trips (csrc/refine.cuh): (1) the half-precision
 * tensor-core filter over all N pairs, bound eps1; (2) a split-operand (hi + lo halves, three MMAs per product)
 * tensor-core re-score of the candidates, bound eps2 ~ 1e-5; (3) the bit-exact CUDA-core re-score of the few pairs within
 * 2 eps2 of the B-th value. Both bounds are verified on every query — on the candidates AND on a deterministic stratified
 * sample of >= 4096 pairs outside them — and widened (stages repeated) when they fail. eps1 defaults to the larger of the
 * per-format default and an a-priori statistical bound computed from the heads' norms and the tables' element / row-norm
 * maxima. alink_set_refine_eps2: caller's eps2 (0 = default 2e-5). alink_refine_report_get: what the last query did. */
typedef struct alink_refine_report {
  int32_t version;        /* 2 = three stages; 1 = filter + exact re-score with host-sized grids (fallback); 0 = none */
  int32_t filter_f16;     /* stage-1 operand format: 1 = IEEE half, 0 = bf16 */
  int32_t rounds;         /* 1 unless a bound had to be widened */
  int32_t fallback;       /* 0 stayed in three stages; 1 shape / format not taken; 3 / 4 / 5 stage-1 / stage-2 / both candidate lists
                             overflowed (tied plateaus); 6 bounds not verifiable in 5 rounds */
  int64_t n_stage2;       /* candidates re-scored with split operands (|C1|) */
  int64_t n_exact;        /* candidates re-scored exactly (|C2|) */
  int64_t n_sample;       /* non-candidate sample carried through stage 2 */
  int64_t n_sample_exact; /* part of it carried through stage 3 */
  double eps1_used, eps1_apriori, eps2_used;  /* eps1_apriori: the statistical model (8 sigma, typical norms) */
  double eps1_worst;      /* the same model with worst-case norms: reported, not used */
  float err1_max;         /* max |stage-1 utility - best available reference| over C1, the sample and C2 */
  float err1_rms;
  float err2_max;         /* max |stage-2 utility - exact utility| over C2 and the exact sample */
  float err2_rms;
} alink_refine_report;
int alink_set_refine_eps2(alink_ctx *ctx, double eps2);
int alink_refine_report_get(const alink_ctx *ctx, alink_refine_report *out);

/* Top-B of the context's utility vector. Replaces modAL multi_argmax (code/uncertainty.py:155,183,213).
 * idx_out (B int64), util_out (B fp32), keys_out (B packed u64 keys for cross-shard merging): any may
 * be NULL. */
int alink_select_topk(alink_ctx *ctx, int64_t B, int64_t base, int64_t *idx_out, float *util_out,
                      uint64_t *keys_out, int out_space);

/* Top-B of a caller-held utility vector (N fp32). */
int alink_topk_values(alink_ctx *ctx, const float *util, int64_t N, int64_t base, int64_t B, int in_space,
                      int64_t *idx_out, float *util_out, uint64_t *keys_out, int out_space);

/* Top-B of a utility vector WITHOUT allocating anything. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it). Same result as alink_topk_values, through the multi-launch selection (6 radix passes ->
 * compaction -> one-CTA sort) with its scratch (selection state + the B unsorted keys) in a static 48 KB device array
 * that every GPU loading the library holds until exit. Limit: B <= ALINK_NOALLOC_MAX_B (larger B is refused with
 * ALINK_E_UNSUPPORTED and a message naming the limit); N and base as alink_topk_values. util_dev and the three outputs
 * (B entries each, any may be NULL) are DEVICE pointers; the call enqueues on the context's stream and does not
 * synchronise. No cudaMalloc / cudaFree / allocating library call on this path. The scratch is shared by every call and
 * every context of the process on that GPU and keeps the last call's contents (the state is cleared at the start of
 * each call): calls that overlap on the GPU (two streams, two contexts) race - serialise them. */
#define ALINK_NOALLOC_MAX_B 4096
int alink_topk_values_noalloc(alink_ctx *ctx, const float *util_dev, int64_t N, int64_t base, int64_t B,
                              int64_t *idx_out_dev, float *util_out_dev, uint64_t *keys_out_dev);

/* Top-B of n packed keys (the merge after the per-shard all-gather). keys_out: B keys, descending. */
int alink_topk_keys(alink_ctx *ctx, const uint64_t *keys, int64_t n, int64_t B, int in_space,
                    uint64_t *keys_out, int out_space);

/* The whole step: score -> K-learner reduce -> utility -> top-B. Replaces learner.query(X, n_instances)
 * (code/base.py:179-195 -> code/uncertainty.py:133-159) and the Bagging + selection part of the
 * A-LINK loop (code/ALINK.py:160). timing may be NULL. */
int alink_query(alink_ctx *ctx, int group, int slot_a, int slot_b, const int32_t *ia, const int32_t *ib,
                int64_t N, int in_space, int measure, int64_t B, int64_t base, int64_t *idx_out,
                float *util_out, uint64_t *keys_out, int out_space, alink_timing *timing);

/* All pairs: probe rows [row0, row0 + nrows) of slot_probe against EVERY row of slot_gallery (G rows), without
 * index arrays; pair index = i * G + j (row-major, the order of the reference's loops), (row0 + nrows) * G
 * <= 2^32 - 2. Replaces the row-at-a-time scoring of utilities/generateMatrixDFW.py:27-35 (DFW score
 * matrix), code/ALINK_MTP.py:279-288 and code/siamese.py:60-79. alink_score_allpairs: consensus is the
 * [nrows * G][C] score matrix block; alink_query_allpairs: top-B of the utility over the block (sharding by
 * probe-row block across GPUs: keys carry the global pair index). */
int alink_score_allpairs(alink_ctx *ctx, int group, int slot_probe, int slot_gallery, int64_t row0, int64_t nrows,
                         float *per_learner, float *consensus, int out_space);
int alink_query_allpairs(alink_ctx *ctx, int group, int slot_probe, int slot_gallery, int64_t row0, int64_t nrows,
                         int measure, int64_t B, int64_t *idx_out, float *util_out, uint64_t *keys_out, int out_space,
                         alink_timing *timing);

/* A-LINK disparity selector on column-extracted scores (code/ALINK.py:170-201):
 * p2 (T*N), p1 (N), y (N fp32 pseudo-oracle labels, may be NULL). idx_out: capacity N int64,
 * ascending pair indices; n_out / active_count: host scalars. */
int alink_disparity_select(alink_ctx *ctx, const float *p2, const float *p1, int64_t N, int T,
                           double ratio, double eps, int blind, const float *y, int in_space,
                           int64_t *idx_out, int out_space, int64_t *n_out, int64_t *active_count);

/* The same rule on a pool sharded by pair index (one process per GPU). The per-noise prefix
 * argsort(d)[:int(N_total*ratio)] (code/ALINK.py:181-183) is a GLOBAL order statistic:
 *   alink_disparity_hist  : one radix-select pass (pass 0..5: 11/11/10 bits of |q - p1|, 11/11/10 bits of the
 *                           index part) over this shard's keys (|q[j] - p1[j]|, base + j) that match `prefix`
 *                           (ignored in pass 0); hist_out = 2048 host counters. The ranks sum the histograms
 *                           (all-reduce), pick the bin, extend the prefix; after pass 5 the prefix IS the
 *                           global threshold key of that noise type (alink_b200/dist.py).
 *   alink_disparity_select_sharded : thr_keys[t] (host, T of them; 0 = nothing passes) replace the local
 *                           selection; idx_out holds GLOBAL indices base + j, ascending. */
int alink_disparity_hist(alink_ctx *ctx, const float *q, const float *p1, int64_t N, int64_t base, uint64_t prefix,
                         int pass, int in_space, uint32_t *hist_out);
int alink_disparity_select_sharded(alink_ctx *ctx, const float *p2, const float *p1, int64_t N, int T, int64_t base,
                                   const uint64_t *thr_keys, double eps, int blind, const float *y, int in_space,
                                   int64_t *idx_out, int out_space, int64_t *n_out, int64_t *active_count);

/* One finetune ("teach") minibatch step of learner `learner` of a head group on the pairs (ia[n], ib[n]), labels
 * y[n] in {0, 1}. Replaces what Keras runs per batch in code/siamese.py:52-58 (finetune ->
 * fit(batch_size=16), called from code/ALINK.py:250) and :86-112 (train_on_batch / test_on_batch): forward,
 * binary cross-entropy + binary accuracy, backward, Adadelta(lr, rho, eps; Keras 2.1.2 defaults rho = 0.95, eps = 1e-8) on the group's fp32
 * master weights, which stay on the device (apply = 0: evaluate only, like test_on_batch). loss_out / acc_out:
 * host scalars (batch means, before the update); probs_out: host [nb][C] or NULL. ia, ib, y follow in_space.
 * alink_get_train_state / alink_set_train_state move one learner's six tensors (Keras order W1,b1,W2,b2,W3,b3;
 * NULL entries are skipped) between host and device: which = 0 parameters (get only), 1 = Adadelta E[g^2],
 * 2 = Adadelta E[dx^2]. Accumulators start at zero when a group is (re)set. */
int alink_train_step(alink_ctx *ctx, int group, int learner, int slot_a, int slot_b, const int32_t *ia,
                     const int32_t *ib, const float *y, int nb, int in_space, double lr, double rho, double eps,
                     int apply, float *loss_out, float *acc_out, float *probs_out);
/* A whole pass over n pairs in minibatches of `batch` (the last one ragged) without a host round trip per batch: what
 * Keras' fit loop does per epoch (code/siamese.py:52-58 -> fit(batch_size=16); apply = 0: the validation pass). Same
 * kernels in the same order as alink_train_step, so parameters, accumulators and every batch's loss are bit-identical
 * to the step loop. ia, ib, y: n entries in the epoch's (shuffled) order, in_space as usual; loss_acc_out: host
 * [ceil(n / batch)][2] = each batch's (mean loss, mean accuracy), or NULL. */
int alink_train_epoch(alink_ctx *ctx, int group, int learner, int slot_a, int slot_b, const int32_t *ia,
                      const int32_t *ib, const float *y, int64_t n, int batch, int in_space, double lr, double rho,
                      double eps, int apply, float *loss_acc_out);
int alink_get_train_state(alink_ctx *ctx, int group, int learner, int which, float *const *host_out);
int alink_set_train_state(alink_ctx *ctx, int group, int learner, int which, const float *const *host_in);

/* Verification-curve counts on a block of the all-pairs score matrix. Replaces the loops of
 * utilities/ROC_precompute.py:19-63: over the strict upper triangle (j > row0 + r) of scores[P][G], mask codes
 * pick the genuine / impostor pairs of a protocol (roc_case 1: code 1 vs 3; 2: 2 vs 4; 3: {1,2} vs {3,4});
 * tp[t] / fp[t] = number of genuine / impostor scores >= thresholds[t] (compared in double, like the reference's
 * float64 matrix); n_genuine / n_impostor = the denominators of TPR / FPR. thresholds, tp, fp and the scalars
 * are host arrays (nthr <= 2047, any order); scores / mask follow in_space. Blocks of probe rows add up. */
int alink_roc_counts(alink_ctx *ctx, const float *scores, const int32_t *mask, int64_t P, int64_t G, int64_t row0,
                     int roc_case, const double *thresholds, int nthr, int in_space, uint64_t *tp, uint64_t *fp,
                     uint64_t *n_genuine, uint64_t *n_impostor);

/* Multi-GPU: the cross-shard exchange behind the C ABI. One process per GPU; pairs shard by index;
 * the only message of the query path is B packed keys per rank. NCCL is bound at run time (dlopen of libnccl.so.2 by
 * soname: a process that already loaded NCCL, e.g. through PyTorch, shares that copy), so a host needs nothing but this
 * header: rank 0 calls alink_comm_unique_id and ships the 128 bytes to the other ranks by any means (file, socket, MPI),
 * every rank calls alink_comm_init on its context, then alink_query_sharded replaces alink_query: the shard's top-B
 * (indices carry `base`), ncclAllGather over NVLink on the context's stream, the same deterministic merge on every rank —
 * the result equals the single-GPU selection bit for bit, ties included. A zero-padded key sorts below every real key, so
 * shards smaller than B are fine. Without alink_comm_init (or world = 1) alink_query_sharded is alink_query. */
int alink_comm_unique_id(void *id128);
int alink_comm_init(alink_ctx *ctx, const void *id128, int rank, int world);
int alink_comm_destroy(alink_ctx *ctx);
int alink_query_sharded(alink_ctx *ctx, int group, int slot_a, int slot_b, const int32_t *ia, const int32_t *ib, int64_t N,
                        int in_space, int measure, int64_t B, int64_t base, int64_t *idx_out, float *util_out, uint64_t *keys_out,
                        int out_space, alink_timing *timing);

/* The A-LINK rule on a pool sharded over the communicator's ranks, thresholds selected ON THE DEVICE: per noise type six
 * histogram passes over this shard's keys, each summed across the ranks by ncclAllReduce on the context's stream, the bin
 * picked by one CTA (no host round trip per pass: round 1 did 6 x T of them from Python), then marks, intersection, grey zone
 * and oracle agreement locally. p2 (T x N), p1, y: this rank's pairs [base, base + N) of a pool of n_total pairs. idx_out:
 * this rank's selected pairs as GLOBAL indices, ascending (capacity N); n_out: their count; active_count: ACTIVE_COUNT of the
 * WHOLE pool (all-reduced). Every rank must call it (collective). */
int alink_disparity_select_comm(alink_ctx *ctx, const float *p2, const float *p1, int64_t N, int T, int64_t base, int64_t n_total,
                                double ratio, double eps, int blind, const float *y, int in_space, int64_t *idx_out, int out_space,
                                int64_t *n_out, int64_t *active_count);

/* Whether alink_query_sharded exchanges the keys through the peer-push kernel (every rank mapped every peer's window
 * through CUDA IPC at alink_comm_init; only with ALINK_COMM_PUSH=1 - measured slower than the NCCL all-gather, which is the default) and whether
 * the last call did. SURVEY.md 5 ("one-shot peer-mapped push"); the reference is single-process. */
int alink_comm_push_state(const alink_ctx *ctx, int *enabled, int *used_last);

/* The A2-LINK noise-augmented pool generated on the device (BASELINE configs[3]). Replaces the
 * per-noise re-featurisation of code/ALINK.py:163-170 (bag.attackModel -> conversionModel.process) with its embedding-space
 * analogue: row n of slot_dst = normalize(e + delta), e = row idx[n] (or n) of slot_src, delta = `ncoord` coordinates set to
 * U(-vmax, vmax) (40 in code/attack.py:91), from a counter-based stream keyed on (seed, pair_base + n, side, t, draw), so the
 * host oracle regenerates the same rows bit for bit and nothing crosses PCIe. slot_dst becomes a context-owned fp32 table
 * of N rows (pair n = row n: score it with ia = ib = NULL).
 *   alink_score_column   column `col` (0 | 1; the C = 1 head is widened) of learner `learner` (-1 = the Bagging mean) of the last scores -> out[N]:
 *                        the p2[t] row / the p1 vector of the rule (ALINK.py:175-176), without a host hop.
 *   alink_min_disparity  util[j] = min_t |p2[t][j] - p1[j]| (kept in the context: alink_select_topk then ranks it). */
int alink_perturb_rows(alink_ctx *ctx, int slot_src, const int32_t *idx, int64_t N, int in_space, int side, int t, uint64_t seed,
                       int64_t pair_base, int ncoord, double vmax, int normalize, int slot_dst);
int alink_get_table_rows(alink_ctx *ctx, int slot, int64_t row0, int64_t nrows, float *out, int out_space);  /* fp32 rows of a slot (e.g. the perturbed rows, for a host driver: code/ALINK.py:219-224) */
int alink_score_column(alink_ctx *ctx, int learner, int col, float *out, int out_space);
int alink_min_disparity(alink_ctx *ctx, const float *p2, const float *p1, int64_t N, int T, int in_space, float *util, int out_space);

/* Embedding-space adversarial pool generator: batched differential evolution with the siamese head as the objective
 * (SURVEY.md 8f.3). Replaces the solver loop of code/differential_evolution.py:540-760 (SciPy fork, whole population scored
 * in one call :630-645, :692-714) as code/attack.py:59-103 drives it (best1bin, dither [0.5, 1), recombination 1, no
 * polish). P problems (one per pair) x n members x G genes in [0, 1]; genes come in (coordinate index, value) pairs scaled
 * to [lo, hi] like _scale_parameters (:763-767); a member assigns left_row[int(index)] = (float) value for each pair
 * (perturb_image, attack.py:5-29, in embedding space) and its energy is 1 - P(target class) of the head group's mean.
 *   alink_de_init      population (host, [P][n][G] float64), table slots, per-problem left / right rows and target class,
 *                      per-gene bounds.
 *   alink_de_mutate    trial = where(cross, pop[0] + scale (pop[r0] - pop[r1]), pop[c]) in float64. r0 != NULL: the caller
 *                      supplies the reference's random stream (scale [P], r0 / r1 [P][n], cross [P][n][G] or NULL = all
 *                      true); genes outside [0, 1] are counted per problem (nviol_out, host [P]) and wait for
 *                      alink_de_repair. r0 == NULL: a counter-based device stream keyed on (seed, problem, gen, member, gene)
 *                      draws everything, repairs in place, and nothing returns to the host.
 *   alink_de_repair    _ensure_constraint (:775-780): the k-th flagged gene of problem p, in (member, gene) order, takes
 *                      rnd[off[p] + k] (host arrays; off has P + 1 entries).
 *   alink_de_evaluate  which = 0 population (first call: the best member swaps into slot 0, :660-668), 1 trials, 2 the best
 *                      member of every problem only -> probs_out (host [P][K][C]) and rows_out (host [P][D], the
 *                      perturbed left rows: the generator's product). Scores with `group` in its compute mode.
 *   alink_de_select    :704-714 for every problem; active (host [P] bytes, may be NULL) freezes problems whose callback
 *                      asked to stop (attack.py:83-84).
 *   alink_de_get       population / energies back to the host (either may be NULL). */
int alink_de_init(alink_ctx *ctx, int P, int n, int G, const double *pop, int slot_left, int slot_right, const int32_t *il,
                  const int32_t *ir, const int32_t *targets, const double *lo, const double *hi);
int alink_de_mutate(alink_ctx *ctx, const double *scale, const int32_t *r0, const int32_t *r1, const uint8_t *cross, uint64_t seed,
                    uint64_t gen, double dither_lo, double dither_hi, double cr, uint32_t *nviol_out);
int alink_de_repair(alink_ctx *ctx, const double *rnd, const int64_t *off);
int alink_de_evaluate(alink_ctx *ctx, int group, int which, float *probs_out, float *rows_out);
int alink_de_select(alink_ctx *ctx, const uint8_t *active);
int alink_de_get(alink_ctx *ctx, double *pop_out, float *e_out);

#ifdef __cplusplus
}
#endif
#endif /* ALINK_B200_H */
