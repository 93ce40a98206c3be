/* alink_oracle.c — CPU restatement of A-LINK's query-selection path. TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library, and only as the checker or the timed CPU baseline. The product
 * (alink_b200/) never links, imports or executes anything under oracle/.
 *
 * PARITY STATUS: pinned to the reference's OWN code, executed in the build container, for every row except the Keras
 * graph. tests/golden/make_ref_fixtures.py imports the unmodified /root/reference modules (uncertainty.py, committee.py,
 * learners.py + base.py, keras_wrapper.py, readDFW.py, readMTP.py) under a stub of the three packages that cannot be
 * installed here (modAL, keras, noise), exec's the text of ALINK.py:171-201, and writes tests/golden/ref_*.npz;
 * tests/golden/make_ref_de_fixture.py does the same with differential_evolution.py. tests/test_ref_pinned.py checks this
 * oracle against those fixtures (bit-exact; scipy's entropy to 2e-7), tests/test_gpu_ref_pinned.py the CUDA path.
 * STILL "parity unpinned": the Dense / relu / softmax / sigmoid graph itself (siamese.py:24-33, siamese3.py:18-26), which
 * exists only inside Keras 2.1.2 / TensorFlow 1.15 (not installable: no network). For it this restatement is pinned
 * against an independent float64 numpy/scipy transcription (oracle/ref_numpy.py; tests/test_oracle.py) and the golden
 * vectors generated from it (tests/golden/{softmax,sigmoid}_*.npz).
 *
 * Reference lines restated (paths relative to /root/reference):
 *   pair feature   x = |l - r| ............................ code/siamese.py:27-28, code/siamese3.py:21-22
 *   head           relu(xW1+b1) -> relu(.W2+b2) -> .W3+b3 .. code/siamese.py:29-31, code/siamese3.py:23-25
 *   activation     softmax(2) / sigmoid(1) ................ code/siamese.py:32, code/siamese3.py:25
 *   predict        batched forward, rows independent ...... code/siamese.py:130-131
 *   Bagging mean   sum_k P_k / K .......................... code/committee.py:13-20
 *   widen C=1      hstack([1-p, p]) ....................... code/keras_wrapper.py:263-270
 *   vote / proba   per-learner labels / probabilities ..... code/learners.py:365-416, keras_wrapper.py:235-242
 *   consensus      mean over learners ..................... code/learners.py:334-345
 *   measures       1-max, margin, entropy ................. code/uncertainty.py:26,43-44,60,82,107-108,130
 *   selection      top-n of the utility ................... code/uncertainty.py:154-159 (modAL multi_argmax)
 *   disparity rule per-noise argsort[:ratio*N], intersect .. code/ALINK.py:170-189
 *   grey zone + oracle agreement ......................... code/ALINK.py:192-201
 *
 * Arithmetic is DEFINED here (and in include/alink_math.h) as: every dense output neuron is one
 * sequential chain acc = fmaf(x[k], W[k][j], acc), k = 0..fan_in-1, acc0 = 0, then acc + b[j], then
 * max(.,0) for the hidden layers. The loops below vectorise ACROSS output neurons j and pairs, never
 * across k, so the chain order is the same as the CUDA fp32 kernel's. Build: -ffp-contract=off.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>
#include "../include/alink_math.h"

#define PB 4  /* pairs per register block */
#define JB 16 /* output neurons per register block */

static int g_threads = 0;
void alink_oracle_set_threads(int n) { g_threads = n; }
int alink_oracle_get_threads(void) {
  if (g_threads > 0) return g_threads;
  long n = sysconf(_SC_NPROCESSORS_ONLN);
  return n > 0 ? (n > 256 ? 256 : (int)n) : 1;
}

/* Columns [j0, Dout) of the layer as plain scalar chains (k ascending, one fmaf per term). */
static void dense_scalar(const float *X, int ldx, int P, int Din, const float *W, const float *b, int Dout,
                         int relu, float *Y, int ldy, int j0) {
  for (int p = 0; p < P; ++p)
    for (int j = j0; j < Dout; ++j) {
      float acc = 0.0f;
      for (int k = 0; k < Din; ++k) acc = fmaf(X[(size_t)p * ldx + k], W[(size_t)k * Dout + j], acc);
      float v = acc + b[j];
      if (relu && !(v > 0.0f)) v = 0.0f;
      Y[(size_t)p * ldy + j] = v;
    }
}

/* Y[p][j] = act(sum_k X[p][k] W[k][j] + b[j]) for p < P (P <= PB). Row strides ldx / ldy. */
static void dense_block(const float *X, int ldx, int P, int Din, const float *W, const float *b,
                        int Dout, int relu, float *Y, int ldy) {
  int j0 = 0;
  if (P == PB) {
    for (; j0 + JB <= Dout; j0 += JB) {
      float acc[PB][JB];
      for (int p = 0; p < PB; ++p)
        for (int j = 0; j < JB; ++j) acc[p][j] = 0.0f;
      for (int k = 0; k < Din; ++k) {
        const float *w = W + (size_t)k * Dout + j0;
#pragma GCC unroll 4
        for (int p = 0; p < PB; ++p) {
          const float xk = X[(size_t)p * ldx + k];
#pragma GCC unroll 16
          for (int j = 0; j < JB; ++j) acc[p][j] = fmaf(xk, w[j], acc[p][j]);
        }
      }
      for (int p = 0; p < PB; ++p)
        for (int j = 0; j < JB; ++j) {
          float v = acc[p][j] + b[j0 + j];
          if (relu && !(v > 0.0f)) v = 0.0f;
          Y[(size_t)p * ldy + j0 + j] = v;
        }
    }
  }
  /* remainder columns / short blocks: plain scalar chains (same order) */
  dense_scalar(X, ldx, P, Din, W, b, Dout, relu, Y, ldy, j0);
}

/* Score N pairs with K learners.
 *   ta/tb  : embedding tables, row-major fp32, D columns. ia/ib: row indices per pair, or NULL for
 *            the identity (pair i = rows i of ta and tb: the dense [X_left, X_right] form).
 *   feat_mode 0: x = |a-b| (fan-in D, the reference). 1: x = [|a-b| ; a*b] (fan-in 2D, north_star opt-in).
 *   W1[k] (Din,H1), W2[k] (H1,H2), W3[k] (H2,C): Keras (in,out) row-major; b*: (out,).
 *   per_learner: N*K*C (layout [n][k][c]) or NULL. consensus: N*C Bagging mean or NULL.
 *   util: N utilities of `measure` (computed on rows widened to 2 columns) or NULL.
 * Returns 0, or -1 on bad arguments. */
typedef struct {
  const float *ta, *tb;
  const int32_t *ia, *ib;
  int64_t N;
  int D, feat_mode, K, H1, H2, C, act, measure;
  const float *const *W1, *const *b1, *const *W2, *const *b2, *const *W3, *const *b3;
  float *per_learner, *consensus, *util;
  int64_t blk_begin, blk_end;
} score_job;

static void *score_worker(void *arg) {
  const score_job *J = (const score_job *)arg;
  const int D = J->D, K = J->K, H1 = J->H1, H2 = J->H2, C = J->C;
  const int Din = J->feat_mode ? 2 * D : D;
  float *x = (float *)malloc(sizeof(float) * (size_t)PB * Din);
  float *h1 = (float *)malloc(sizeof(float) * (size_t)PB * H1);
  float *h2 = (float *)malloc(sizeof(float) * (size_t)PB * H2);
  float z[PB * 2];
  float rows[PB][ALINK_MAX_K * 2]; /* widened 2-column rows per pair */
  for (int64_t blk = J->blk_begin; blk < J->blk_end; ++blk) {
    const int64_t n0 = blk * PB;
    const int P = (int)((J->N - n0) < PB ? (J->N - n0) : PB);
    for (int p = 0; p < P; ++p) {
      const int64_t n = n0 + p;
      const float *a = J->ta + (size_t)(J->ia ? J->ia[n] : n) * D;
      const float *bb = J->tb + (size_t)(J->ib ? J->ib[n] : n) * D;
      float *xp = x + (size_t)p * Din;
      for (int d = 0; d < D; ++d) {
        float df = a[d] - bb[d];
        xp[d] = df < 0.0f ? -df : df; /* K.abs(l - r), code/siamese.py:27 */
      }
      if (J->feat_mode)
        for (int d = 0; d < D; ++d) xp[D + d] = a[d] * bb[d];
    }
    for (int k = 0; k < K; ++k) {
      dense_block(x, Din, P, Din, J->W1[k], J->b1[k], H1, 1, h1, H1);
      dense_block(h1, H1, P, H1, J->W2[k], J->b2[k], H2, 1, h2, H2);
      dense_scalar(h2, H2, P, H2, J->W3[k], J->b3[k], C, 0, z, 2, 0); /* C <= 2 columns: never blocked */
      for (int p = 0; p < P; ++p) {
        float p0, p1;
        if (J->act == ALINK_ACT_SOFTMAX) {
          alink_softmax2(z[p * 2 + 0], z[p * 2 + 1], &p0, &p1);
          if (J->per_learner) {
            J->per_learner[((size_t)(n0 + p) * K + k) * 2 + 0] = p0;
            J->per_learner[((size_t)(n0 + p) * K + k) * 2 + 1] = p1;
          }
        } else {
          p1 = alink_sigmoid(z[p * 2 + 0]);
          p0 = 1.0f - p1; /* keras_wrapper.py:268 */
          if (J->per_learner) J->per_learner[(size_t)(n0 + p) * K + k] = p1;
        }
        rows[p][2 * k + 0] = p0;
        rows[p][2 * k + 1] = p1;
      }
    }
    for (int p = 0; p < P; ++p) {
      float c0, c1;
      alink_consensus2(rows[p], K, &c0, &c1);
      if (J->consensus) {
        if (C == 2) {
          J->consensus[(size_t)(n0 + p) * 2 + 0] = c0;
          J->consensus[(size_t)(n0 + p) * 2 + 1] = c1;
        } else {
          J->consensus[(size_t)(n0 + p)] = c1;
        }
      }
      if (J->util) J->util[n0 + p] = alink_utility2(J->measure, rows[p], K, c0, c1);
    }
  }
  free(x);
  free(h1);
  free(h2);
  return NULL;
}

int alink_oracle_score(const float *ta, const float *tb, const int32_t *ia, const int32_t *ib,
                       int64_t N, int D, int feat_mode, int K, int H1, int H2, int C, int act,
                       const float *const *W1, const float *const *b1, const float *const *W2,
                       const float *const *b2, const float *const *W3, const float *const *b3,
                       float *per_learner, float *consensus, int measure, float *util) {
  if (K < 1 || K > ALINK_MAX_K || (C != 1 && C != 2) || D < 1 || H1 < 1 || H2 < 1) return -1;
  if ((act == ALINK_ACT_SOFTMAX && C != 2) || (act == ALINK_ACT_SIGMOID && C != 1)) return -1;
  const int64_t nblk = (N + PB - 1) / PB;
  int nt = alink_oracle_get_threads();
  if (nt > nblk) nt = nblk > 0 ? (int)nblk : 1;
  score_job jobs[256];
  pthread_t tids[256];
  for (int t = 0; t < nt; ++t) {
    score_job j = {ta, tb, ia, ib, N, D, feat_mode, K, H1, H2, C, act, measure,
                   W1, b1, W2, b2, W3, b3, per_learner, consensus, util,
                   nblk * t / nt, nblk * (t + 1) / nt};
    jobs[t] = j;
  }
  for (int t = 1; t < nt; ++t) pthread_create(&tids[t], NULL, score_worker, &jobs[t]);
  score_worker(&jobs[0]);
  for (int t = 1; t < nt; ++t) pthread_join(tids[t], NULL);
  return 0;
}

/* Utility of N rows from already computed per-learner probabilities (layout [n][k][c]). */
int alink_oracle_utility(const float *per_learner, int64_t N, int K, int C, int measure, float *util,
                         float *consensus) {
  if (K < 1 || K > ALINK_MAX_K || (C != 1 && C != 2)) return -1;
  for (int64_t n = 0; n < N; ++n) {
    float rows[ALINK_MAX_K * 2];
    for (int k = 0; k < K; ++k) {
      if (C == 2) {
        rows[2 * k] = per_learner[((size_t)n * K + k) * 2];
        rows[2 * k + 1] = per_learner[((size_t)n * K + k) * 2 + 1];
      } else {
        float p = per_learner[(size_t)n * K + k];
        rows[2 * k] = 1.0f - p;
        rows[2 * k + 1] = p;
      }
    }
    float c0, c1;
    alink_consensus2(rows, K, &c0, &c1);
    if (consensus) {
      if (C == 2) {
        consensus[2 * n] = c0;
        consensus[2 * n + 1] = c1;
      } else
        consensus[n] = c1;
    }
    if (util) util[n] = alink_utility2(measure, rows, K, c0, c1);
  }
  return 0;
}

static int cmp_key_desc(const void *a, const void *b) {
  uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
  return x < y ? 1 : (x > y ? -1 : 0);
}

/* Top-B of util[0..N) by (utility desc, index asc); index reported = base + position.
 * idx_out/util_out/keys_out: B entries each (any may be NULL). */
int alink_oracle_topk(const float *util, int64_t N, int64_t base, int64_t B, int64_t *idx_out,
                      float *util_out, uint64_t *keys_out) {
  if (B < 0 || B > N || base + N > 0xFFFFFFFFll) return -1;
  uint64_t *keys = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)(N ? N : 1));
  for (int64_t i = 0; i < N; ++i) keys[i] = alink_pack_key(util[i], (uint64_t)(base + i));
  qsort(keys, (size_t)N, sizeof(uint64_t), cmp_key_desc);
  for (int64_t i = 0; i < B; ++i) {
    if (idx_out) idx_out[i] = (int64_t)alink_key_index(keys[i]);
    if (util_out) util_out[i] = alink_key_utility(keys[i]);
    if (keys_out) keys_out[i] = keys[i];
  }
  free(keys);
  return 0;
}

/* Merge per-shard candidate keys: top-B of n keys, sorted descending. */
int alink_oracle_topk_keys(const uint64_t *keys_in, int64_t n, int64_t B, uint64_t *keys_out) {
  if (B < 0 || B > n) return -1;
  uint64_t *keys = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)(n ? n : 1));
  memcpy(keys, keys_in, sizeof(uint64_t) * (size_t)n);
  qsort(keys, (size_t)n, sizeof(uint64_t), cmp_key_desc);
  memcpy(keys_out, keys, sizeof(uint64_t) * (size_t)B);
  free(keys);
  return 0;
}

/* A-LINK disparity selector (code/ALINK.py:170-201).
 *   p2: T*N  — column `col` of the disguised-faces model on each of T noisy copies of the pool.
 *   p1: N    — column `col` of the committee mean on the clean pool.
 *   per noise t: blind==0: keep the int(N*ratio) pairs with the LARGEST |p2-p1| (argsort of
 *   -|.| ascending, ALINK.py:181-183; ties -> lowest index); blind!=0: keep pairs whose 0.5
 *   decisions differ (ALINK.py:177-179). Intersect over t (ALINK.py:186-189). Then keep j only if
 *   p1[j] <= .5-eps or >= .5+eps (each such j is one oracle query, ALINK.py:195-198) and, when y is
 *   given, (p1[j] >= .5) == (y[j] >= .5) (ALINK.py:196-200).
 * idx_out: ascending pair indices (capacity N). n_out, active_count: scalars. */
int alink_oracle_disparity_select(const float *p2, const float *p1, int64_t N, int T, double ratio,
                                  double eps, int blind, const float *y, int64_t *idx_out,
                                  int64_t *n_out, int64_t *active_count) {
  if (T < 1 || N < 0) return -1;
  int32_t *cnt = (int32_t *)calloc((size_t)(N ? N : 1), sizeof(int32_t));
  uint64_t *keys = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)(N ? N : 1));
  const int64_t m = (int64_t)((double)N * ratio); /* int(len * ratio): python float64 */
  for (int t = 0; t < T; ++t) {
    const float *q = p2 + (size_t)t * N;
    if (blind) {
      for (int64_t j = 0; j < N; ++j) cnt[j] += ((q[j] >= 0.5f) != (p1[j] >= 0.5f)) ? 1 : 0;
    } else {
      for (int64_t j = 0; j < N; ++j) {
        float d = q[j] - p1[j];
        d = d < 0.0f ? -d : d;
        keys[j] = alink_pack_key(d, (uint64_t)j);
      }
      qsort(keys, (size_t)N, sizeof(uint64_t), cmp_key_desc);
      for (int64_t i = 0; i < m; ++i) cnt[alink_key_index(keys[i])] += 1;
    }
  }
  int64_t n = 0, active = 0;
  const double lo = 0.5 - eps, hi = 0.5 + eps; /* python compares the f32 score with f64 bounds */
  for (int64_t j = 0; j < N; ++j) {
    if (cnt[j] != T) continue;
    if ((double)p1[j] <= lo || (double)p1[j] >= hi) {
      ++active;
      if (!y || ((p1[j] >= 0.5f) == (y[j] >= 0.5f))) idx_out[n++] = j;
    }
  }
  *n_out = n;
  if (active_count) *active_count = active;
  free(cnt);
  free(keys);
  return 0;
}

/* scalar probes so tests can pin include/alink_math.h against libm / scipy */
float alink_oracle_expf(float x) { return alink_expf(x); }
float alink_oracle_logf(float x) { return alink_logf(x); }
uint64_t alink_oracle_pack_key(float u, uint64_t idx) { return alink_pack_key(u, idx); }

/* Verification-curve counts on an all-pairs score matrix — the loops of
 * /root/reference/utilities/ROC_precompute.py:19-63 as written: collect the genuine / impostor scores of the
 * strict upper triangle (for i: for j in range(i + 1, n), :22-23) by mask code and protocol (:24-41), then for
 * every threshold count the scores >= it (:52-58). scores[P][G] fp32 (the reference reads the same values back
 * as float64 with np.loadtxt, so the comparison is done in double), row0 = global index of the block's first
 * probe row. tp / fp: nthr counters; n_gen / n_imp: len(Genuine_score), len(Imposter_score) (:59,:61). */
int alink_oracle_roc_counts(const float *scores, const int32_t *mask, int64_t P, int64_t G, int64_t row0,
                            int roc_case, const double *thr, int nthr, uint64_t *tp, uint64_t *fp,
                            uint64_t *n_gen, uint64_t *n_imp) {
  if (roc_case < 1 || roc_case > 3 || P < 0 || G < 0 || nthr < 0) return -1;
  uint64_t ng = 0, ni = 0;
  for (int t = 0; t < nthr; ++t) tp[t] = fp[t] = 0;
  for (int64_t r = 0; r < P; ++r) {
    for (int64_t j = row0 + r + 1; j < G; ++j) {
      const int code = mask[r * G + j];
      int cls = -1;
      if (roc_case == 3) cls = (code == 1 || code == 2) ? 0 : ((code == 3 || code == 4) ? 1 : -1);
      else if (roc_case == 2) cls = code == 2 ? 0 : (code == 4 ? 1 : -1);
      else cls = code == 1 ? 0 : (code == 3 ? 1 : -1);
      if (cls < 0) continue;
      const double s = (double)scores[r * G + j];
      if (cls == 0) ++ng; else ++ni;
      uint64_t *out = cls == 0 ? tp : fp;
      for (int t = 0; t < nthr; ++t)
        if (s >= thr[t]) ++out[t];
    }
  }
  if (n_gen) *n_gen = ng;
  if (n_imp) *n_imp = ni;
  return 0;
}

/* ---- finetune ("teach") minibatch step: the CPU restatement csrc/train.cuh is checked against -------------
 * What Keras runs per batch in /root/reference/code/siamese.py:52-58 (finetune -> fit, batch 16) and :86-112
 * (train_on_batch / test_on_batch) for ONE learner: forward of the head (siamese.py:24-33), binary cross-entropy
 * and binary accuracy, backprop through the three Dense layers, Adadelta (siamese.py:35). Arithmetic of the
 * loss / optimizer: include/alink_train_math.h. Reduction orders (they define the bits): forward chains over k
 * ascending then + bias; input gradients over j ascending; weight and bias gradients over the batch index
 * ascending; batch means = sequential sums / nb. Parameters and accumulators are updated in place when apply.
 * params / acc_g / acc_d: six arrays in Keras order W1,b1,W2,b2,W3,b3. probs: nb*C or NULL. */
#include "../include/alink_train_math.h"

int alink_oracle_train_step(const float *ta, const float *tb, const int32_t *ia, const int32_t *ib, const float *y,
                            int nb, int D, int feat_mode, int H1, int H2, int C, float *const *params,
                            float *const *acc_g, float *const *acc_d, float lr, float rho, float eps, int apply,
                            float *loss_out, float *acc_out, float *probs) {
  if (nb < 1 || (C != 1 && C != 2) || !params || !y) return -1;
  const int Din = feat_mode ? 2 * D : D;
  float *W1 = params[0], *b1 = params[1], *W2 = params[2], *b2 = params[3], *W3 = params[4], *b3 = params[5];
  float *X = (float *)malloc(sizeof(float) * (size_t)nb * Din), *h1 = (float *)malloc(sizeof(float) * (size_t)nb * H1);
  float *h2 = (float *)malloc(sizeof(float) * (size_t)nb * H2), *dz1 = (float *)malloc(sizeof(float) * (size_t)nb * H1);
  float *dz2 = (float *)malloc(sizeof(float) * (size_t)nb * H2), *dz3 = (float *)malloc(sizeof(float) * (size_t)nb * 2);
  for (int n = 0; n < nb; ++n) {
    const int64_t ra = ia ? ia[n] : n, rb = ib ? ib[n] : n;
    for (int k = 0; k < Din; ++k) {
      const int kk = k < D ? k : k - D;
      const float a = ta[ra * D + kk], b = tb[rb * D + kk];
      X[(size_t)n * Din + k] = k < D ? fabsf(a - b) : a * b;
    }
  }
  for (int n = 0; n < nb; ++n)
    for (int j = 0; j < H1; ++j) {
      float acc = 0.0f;
      for (int k = 0; k < Din; ++k) acc = fmaf(X[(size_t)n * Din + k], W1[(size_t)k * H1 + j], acc);
      acc = acc + b1[j];
      h1[(size_t)n * H1 + j] = acc > 0.0f ? acc : 0.0f;
    }
  for (int n = 0; n < nb; ++n)
    for (int j = 0; j < H2; ++j) {
      float acc = 0.0f;
      for (int k = 0; k < H1; ++k) acc = fmaf(h1[(size_t)n * H1 + k], W2[(size_t)k * H2 + j], acc);
      acc = acc + b2[j];
      h2[(size_t)n * H2 + j] = acc > 0.0f ? acc : 0.0f;
    }
  const float inv_nb = 1.0f / (float)nb;
  float sl = 0.0f, sa = 0.0f;
  for (int n = 0; n < nb; ++n) {
    float z0 = 0.0f, z1 = 0.0f;
    for (int j = 0; j < H2; ++j) {
      z0 = fmaf(h2[(size_t)n * H2 + j], W3[j * C + 0], z0);
      if (C == 2) z1 = fmaf(h2[(size_t)n * H2 + j], W3[j * C + 1], z1);
    }
    z0 = z0 + b3[0];
    if (C == 2) z1 = z1 + b3[1];
    float p0, p1, a, d0, d1;
    const float l = alink_head_loss(C, z0, z1, y[n], &p0, &p1, &a, &d0, &d1);
    sl = sl + l;
    sa = sa + a;
    dz3[n * C + 0] = d0 * inv_nb;
    if (C == 2) dz3[n * C + 1] = d1 * inv_nb;
    if (probs) {
      if (C == 2) { probs[n * 2] = p0; probs[n * 2 + 1] = p1; }
      else probs[n] = p1;
    }
  }
  if (loss_out) *loss_out = sl / (float)nb;
  if (acc_out) *acc_out = sa / (float)nb;
  if (apply) {
    for (int n = 0; n < nb; ++n)
      for (int k = 0; k < H2; ++k) {
        float acc = 0.0f;
        if (h2[(size_t)n * H2 + k] > 0.0f)
          for (int j = 0; j < C; ++j) acc = fmaf(dz3[n * C + j], W3[k * C + j], acc);
        dz2[(size_t)n * H2 + k] = acc;
      }
    for (int n = 0; n < nb; ++n)
      for (int k = 0; k < H1; ++k) {
        float acc = 0.0f;
        if (h1[(size_t)n * H1 + k] > 0.0f)
          for (int j = 0; j < H2; ++j) acc = fmaf(dz2[(size_t)n * H2 + j], W2[(size_t)k * H2 + j], acc);
        dz1[(size_t)n * H1 + k] = acc;
      }
    /* weight / bias gradients + Adadelta, last layer first (every input gradient above used the OLD weights) */
    const float *As[3] = {h2, h1, X};
    const float *dZs[3] = {dz3, dz2, dz1};
    const int Kd[3] = {H2, H1, Din}, Jd[3] = {C, H2, H1}, wi[3] = {4, 2, 0};
    for (int L = 0; L < 3; ++L) {
      float *W = params[wi[L]], *b = params[wi[L] + 1];
      for (int k = 0; k < Kd[L]; ++k)
        for (int j = 0; j < Jd[L]; ++j) {
          float g = 0.0f;
          for (int n = 0; n < nb; ++n) g = fmaf(As[L][(size_t)n * Kd[L] + k], dZs[L][(size_t)n * Jd[L] + j], g);
          const size_t e = (size_t)k * Jd[L] + j;
          W[e] = alink_adadelta(W[e], g, &acc_g[wi[L]][e], &acc_d[wi[L]][e], lr, rho, eps);
        }
      for (int j = 0; j < Jd[L]; ++j) {
        float g = 0.0f;
        for (int n = 0; n < nb; ++n) g = g + dZs[L][(size_t)n * Jd[L] + j];
        b[j] = alink_adadelta(b[j], g, &acc_g[wi[L] + 1][j], &acc_d[wi[L] + 1][j], lr, rho, eps);
      }
    }
  }
  free(X); free(h1); free(h2); free(dz1); free(dz2); free(dz3);
  return 0;
}
