/* alink_math.h — the scalar fp32 arithmetic of the query-selection path, written ONCE and
 * compiled by both gcc (oracle/, plain C) and nvcc (alink_b200/csrc, device code).
 *
 * Why this exists: north_star asks for "selected index set bit-exact in fp32 mode". At N=1e6,
 * B=1e3 adjacent utilities near the B-th rank differ by ~1e-7, so set identity is only attainable
 * if the GPU fp32 path reproduces the CPU oracle bit for bit. Every operation below is therefore
 * an explicitly rounded IEEE-754 binary32 op (fma / mul / add / div, round-to-nearest-even) or an
 * integer bit operation; nothing depends on libm, on compiler contraction, or on fast-math.
 *   host  : build with -ffp-contract=off (oracle/Makefile); fmaf() must be a correctly rounded fma
 *           (-mfma inlines vfmadd; glibc's software fmaf is also correctly rounded).
 *   device: __fmaf_rn/__fmul_rn/__fadd_rn/__fdiv_rn are never contracted or reassociated by nvcc.
 *
 * What it restates (reference = iamgroot42/A-LINK, paths relative to /root/reference):
 *   softmax over the 2 logits ......... code/siamese.py:31-32   (Dense(2) + Activation('softmax'))
 *   sigmoid over the 1 logit .......... code/siamese3.py:25     (Dense(1, activation='sigmoid'))
 *   1 - max_c p ....................... code/uncertainty.py:26,82
 *   margin p(1) - p(2) ................ code/uncertainty.py:43-44,107-108
 *   entropy -sum p^ ln p^ ............. code/uncertainty.py:60,130 (scipy.stats.entropy: normalise, ln, entr(0)=0)
 *   KL(p_k || pbar) ................... modAL.disagreement.KL_max_disagreement (external)
 *   top-k key ......................... code/uncertainty.py:155 (multi_argmax) + north_star tie rule
 */
#ifndef ALINK_MATH_H
#define ALINK_MATH_H

#include <stdint.h>

#if defined(__CUDA_ARCH__)
#define ALINK_HD __host__ __device__ __forceinline__
#define ALINK_FMA(a, b, c) __fmaf_rn((a), (b), (c))
#define ALINK_MUL(a, b) __fmul_rn((a), (b))
#define ALINK_ADD(a, b) __fadd_rn((a), (b))
#define ALINK_SUB(a, b) __fsub_rn((a), (b))
#define ALINK_DIV(a, b) __fdiv_rn((a), (b))
#define ALINK_F2I(f) __float_as_int(f)
#define ALINK_I2F(i) __int_as_float(i)
#else
#include <math.h>
#include <string.h>
#if defined(__CUDACC__)
#define ALINK_HD __host__ __device__ inline
#else
#define ALINK_HD static inline
#endif
#define ALINK_FMA(a, b, c) fmaf((a), (b), (c))
#define ALINK_MUL(a, b) ((a) * (b))
#define ALINK_ADD(a, b) ((a) + (b))
#define ALINK_SUB(a, b) ((a) - (b))
#define ALINK_DIV(a, b) ((a) / (b))
static inline int32_t alink__f2i(float f) { int32_t i; memcpy(&i, &f, 4); return i; }
static inline float alink__i2f(int32_t i) { float f; memcpy(&f, &i, 4); return f; }
#define ALINK_F2I(f) alink__f2i(f)
#define ALINK_I2F(i) alink__i2f(i)
#endif

/* ---- enums shared by the C ABI, the oracle and the kernels --------------------------------- */
enum { ALINK_ACT_SOFTMAX = 0, ALINK_ACT_SIGMOID = 1 };
enum {
  ALINK_UTIL_UNCERTAINTY = 0,       /* 1 - max_c p            code/uncertainty.py:82            */
  ALINK_UTIL_NEG_MARGIN = 1,        /* -(p(1) - p(2))         code/uncertainty.py:107-108,183   */
  ALINK_UTIL_ENTROPY = 2,           /* H(p)                   code/uncertainty.py:130           */
  ALINK_UTIL_VOTE_ENTROPY = 3,      /* H(vote counts / K)     modAL.disagreement (learners.py:12,287) */
  ALINK_UTIL_CONSENSUS_ENTROPY = 4, /* H(mean_k p_k)          modAL.disagreement                */
  ALINK_UTIL_KL_MAX = 5,            /* max_k KL(p_k || pbar)  modAL.disagreement                */
  ALINK_UTIL_COUNT = 6
};

/* ---- exp: Cody-Waite reduction + degree-5 minimax on r^2 (Cephes expf coefficients) --------
 * Max error ~1 ulp over the clamped range; exp(0) == 1 exactly. Inputs are clamped to [-86, 88]
 * so the result is always a normal float (no subnormal / inf paths to disagree on). */
ALINK_HD float alink_expf(float x) {
  if (x < -86.0f) x = -86.0f;
  if (x > 88.0f) x = 88.0f;
  const float magic = 12582912.0f; /* 1.5 * 2^23: adding it rounds to nearest-even integer */
  float t = ALINK_FMA(x, 1.44269504088896341f, magic);
  float n = ALINK_SUB(t, magic);
  float r = ALINK_FMA(n, -0.693359375f, x);      /* ln2 hi (8 significant bits: n*hi exact) */
  r = ALINK_FMA(n, 2.12194440e-4f, r);           /* ln2 lo */
  float p = 1.9875691500e-4f;
  p = ALINK_FMA(p, r, 1.3981999507e-3f);
  p = ALINK_FMA(p, r, 8.3334519073e-3f);
  p = ALINK_FMA(p, r, 4.1665795894e-2f);
  p = ALINK_FMA(p, r, 1.6666665459e-1f);
  p = ALINK_FMA(p, r, 5.0000001201e-1f);
  float r2 = ALINK_MUL(r, r);
  float y = ALINK_FMA(p, r2, r);
  y = ALINK_ADD(y, 1.0f);
  int32_t ni = (int32_t)n; /* n is integral: conversion is exact */
  return ALINK_I2F(ALINK_F2I(y) + (int32_t)((uint32_t)ni << 23));
}

/* ---- ln for normal positive x (Cephes logf polynomial on [sqrt(1/2), sqrt(2))) -------------- */
ALINK_HD float alink_logf(float x) {
  int32_t ix = ALINK_F2I(x);
  int32_t e = (ix - 0x3f3504f3) >> 23; /* arithmetic shift; 0x3f3504f3 = bits(sqrt(0.5)) */
  float m = ALINK_I2F(ix - (int32_t)((uint32_t)e << 23));
  float f = ALINK_SUB(m, 1.0f);
  float fe = (float)e;
  float z = ALINK_MUL(f, f);
  float p = 7.0376836292e-2f;
  p = ALINK_FMA(p, f, -1.1514610310e-1f);
  p = ALINK_FMA(p, f, 1.1676998740e-1f);
  p = ALINK_FMA(p, f, -1.2420140846e-1f);
  p = ALINK_FMA(p, f, 1.4249322787e-1f);
  p = ALINK_FMA(p, f, -1.6668057665e-1f);
  p = ALINK_FMA(p, f, 2.0000714765e-1f);
  p = ALINK_FMA(p, f, -2.4999993993e-1f);
  p = ALINK_FMA(p, f, 3.3333331174e-1f);
  float y = ALINK_MUL(ALINK_MUL(p, f), z);
  y = ALINK_FMA(fe, -2.12194440e-4f, y);
  y = ALINK_FMA(z, -0.5f, y);
  float res = ALINK_ADD(f, y);
  res = ALINK_FMA(fe, 0.693359375f, res);
  return res;
}

/* entr(p) = -p ln p, entr(0) = 0 (scipy.special.entr) */
ALINK_HD float alink_entr(float p) {
  if (!(p > 0.0f)) return 0.0f;
  return -ALINK_MUL(p, alink_logf(p));
}

/* rel_entr(p, q) = p ln(p/q); 0 when p == 0; +inf when q == 0 < p (cannot happen: probabilities
 * here are bounded below by exp(-86)/2). */
ALINK_HD float alink_rel_entr(float p, float q) {
  if (!(p > 0.0f)) return 0.0f;
  return ALINK_MUL(p, alink_logf(ALINK_DIV(p, q)));
}

/* ---- output activations ---------------------------------------------------------------------
 * softmax over two logits, the TF/Keras formulation: exp(z - max) / sum  (code/siamese.py:32). */
ALINK_HD void alink_softmax2(float z0, float z1, float *p0, float *p1) {
  float m = z0 > z1 ? z0 : z1;
  float e0 = alink_expf(ALINK_SUB(z0, m));
  float e1 = alink_expf(ALINK_SUB(z1, m));
  float s = ALINK_ADD(e0, e1);
  *p0 = ALINK_DIV(e0, s);
  *p1 = ALINK_DIV(e1, s);
}

/* sigmoid: 1 / (1 + exp(-z))  (code/siamese3.py:25) */
ALINK_HD float alink_sigmoid(float z) {
  float e = alink_expf(-z);
  return ALINK_DIV(1.0f, ALINK_ADD(1.0f, e));
}

/* ---- per-row measures on a 2-column probability row (C==1 heads are widened to [1-p, p] the way
 * code/keras_wrapper.py:263-270 does before any measure sees them) --------------------------- */
ALINK_HD float alink_uncertainty2(float p0, float p1) { /* code/uncertainty.py:82 */
  float m = p0 > p1 ? p0 : p1;
  return ALINK_SUB(1.0f, m);
}
ALINK_HD float alink_neg_margin2(float p0, float p1) { /* code/uncertainty.py:107-108 + :183 sign */
  float hi = p0 > p1 ? p0 : p1;
  float lo = p0 > p1 ? p1 : p0;
  return -ALINK_SUB(hi, lo);
}
ALINK_HD float alink_entropy2(float p0, float p1) { /* code/uncertainty.py:130 */
  float s = ALINK_ADD(p0, p1);
  float q0 = ALINK_DIV(p0, s), q1 = ALINK_DIV(p1, s);
  return ALINK_ADD(alink_entr(q0), alink_entr(q1));
}
ALINK_HD float alink_kl2(float p0, float p1, float q0, float q1) {
  float sp = ALINK_ADD(p0, p1), sq = ALINK_ADD(q0, q1);
  float a0 = ALINK_DIV(p0, sp), a1 = ALINK_DIV(p1, sp);
  float b0 = ALINK_DIV(q0, sq), b1 = ALINK_DIV(q1, sq);
  return ALINK_ADD(alink_rel_entr(a0, b0), alink_rel_entr(a1, b1));
}

/* Utility of one pair from the K learners' 2-column rows (p[k*2 + c]) and their Bagging mean
 * (code/committee.py:18: sum in k order, then / K). K <= ALINK_MAX_K. */
#define ALINK_MAX_K 16
ALINK_HD void alink_consensus2(const float *p, int K, float *c0, float *c1) {
  float s0 = 0.0f, s1 = 0.0f;
  for (int k = 0; k < K; ++k) {
    s0 = ALINK_ADD(s0, p[2 * k + 0]);
    s1 = ALINK_ADD(s1, p[2 * k + 1]);
  }
  float fk = (float)K;
  *c0 = ALINK_DIV(s0, fk);
  *c1 = ALINK_DIV(s1, fk);
}
ALINK_HD float alink_utility2(int measure, const float *p, int K, float c0, float c1) {
  switch (measure) {
    case ALINK_UTIL_UNCERTAINTY: return alink_uncertainty2(c0, c1);
    case ALINK_UTIL_NEG_MARGIN: return alink_neg_margin2(c0, c1);
    case ALINK_UTIL_ENTROPY:
    case ALINK_UTIL_CONSENSUS_ENTROPY: return alink_entropy2(c0, c1);
    case ALINK_UTIL_VOTE_ENTROPY: {
      /* hard vote = argmax_c p (first index on ties, np.argmax; keras_wrapper.py:238-241) */
      int n1 = 0;
      for (int k = 0; k < K; ++k) n1 += (p[2 * k + 1] > p[2 * k + 0]) ? 1 : 0;
      float fk = (float)K;
      return alink_entropy2(ALINK_DIV((float)(K - n1), fk), ALINK_DIV((float)n1, fk));
    }
    case ALINK_UTIL_KL_MAX: {
      float best = alink_kl2(p[0], p[1], c0, c1);
      for (int k = 1; k < K; ++k) {
        float v = alink_kl2(p[2 * k], p[2 * k + 1], c0, c1);
        best = v > best ? v : best;
      }
      return best;
    }
    default: return 0.0f;
  }
}

/* ---- top-k ordering key ----------------------------------------------------------------------
 * Descending unsigned order on the key == (utility descending, then pair index ascending): the
 * north_star tie rule applied to modAL's multi_argmax (code/uncertainty.py:155). -0.0 is folded
 * onto +0.0 so that equal utilities always tie. idx < 2^32 - 1. */
ALINK_HD uint32_t alink_orderable(float u) {
  uint32_t b = (uint32_t)ALINK_F2I(u);
  if (b == 0x80000000u) b = 0u;
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
ALINK_HD float alink_unorderable(uint32_t o) {
  uint32_t b = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
  return ALINK_I2F((int32_t)b);
}
ALINK_HD uint64_t alink_pack_key(float u, uint64_t idx) {
  return ((uint64_t)alink_orderable(u) << 32) | (uint64_t)(0xFFFFFFFFu - (uint32_t)idx);
}
ALINK_HD uint64_t alink_key_index(uint64_t key) { return (uint64_t)(0xFFFFFFFFu - (uint32_t)key); }
ALINK_HD float alink_key_utility(uint64_t key) { return alink_unorderable((uint32_t)(key >> 32)); }

#endif /* ALINK_MATH_H */
