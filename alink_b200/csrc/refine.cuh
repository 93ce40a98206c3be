// refine.cuh — FP32_REFINE: bit-exact fp32 selection at tensor-core speed (filter and refine).
//
// The fp32 selection of /root/reference/code/uncertainty.py:154-159 (top-n of the utility) only depends
// on the pairs near the top. So: (1) the tcgen05 bf16 head scores EVERY pair and yields approximate
// utilities u~ with |u~ - u| <= eps; (2) radix-select the B-th largest u~ (tau); (3) keep the candidates
// {n : u~[n] >= tau - 2 eps} — every pair of the exact top-B is among them, because the exact B-th value
// is >= tau - eps and a pair outside has u < tau - 2 eps + eps; (4) re-score only the candidates with the
// bit-exact FFMA head (head_ffma.cuh, the oracle's summation order) and (5) select the top-B of the exact
// utilities with the pairs' ORIGINAL indices in the keys. The result is identical to FP32_EXACT whenever
// the error bound holds. A PILOT narrows the band first: the pairs with u~ >= tau (at least B of them) are
// re-scored exactly; the smallest of their exact utilities, L, is a lower bound of the exact B-th largest value
// (B pairs reach it), so a pair of the exact top-B has u >= L and u~ >= L - eps — the candidates are
// {u~ >= L - eps}, a band of about eps + |error| below tau instead of 2 eps (a third fewer re-scored pairs).
// The bound is checked on every query: the observed max |u~ - u| over the
// candidates must stay below it, otherwise it is widened and steps (3)-(5) are repeated (alink_refine_stats
// / alink_refine_detail report survivors, max and rms error, the bound used and the number of rounds).
//
// Filter operand format: IEEE half by default (11 significant bits: the rounding error of the features,
// weights and hidden activations is 8x smaller than bf16's, so the candidate band is 8x narrower and the
// exact re-score 8x cheaper) — used only when it provably cannot overflow: |x| <= |a|max + |b|max and
// |h1| <= xmax * max_j sum_k |W1[k][j]| + |b1|max must stay below 6e4; otherwise the bf16 images are used.
#pragma once
#include "common.cuh"
#include "topk.cuh"

// ---- where the candidate band ends, in the space the error bound lives in -------------------------------------------------
// For the measures that are monotone functions of u = min(pbar, 1 - pbar) of the two-column consensus — uncertainty U = u,
// neg-margin U = 2u - 1, entropy / consensus entropy U = H(u) — the filter's error is bounded in u, not in U:
//   additive        |u~ - u| <= eps                        (probability error; eps = the stage's bound)
//   multiplicative  u e^-ez <= u~ <= u e^ez                 (each learner's LOGIT is off by at most ez, and sigma(d + e) lies
//                                                            between sigma(d) e^-|e| and sigma(d) e^|e|; so does the mean)
// Let a0 = g^-1(tau - nu) be the u of the B-th largest approximate utility (nu = evaluation noise of the fp32 measure).
// The approximate top-B members have true u >= a1 = max(a0 - eps, a0 e^-ez), so the exact B-th utility is >= g(a1) - nu; a
// pair at or above it has u >= a2 = g^-1(g(a1) - 2 nu) and therefore u~ >= a3 = max(a2 - eps, a2 e^-ez):
//   threshold = g(a3) - nu.
// For uncertainty this is tau - 2 eps (the additive band) unless the head is so confident that the B-th utility is small,
// where the multiplicative band takes over (the additive one would swallow the pool); for the entropy measures, whose
// utility is flat at the top (H ~ ln 2 - 2 (u - 1/2)^2), it is a band of 2 eps in u instead of 2.5 eps in H — 10-30 % of a
// pool. Other measures (KL-max) keep the additive band in utility space. Double precision, once per CTA.
__device__ __forceinline__ double refine_H(double u) {
  if (u <= 0.0) return 0.0;
  if (u >= 0.5) return 0.6931471805599453;
  return -(u * log(u) + (1.0 - u) * log(1.0 - u));
}
__device__ __forceinline__ double refine_Hinv(double h) {  // u in [0, 1/2] with H(u) = h
  if (h <= 0.0) return 0.0;
  if (h >= 0.6931471805599453) return 0.5;
  double lo = 0.0, hi = 0.5;
  for (int i = 0; i < 60; ++i) {
    const double mid = 0.5 * (lo + hi);
    if (refine_H(mid) < h) lo = mid; else hi = mid;
  }
  return lo;  // the lower end: a smaller u only widens the band
}
__device__ __forceinline__ double refine_g(int measure, double u) {
  return measure == ALINK_UTIL_UNCERTAINTY ? u : (measure == ALINK_UTIL_NEG_MARGIN ? 2.0 * u - 1.0 : refine_H(u));
}
__device__ __forceinline__ double refine_ginv(int measure, double U) {
  const double u = measure == ALINK_UTIL_UNCERTAINTY ? U : (measure == ALINK_UTIL_NEG_MARGIN ? 0.5 * (U + 1.0) : refine_Hinv(U));
  return u < 0.0 ? 0.0 : (u > 0.5 ? 0.5 : u);
}
__device__ __forceinline__ float refine_threshold(int measure, float tau, float margin, double eps, double ez) {
  const bool ent = measure == ALINK_UTIL_ENTROPY || measure == ALINK_UTIL_CONSENSUS_ENTROPY;
  if (!(measure == ALINK_UTIL_UNCERTAINTY || measure == ALINK_UTIL_NEG_MARGIN || ent) || !(eps > 0.0)) return __fsub_rn(tau, margin);
  const double nu = ent ? 4.0e-7 : 1.2e-7;
  const double shrink = exp(-ez);
  const double a0 = refine_ginv(measure, (double)tau - nu);
  const double a1 = fmax(a0 - eps, a0 * shrink);
  const double a2 = refine_ginv(measure, refine_g(measure, a1) - 2.0 * nu);
  const double a3 = fmax(a2 - eps, a2 * shrink);
  double thr = refine_g(measure, a3) - nu;
  if (!ent) {  // never wider than the plain additive band in utility space (the entropy band is the point of all this)
    const double add = (double)tau - (double)margin;
    if (add > thr) thr = add;
  }
  float t = (float)thr;
  if ((double)t > thr) t = nextafterf(t, -INFINITY);
  return t;
}

// candidates = pairs whose approximate utility is within `margin` of a reference value: the B-th largest
// approximate utility (st->prefix = the high 32 bits of its key: select_threshold(..., passes = 3)) or, when `pilot_min` is given, the smallest EXACT utility of the
// pilot set (an order-preserving uint, refine_min_kernel)
__global__ void __launch_bounds__(256) refine_mark_kernel(const float *__restrict__ util, int64_t N, const SelectState *st,
                                                          const unsigned int *__restrict__ pilot_min, float margin,
                                                          uint32_t *__restrict__ cand, unsigned int *__restrict__ count,
                                                          unsigned int cap, int measure = -1, double eps = 0.0, double ez = 0.0) {
  const float tau = alink_key_utility(st->prefix << 32);
  // never above tau: the pilot set itself always stays in (it matters when the caller's bound is too tight)
  const float thr = pilot_min ? fminf(__fsub_rn(alink_unorderable(*pilot_min), margin), tau)
                              : (measure >= 0 ? refine_threshold(measure, tau, margin, eps, ez) : __fsub_rn(tau, margin));
  const int lane = threadIdx.x & 31;
  // warp-uniform loop bound: every lane takes part in the ballot
  for (int64_t n0 = (int64_t)blockIdx.x * 256 + (threadIdx.x & ~31); n0 < N; n0 += (int64_t)gridDim.x * 256) {
    const int64_t n = n0 + lane;
    const bool keep = n < N && __ldg(util + n) >= thr;
    const unsigned int m = __ballot_sync(0xffffffffu, keep);  // one atomic per warp
    if (m) {
      const int first = __ffs(m) - 1;
      unsigned int basepos = 0;
      if (lane == first) basepos = atomicAdd(count, (unsigned int)__popc(m));
      basepos = __shfl_sync(0xffffffffu, basepos, first);
      const unsigned int pos = basepos + __popc(m & ((1u << lane) - 1u));
      if (keep && pos < cap) cand[pos] = (uint32_t)n;  // *count may exceed cap: the caller sees the overflow
    }
  }
}

// table rows of the candidates: cia[j] = ia[cand[j]] (or cand[j] for identity pairs)
__global__ void __launch_bounds__(256) refine_compose_kernel(const uint32_t *__restrict__ cand, unsigned int M,
                                                             const int32_t *__restrict__ ia, const int32_t *__restrict__ ib,
                                                             uint32_t G, int64_t pair0, int32_t *__restrict__ cia,
                                                             int32_t *__restrict__ cib) {
  for (unsigned int j = blockIdx.x * 256 + threadIdx.x; j < M; j += gridDim.x * 256) {
    const uint32_t n = cand[j];
    if (ia) {
      cia[j] = __ldg(ia + n);
      cib[j] = __ldg(ib + n);
    } else if (G) {
      const int64_t p = pair0 + n;
      const int64_t ra = p / G;
      cia[j] = (int32_t)ra;
      cib[j] = (int32_t)(p - ra * G);
    } else {
      cia[j] = cib[j] = (int32_t)n;
    }
  }
}

// max and sum of squares over the candidates of |exact - approximate| utility
// stat: [0] candidate count, [1] max error (ordered uint bits: values are >= 0), [2..3] double sum of squares
__global__ void __launch_bounds__(256) refine_maxerr_kernel(const float *__restrict__ exact, const float *__restrict__ approx,
                                                            const uint32_t *__restrict__ cand, unsigned int M,
                                                            unsigned int *__restrict__ stat) {
  float m = 0.0f, ss = 0.0f;
  for (unsigned int j = blockIdx.x * 256 + threadIdx.x; j < M; j += gridDim.x * 256) {
    const float e = fabsf(__fsub_rn(exact[j], __ldg(approx + cand[j])));
    m = fmaxf(m, e);
    ss = __fmaf_rn(e, e, ss);
  }
  for (int o = 16; o > 0; o >>= 1) {
    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    ss += __shfl_xor_sync(0xffffffffu, ss, o);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax(stat + 1, __float_as_uint(m));
    atomicAdd(reinterpret_cast<double *>(stat + 2), (double)ss);
  }
}

// smallest exact utility of the pilot set, as an order-preserving uint (out starts at 0xFFFFFFFF)
__global__ void __launch_bounds__(256) refine_min_kernel(const float *__restrict__ exact, unsigned int M,
                                                         unsigned int *__restrict__ out) {
  unsigned int m = 0xFFFFFFFFu;
  for (unsigned int j = blockIdx.x * 256 + threadIdx.x; j < M; j += gridDim.x * 256) m = min(m, alink_orderable(exact[j]));
  for (int o = 16; o > 0; o >>= 1) m = min(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) atomicMin(out, m);
}

// max |element| of a table (fp32): decides whether the half-precision filter can overflow (65504)
__global__ void __launch_bounds__(256) absmax_kernel(const float4 *__restrict__ x, int64_t n4, unsigned int *__restrict__ out_bits) {
  float m = 0.0f;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (int64_t)gridDim.x * 256) {
    const float4 v = __ldg(x + i);
    m = fmaxf(m, fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
  }
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  // NaN/Inf elements: fmaxf drops NaN, so test the bit pattern of what remains and flag non-finite separately
  if ((threadIdx.x & 31) == 0) atomicMax(out_bits, __float_as_uint(m));
}


// ======================================================================================================================
// FP32_REFINE, three stages without host round trips (refine_query_v2 in alink_b200.cu):
//   stage 1  half-precision tensor-core filter over all N pairs (u~, error <= eps1) -> candidates C1 = {u~ >= tau1 - 2 eps1}
//   stage 2  split-operand tensor-core re-score of C1 (u2, error <= eps2 ~ 1e-5, head_tc_split.cuh)
//            -> C2 = {n in C1 : u2 >= tau2 - 2 eps2}   (tau = the B-th largest value of the stage)
//   stage 3  bit-exact FFMA re-score of C2 (u), exact top-B with the pairs' original indices in the keys.
// Both bounds are CHECKED on every query, also on pairs OUTSIDE the candidate sets: a deterministic stratified sample S
// of the pool (>= 4096 pairs, as many as fill the last wave of the split kernel) rides through stage 2 — it verifies
// eps1 against u2 — and every 16th of those rides through stage 3 — it verifies eps2 against the exact value.
// Counts stay on the device (grids are sized for the capacities); the host reads 64 bytes of statistics with the result.
// rs[] words: 0 count1 | 1 M1 = n1 + |S| | 2 count2 | 3 M2 = n2 + |S2| | 4 flags (1: C1 overflow, 2: C2 overflow)
//             5 max |u~ - u2| over C1 u S | 6 max |u2 - u| over C2 u S2 | 7 max |u~ - u| over C2 u S2 | 8-9 sum sq of |u~ - u2| over C1 (double)
//             10-11 sum sq of 6 (double) | 12 n1 = min(count1, cap1) | 13 n2 = min(count2, cap2)
//             14 tau1 bits | 15 max relative error | 16 ok: the verification's verdict, evaluated ON THE DEVICE (refine_verify_kernel)
enum { RS_COUNT1 = 0, RS_M1 = 1, RS_COUNT2 = 2, RS_M2 = 3, RS_FLAGS = 4, RS_ERR1 = 5, RS_ERR2 = 6, RS_ERRA = 7, RS_SS1 = 8,
       RS_SS2 = 10, RS_N1 = 12, RS_N2 = 13, RS_TAU = 14, RS_REL1 = 15, RS_OK = 16, RS_WORDS = 32 };

// The host's acceptance test of a three-stage query (refine_query_v2, after its synchronisation), restated on the device so
// that a SHARDED step can ship the verdict with its keys instead of stopping for it: ok = no list overflowed, stage 2 within
// eps2 of the exact values, stage 1 within eps1 of stage 2 (+ eps2) and of the exact values, 8 sigma of the candidates'
// error inside eps1, the multiplicative band's premise intact. Double arithmetic, IEEE sqrt: the same verdict as the host's.
struct VerifyConsts {
  double eps1, eps2, scv, relmax;  // relmax = expm1(ez) for the banded measures, < 0: not banded
};
__global__ void refine_verify_kernel(unsigned int *__restrict__ rs, const VerifyConsts c) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const double e1 = (double)__uint_as_float(rs[RS_ERR1]), e2 = (double)__uint_as_float(rs[RS_ERR2]);
  const double ea = (double)__uint_as_float(rs[RS_ERRA]), rel1 = (double)__uint_as_float(rs[RS_REL1]);
  const double ss1 = *reinterpret_cast<const double *>(rs + RS_SS1);
  const unsigned int n1 = rs[RS_N1];
  const double seen2 = e2 / c.scv;
  const double s1a = e1 / c.scv + c.eps2, s1b = ea / c.scv;
  const double seen1 = s1a > s1b ? s1a : s1b;
  const double rms1 = n1 ? (double)(float)sqrt(ss1 / (double)n1) : 0.0;  // (the host rounds the rms to float first)
  const double tail1 = 8.0 * rms1 / c.scv;
  const bool badrel = c.relmax >= 0.0 && rel1 > c.relmax;
  const bool bad = rs[RS_FLAGS] != 0u || seen2 > c.eps2 || seen1 > c.eps1 || tail1 > c.eps1 || badrel;
  rs[RS_OK] = bad ? 0u : 1u;
}

// after refine_mark_kernel: clamp the count, append the stratified sample, translate rows to table rows
__global__ void __launch_bounds__(256) refine_stage1_finish_kernel(unsigned int *__restrict__ rs, unsigned int cap1, int64_t N,
                                                                   unsigned int ns_min, unsigned int ns_max, unsigned int wave_rows,
                                                                   uint32_t *__restrict__ cand1, const int32_t *__restrict__ ia,
                                                                   const int32_t *__restrict__ ib, uint32_t G, int64_t pair0,
                                                                   int32_t *__restrict__ cia, int32_t *__restrict__ cib) {
  const unsigned int count1 = rs[RS_COUNT1];
  const unsigned int n1 = count1 < cap1 ? count1 : cap1;
  // the sample fills the split kernel's last round of (tile pair, learner) units (at least ns_min pairs, at most ns_max):
  // wave_rows = clusters << 16 | K; a round is `clusters` units = clusters / K tile pairs of 256 rows
  const unsigned int ncl = wave_rows >> 16, Kl = wave_rows & 0xFFFFu;
  const unsigned int pairs_min = (n1 + ns_min + 255u) / 256u;
  const unsigned int rounds = (pairs_min * Kl + ncl - 1u) / ncl;
  unsigned int ns = (rounds * ncl / Kl) * 256u - n1;
  if (ns > ns_max) ns = ns_max;
  if ((int64_t)ns > N) ns = (unsigned int)N;
  const unsigned int M1 = n1 + ns;
  for (unsigned int r = blockIdx.x * 256 + threadIdx.x; r < M1; r += gridDim.x * 256) {
    uint32_t n;
    if (r < n1) {
      n = cand1[r];
    } else {  // pair floor((j + 1/2) N / ns): one per stratum of N / ns consecutive pairs
      const unsigned long long j = r - n1;
      n = (uint32_t)(((2ull * j + 1ull) * (unsigned long long)N) / (2ull * ns));
      cand1[r] = n;
    }
    if (ia) {
      cia[r] = __ldg(ia + n);
      cib[r] = __ldg(ib + n);
    } else if (G) {
      const int64_t p = pair0 + n;
      const int64_t ra = p / G;
      cia[r] = (int32_t)ra;
      cib[r] = (int32_t)(p - ra * G);
    } else {
      cia[r] = cib[r] = (int32_t)n;
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    rs[RS_M1] = M1;
    rs[RS_N1] = n1;
    if (count1 > cap1) atomicOr(rs + RS_FLAGS, 1u);
  }
}

// Stage 1's bookkeeping in ONE cooperative launch (was: select_init + 3 radix passes + mark + finish = 6 launches, each
// ~10 us of launch + tail latency for 4 MB of L2-resident data): three histogram passes over the utility bits with a
// grid-wide barrier after each (every CTA then scans the 2048-bin global histogram redundantly, so nothing has to be
// broadcast), the candidate compaction, a barrier, and the clamp / sample / table-row translation.
// hist: 3 x 2048 zeroed counters. tau_out: the B-th largest approximate utility's orderable bits (for the record).
struct Stage1Args {
  const float *util;
  int64_t N;
  unsigned long long B;
  float margin;
  int measure;          // refine_threshold: the band lives in u-space for the consensus measures (-1: additive)
  double eps, ez;
  unsigned int cap1, ns_min, ns_max, wave_rows;
  unsigned int *hist, *rs;
  uint32_t *cand1;
  const int32_t *ia, *ib;
  uint32_t G;
  int64_t pair0;
  int32_t *cia, *cib;
};

__global__ void __launch_bounds__(256) refine_stage1_coop_kernel(const Stage1Args a) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  __shared__ unsigned int sh[2048];
  __shared__ unsigned int part[256];
  const int tid = threadIdx.x, lane = tid & 31;
  const int64_t stride = (int64_t)gridDim.x * 256;
  unsigned long long remaining = a.B;
  unsigned int prefix = 0;  // decided high bits of the orderable utility
  const int shifts[3] = {21, 10, 0}, nbits[3] = {11, 11, 10};
#pragma unroll 1
  for (int p = 0; p < 3; ++p) {
    const int nb = 1 << nbits[p];
    for (int i = tid; i < nb; i += 256) sh[i] = 0;
    __syncthreads();
    const int hi = shifts[p] + nbits[p];
    for (int64_t i = (int64_t)blockIdx.x * 256 + tid; i < a.N; i += stride) {
      const unsigned int o = alink_orderable(__ldg(a.util + i));
      if (p == 0 || (o >> hi) == prefix) atomicAdd(&sh[(o >> shifts[p]) & (unsigned)(nb - 1)], 1u);
    }
    __syncthreads();
    unsigned int *gh = a.hist + p * 2048;
    for (int i = tid; i < nb; i += 256)
      if (sh[i]) atomicAdd(gh + i, sh[i]);
    __threadfence();
    grid.sync();
    unsigned int bin;
    stage1_pick(gh, nb, remaining, bin, sh, part);
    prefix = p == 0 ? bin : ((prefix << nbits[p]) | bin);
  }
  // candidates: u~ >= tau - margin (ballot compaction, one atomic per warp)
  const float tau = alink_unorderable(prefix);
  __shared__ float s_thr;
  if (tid == 0) s_thr = a.measure >= 0 ? refine_threshold(a.measure, tau, a.margin, a.eps, a.ez) : __fsub_rn(tau, a.margin);
  __syncthreads();
  const float thr = s_thr;
  unsigned int *count = a.rs + RS_COUNT1;
  for (int64_t n0 = (int64_t)blockIdx.x * 256 + (tid & ~31); n0 < a.N; n0 += stride) {
    const int64_t n = n0 + lane;
    const bool keep = n < a.N && __ldg(a.util + n) >= thr;
    const unsigned int m = __ballot_sync(0xffffffffu, keep);
    if (m) {
      const int first = __ffs(m) - 1;
      unsigned int basepos = 0;
      if (lane == first) basepos = atomicAdd(count, (unsigned int)__popc(m));
      basepos = __shfl_sync(0xffffffffu, basepos, first);
      const unsigned int pos = basepos + __popc(m & ((1u << lane) - 1u));
      if (keep && pos < a.cap1) a.cand1[pos] = (uint32_t)n;
    }
  }
  __threadfence();
  grid.sync();
  // clamp, sample, table rows (refine_stage1_finish_kernel's body)
  const unsigned int count1 = __ldcg(count);
  const unsigned int n1 = count1 < a.cap1 ? count1 : a.cap1;
  const unsigned int ncl = a.wave_rows >> 16, Kl = a.wave_rows & 0xFFFFu;
  const unsigned int pairs_min = (n1 + a.ns_min + 255u) / 256u;
  const unsigned int rounds = (pairs_min * Kl + ncl - 1u) / ncl;
  unsigned int ns = (rounds * ncl / Kl) * 256u - n1;
  if (ns > a.ns_max) ns = a.ns_max;
  if ((int64_t)ns > a.N) ns = (unsigned int)a.N;
  const unsigned int M1 = n1 + ns;
  for (unsigned int r = blockIdx.x * 256 + tid; r < M1; r += gridDim.x * 256) {
    uint32_t n;
    if (r < n1) {
      n = __ldcg(a.cand1 + r);
    } else {
      const unsigned long long j = r - n1;
      n = (uint32_t)(((2ull * j + 1ull) * (unsigned long long)a.N) / (2ull * ns));
      a.cand1[r] = n;
    }
    if (a.ia) {
      a.cia[r] = __ldg(a.ia + n);
      a.cib[r] = __ldg(a.ib + n);
    } else if (a.G) {
      const int64_t pp = a.pair0 + n;
      const int64_t ra = pp / a.G;
      a.cia[r] = (int32_t)ra;
      a.cib[r] = (int32_t)(pp - ra * a.G);
    } else {
      a.cia[r] = a.cib[r] = (int32_t)n;
    }
  }
  if (blockIdx.x == 0 && tid == 0) {
    a.rs[RS_M1] = M1;
    a.rs[RS_N1] = n1;
    a.rs[RS_TAU] = prefix;
    if (count1 > a.cap1) atomicOr(a.rs + RS_FLAGS, 1u);
  }
}

// stage 2 -> stage 3 hand-over in ONE CTA: B-th largest split-operand utility among the n1 candidates (radix select on
// the utility bits), C2 = rows within `margin` of it, plus every (|S| / ns2)-th sample row; emits the table rows, the
// original pair index and the stage-2 row of every selected row.
__global__ void __launch_bounds__(1024) refine_stage2_kernel(unsigned int *__restrict__ rs, const float *__restrict__ util2,
                                                             unsigned long long B, float margin, int measure, double eps,
                                                             double ez, unsigned int cap2,
                                                             unsigned int ns2_max, const uint32_t *__restrict__ cand1,
                                                             const int32_t *__restrict__ cia1, const int32_t *__restrict__ cib1,
                                                             int32_t *__restrict__ cia2, int32_t *__restrict__ cib2,
                                                             uint32_t *__restrict__ map2, uint32_t *__restrict__ src2) {
  __shared__ unsigned int sh[2048];
  __shared__ unsigned int part[256];
  __shared__ unsigned long long s_prefix, s_remaining;
  __shared__ unsigned int s_count;
  const int tid = threadIdx.x;
  const unsigned int n1 = rs[RS_N1], M1 = rs[RS_M1];
  if (tid == 0) s_count = 0;
  KeySrc src{util2, nullptr, 0, nullptr};
  radix_select_cta(src, (int)n1, B < n1 ? B : (unsigned long long)(n1 ? n1 : 1), 3, sh, part, s_prefix, s_remaining);
  __shared__ float s_thr;
  if (tid == 0) {
    const float tau2 = alink_key_utility(s_prefix << 32);
    s_thr = measure >= 0 ? refine_threshold(measure, tau2, margin, eps, ez) : __fsub_rn(tau2, margin);
  }
  __syncthreads();
  const float thr = s_thr;
  for (unsigned int r = tid; r < n1; r += 1024) {
    if (util2[r] >= thr) {
      const unsigned int pos = atomicAdd(&s_count, 1u);
      if (pos < cap2) {
        src2[pos] = r;
        map2[pos] = cand1[r];
        cia2[pos] = cia1[r];
        cib2[pos] = cib1[r];
      }
    }
  }
  __syncthreads();
  const unsigned int count2 = s_count, n2 = count2 < cap2 ? count2 : cap2;
  const unsigned int nsamp = M1 - n1;
  const unsigned int ns2 = nsamp < ns2_max ? nsamp : ns2_max;
  for (unsigned int j = tid; j < ns2; j += 1024) {
    const unsigned int r = n1 + (unsigned int)(((unsigned long long)j * nsamp) / ns2);
    src2[n2 + j] = r;
    map2[n2 + j] = cand1[r];
    cia2[n2 + j] = cia1[r];
    cib2[n2 + j] = cib1[r];
  }
  if (tid == 0) {
    rs[RS_COUNT2] = count2;
    rs[RS_N2] = n2;
    rs[RS_M2] = n2 + ns2;
    if (count2 > cap2 || n1 < B) atomicOr(rs + RS_FLAGS, 2u);
  }
}

// exact utilities of the stage-3 rows from their per-learner probabilities, and the error statistics of both filters
// u = min(pbar, 1 - pbar) of the two-column consensus of one row of per-learner probabilities
template <int KT>
__device__ __forceinline__ float consensus_u(const float *__restrict__ probs, size_t row, int K, int C, float *rows) {
  const int KK = KT > 0 ? KT : K;
  if (C == 2) {
    const float2 *src = reinterpret_cast<const float2 *>(probs) + row * KK;
#pragma unroll
    for (int k = 0; k < KK; ++k) {
      const float2 v = src[k];
      rows[2 * k] = v.x;
      rows[2 * k + 1] = v.y;
    }
  } else {
#pragma unroll
    for (int k = 0; k < KK; ++k) {
      const float p = probs[row * KK + k];
      rows[2 * k] = __fsub_rn(1.0f, p);  // keras_wrapper.py:268
      rows[2 * k + 1] = p;
    }
  }
  float c0, c1;
  alink_consensus2(rows, KK, &c0, &c1);
  return fminf(c0, c1);
}

// Exact utilities of the stage-3 rows, and the error statistics of both filters. band >= 0 (a consensus measure): the errors
// are taken in PROBABILITY space, u = min(pbar, 1 - pbar) — from the per-learner probabilities of stages 2 and 3, and from
// the inverse of the measure for the filter's fused utility — because that is where the bounds live; in utility space the
// entropy measures, flat at the top, would hide a probability error exactly where the candidates sit. band < 0: utility space.
// rs[RS_REL1]: max relative error |u~ - u2| / u2 over rows with u2 >= 1e-3 (the multiplicative band's premise).
template <int KT>
__global__ void __launch_bounds__(256) refine_finish_kernel(unsigned int *__restrict__ rs, const float *__restrict__ probs, int K,
                                                            int C, int measure, const float *__restrict__ util_a,
                                                            const float *__restrict__ util2, const uint32_t *__restrict__ cand1,
                                                            const uint32_t *__restrict__ map2, const uint32_t *__restrict__ src2,
                                                            float *__restrict__ util_c, const float *__restrict__ probs2, int band) {
  const int KK = KT > 0 ? KT : K;
  const unsigned int M1 = rs[RS_M1], M2 = rs[RS_M2], n1 = rs[RS_N1];
  float m1 = 0.0f, m2 = 0.0f, ma = 0.0f, ss1 = 0.0f, ss2 = 0.0f, rel = 0.0f;
  float rows[2 * (KT > 0 ? KT : ALINK_MAX_K)];
  for (unsigned int j = blockIdx.x * 256 + threadIdx.x; j < M2; j += gridDim.x * 256) {
    const float ux = consensus_u<KT>(probs, j, K, C, rows);
    float c0, c1;
    alink_consensus2(rows, KK, &c0, &c1);
    const float u = alink_utility2(measure, rows, KK, c0, c1);
    util_c[j] = u;
    float e2, ea;
    if (band >= 0) {
      float r2[2 * (KT > 0 ? KT : ALINK_MAX_K)];
      e2 = fabsf(__fsub_rn(consensus_u<KT>(probs2, src2[j], K, C, r2), ux));
      ea = fabsf((float)refine_ginv(band, (double)__ldg(util_a + map2[j])) - ux);
    } else {
      e2 = fabsf(__fsub_rn(util2[src2[j]], u));
      ea = fabsf(__fsub_rn(__ldg(util_a + map2[j]), u));
    }
    m2 = fmaxf(m2, e2);
    ma = fmaxf(ma, ea);
    ss2 = __fmaf_rn(e2, e2, ss2);
  }
  for (unsigned int r = blockIdx.x * 256 + threadIdx.x; r < M1; r += gridDim.x * 256) {
    float e1;
    if (band >= 0) {
      const float u2p = consensus_u<KT>(probs2, r, K, C, rows);
      e1 = fabsf((float)refine_ginv(band, (double)__ldg(util_a + cand1[r])) - u2p);
      if (u2p >= 1.0e-3f) rel = fmaxf(rel, __fdiv_rn(e1, u2p));
    } else {
      e1 = fabsf(__fsub_rn(__ldg(util_a + cand1[r]), util2[r]));
    }
    m1 = fmaxf(m1, e1);
    if (r < n1) ss1 = __fmaf_rn(e1, e1, ss1);  // rms over the candidates (the pairs at risk: p near 1/2, largest slope)
  }
  for (int o = 16; o > 0; o >>= 1) {
    m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, o));
    m2 = fmaxf(m2, __shfl_xor_sync(0xffffffffu, m2, o));
    ma = fmaxf(ma, __shfl_xor_sync(0xffffffffu, ma, o));
    rel = fmaxf(rel, __shfl_xor_sync(0xffffffffu, rel, o));
    ss1 += __shfl_xor_sync(0xffffffffu, ss1, o);
    ss2 += __shfl_xor_sync(0xffffffffu, ss2, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (m1 > 0.0f) atomicMax(rs + RS_ERR1, __float_as_uint(m1));
    if (m2 > 0.0f) atomicMax(rs + RS_ERR2, __float_as_uint(m2));
    if (ma > 0.0f) atomicMax(rs + RS_ERRA, __float_as_uint(ma));
    if (rel > 0.0f) atomicMax(rs + RS_REL1, __float_as_uint(rel));
    if (ss1 > 0.0f) atomicAdd(reinterpret_cast<double *>(rs + RS_SS1), (double)ss1);
    if (ss2 > 0.0f) atomicAdd(reinterpret_cast<double *>(rs + RS_SS2), (double)ss2);
  }
}

// max |element| and max row L2 norm of an fp32 table: one warp per row (D % 4 == 0)
// out[0] max |element| bits, out[1] max row sum of squares bits, out[2..3] double: sum over rows of the row sum of squares
__global__ void __launch_bounds__(256) table_stats_kernel(const float *__restrict__ x, int64_t M, int D, unsigned int *__restrict__ out) {
  const int lane = threadIdx.x & 31;
  float amax = 0.0f, nmax = 0.0f, nsum = 0.0f;
  for (int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); row < M; row += (int64_t)gridDim.x * 8) {
    const float4 *p = reinterpret_cast<const float4 *>(x + (size_t)row * D);
    float ss = 0.0f;
    for (int i = lane; i < D / 4; i += 32) {
      const float4 v = __ldg(p + i);
      amax = fmaxf(amax, fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
      ss = __fmaf_rn(v.x, v.x, ss); ss = __fmaf_rn(v.y, v.y, ss); ss = __fmaf_rn(v.z, v.z, ss); ss = __fmaf_rn(v.w, v.w, ss);
    }
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    nmax = fmaxf(nmax, ss);
    nsum += ss;
  }
  for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
  if (lane == 0) {
    atomicMax(out, __float_as_uint(amax));        // non-finite elements: fmaxf drops NaN, inf stays inf (fails the bounds)
    atomicMax(out + 1, __float_as_uint(nmax));    // max over rows of sum of squares (>= 0: uint order == float order)
    atomicAdd(reinterpret_cast<double *>(out + 2), (double)nsum);
  }
}

// ======================================================================================================================
// FP32_REFINE for VOTE ENTROPY (modAL's vote_entropy_sampling, the DEFAULT query strategy of the reference's Committee,
// code/learners.py:287): the utility is a function of the K hard votes, so it is exact wherever every vote is certain.
// The filter also returns vmin = min_k |p1 - p0| per pair; a vote can only flip where vmin is within twice the
// per-learner probability bound. Those pairs are re-scored with split operands, their votes replace the filter's, and
// the handful still within the split stage's bound is re-scored exactly. The utilities of ALL N pairs are then exact
// and the ordinary top-B (ties by lowest index: the plateaus are huge) runs over them.
__global__ void __launch_bounds__(256) vote_mark_kernel(const float *__restrict__ vmin, int64_t N, float thr, uint32_t *__restrict__ cand,
                                                        unsigned int *__restrict__ count, unsigned int cap) {
  const int lane = threadIdx.x & 31;
  for (int64_t n0 = (int64_t)blockIdx.x * 256 + (threadIdx.x & ~31); n0 < N; n0 += (int64_t)gridDim.x * 256) {
    const int64_t n = n0 + lane;
    const bool keep = n < N && __ldg(vmin + n) < thr;
    const unsigned int m = __ballot_sync(0xffffffffu, keep);
    if (m) {
      const int first = __ffs(m) - 1;
      unsigned int basepos = 0;
      if (lane == first) basepos = atomicAdd(count, (unsigned int)__popc(m));
      basepos = __shfl_sync(0xffffffffu, basepos, first);
      const unsigned int pos = basepos + __popc(m & ((1u << lane) - 1u));
      if (keep && pos < cap) cand[pos] = (uint32_t)n;
    }
  }
}

// rows of per-learner probabilities -> widened rows[], returns min_k |p1 - p0|
template <int KT>
__device__ __forceinline__ float vote_rows(const float *__restrict__ probs, size_t row, int K, int C, float *rows) {
  const int KK = KT > 0 ? KT : K;
  float vm = 2.0f;
#pragma unroll
  for (int k = 0; k < KK; ++k) {
    float p0, p1;
    if (C == 2) {
      const float2 v = reinterpret_cast<const float2 *>(probs)[row * KK + k];
      p0 = v.x;
      p1 = v.y;
    } else {
      p1 = probs[row * KK + k];
      p0 = __fsub_rn(1.0f, p1);
    }
    rows[2 * k] = p0;
    rows[2 * k + 1] = p1;
    vm = fminf(vm, fabsf(__fsub_rn(p1, p0)));
  }
  return vm;
}

// after the split stage: the votes of C1 replace the filter's; pairs still within thr2 of a flip go to the exact stage;
// statistics: |vmin(filter) - vmin(split)| / 2 estimates the filter's per-learner probability error on C1 u S; a SAMPLE
// row (a pair the filter called certain) whose vote entropy differs from the split stage's is a violation (flag 4).
template <int KT>
__global__ void __launch_bounds__(256) vote_stage2_kernel(unsigned int *__restrict__ rs, const float *__restrict__ probs2, int K, int C,
                                                          float *__restrict__ util_a, const float *__restrict__ vmin,
                                                          const uint32_t *__restrict__ cand1, float thr1, float thr2,
                                                          unsigned int cap2, const int32_t *__restrict__ cia1,
                                                          const int32_t *__restrict__ cib1, int32_t *__restrict__ cia2,
                                                          int32_t *__restrict__ cib2, uint32_t *__restrict__ map2,
                                                          uint32_t *__restrict__ src2) {
  const int KK = KT > 0 ? KT : K;
  const unsigned int M1 = rs[RS_M1], n1 = rs[RS_N1];
  float m1 = 0.0f, ss1 = 0.0f;
  float rows[2 * (KT > 0 ? KT : ALINK_MAX_K)];
  for (unsigned int r = blockIdx.x * 256 + threadIdx.x; r < M1; r += gridDim.x * 256) {
    const float vm2 = vote_rows<KT>(probs2, r, K, C, rows);
    const float u2 = alink_utility2(ALINK_UTIL_VOTE_ENTROPY, rows, KK, 0.0f, 0.0f);
    const uint32_t n = cand1[r];
    const float vm1 = __ldg(vmin + n);
    const float e1 = 0.5f * fabsf(__fsub_rn(vm1, vm2));
    m1 = fmaxf(m1, e1);
    if (r < n1) {
      ss1 = __fmaf_rn(e1, e1, ss1);
      util_a[n] = u2;
      if (vm2 < thr2) {
        const unsigned int pos = atomicAdd(rs + RS_COUNT2, 1u);
        if (pos < cap2) {
          src2[pos] = r;
          map2[pos] = n;
          cia2[pos] = cia1[r];
          cib2[pos] = cib1[r];
        }
      }
    } else if (vm1 >= thr1 && util_a[n] != u2) {  // (a sample pair that is also in C1 is no witness: C1 rewrites it)
      atomicOr(rs + RS_FLAGS, 4u);
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, o));
    ss1 += __shfl_xor_sync(0xffffffffu, ss1, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (m1 > 0.0f) atomicMax(rs + RS_ERR1, __float_as_uint(m1));
    if (ss1 > 0.0f) atomicAdd(reinterpret_cast<double *>(rs + RS_SS1), (double)ss1);
  }
}

// clamp |C2|, append ns2 evenly spaced sample rows for the exact stage
__global__ void __launch_bounds__(256) vote_stage2_finish_kernel(unsigned int *__restrict__ rs, unsigned int cap2, unsigned int ns2_max,
                                                                 const uint32_t *__restrict__ cand1, const int32_t *__restrict__ cia1,
                                                                 const int32_t *__restrict__ cib1, int32_t *__restrict__ cia2,
                                                                 int32_t *__restrict__ cib2, uint32_t *__restrict__ map2,
                                                                 uint32_t *__restrict__ src2) {
  const unsigned int count2 = rs[RS_COUNT2], n2 = count2 < cap2 ? count2 : cap2;
  const unsigned int n1 = rs[RS_N1], nsamp = rs[RS_M1] - n1;
  const unsigned int ns2 = nsamp < ns2_max ? nsamp : ns2_max;
  for (unsigned int j = threadIdx.x; j < ns2; j += 256) {
    const unsigned int r = n1 + (unsigned int)(((unsigned long long)j * nsamp) / ns2);
    src2[n2 + j] = r;
    map2[n2 + j] = cand1[r];
    cia2[n2 + j] = cia1[r];
    cib2[n2 + j] = cib1[r];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    rs[RS_N2] = n2;
    rs[RS_M2] = n2 + ns2;
    if (count2 > cap2) atomicOr(rs + RS_FLAGS, 2u);
  }
}

// exact votes of C2 replace the split stage's; per-learner |p2 - p| is the split stage's error; an exact SAMPLE row whose
// vote entropy differs from what the earlier stages left in util_a is a violation (flag 8)
template <int KT>
__global__ void __launch_bounds__(256) vote_finish_kernel(unsigned int *__restrict__ rs, const float *__restrict__ probs,
                                                          const float *__restrict__ probs2, int K, int C, float *__restrict__ util_a,
                                                          const uint32_t *__restrict__ map2, const uint32_t *__restrict__ src2,
                                                          float thr2) {
  const int KK = KT > 0 ? KT : K;
  const unsigned int M2 = rs[RS_M2], n2 = rs[RS_N2];
  float m2 = 0.0f, ss2 = 0.0f;
  float rows[2 * (KT > 0 ? KT : ALINK_MAX_K)], r2[2 * (KT > 0 ? KT : ALINK_MAX_K)];
  for (unsigned int j = blockIdx.x * 256 + threadIdx.x; j < M2; j += gridDim.x * 256) {
    vote_rows<KT>(probs, j, K, C, rows);
    const float vm2 = vote_rows<KT>(probs2, src2[j], K, C, r2);
    const float u = alink_utility2(ALINK_UTIL_VOTE_ENTROPY, rows, KK, 0.0f, 0.0f);
    float e2 = 0.0f;
#pragma unroll
    for (int k = 0; k < KK; ++k) e2 = fmaxf(e2, fabsf(__fsub_rn(rows[2 * k + 1], r2[2 * k + 1])));
    m2 = fmaxf(m2, e2);
    ss2 = __fmaf_rn(e2, e2, ss2);
    const uint32_t n = map2[j];
    if (j < n2) util_a[n] = u;
    else if (vm2 >= thr2 && util_a[n] != u) atomicOr(rs + RS_FLAGS, 8u);  // (vm2 < thr2: the pair may sit in C2 itself)
  }
  for (int o = 16; o > 0; o >>= 1) {
    m2 = fmaxf(m2, __shfl_xor_sync(0xffffffffu, m2, o));
    ss2 += __shfl_xor_sync(0xffffffffu, ss2, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (m2 > 0.0f) atomicMax(rs + RS_ERR2, __float_as_uint(m2));
    if (ss2 > 0.0f) atomicAdd(reinterpret_cast<double *>(rs + RS_SS2), (double)ss2);
  }
}
