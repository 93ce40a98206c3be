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

// candidates = pairs whose approximate utility is within `margin` of a reference value: the B-th largest
// approximate utility (st->prefix = the high 32 bits of its key: select_threshold(..., passes = 3)) or, when `pilot_min` is given, the smallest EXACT utility of the
// pilot set (an order-preserving uint, refine_min_kernel)
__global__ void __launch_bounds__(256) refine_mark_kernel(const float *__restrict__ util, int64_t N, const SelectState *st,
                                                          const unsigned int *__restrict__ pilot_min, float margin,
                                                          uint32_t *__restrict__ cand, unsigned int *__restrict__ count) {
  const float tau = alink_key_utility(st->prefix << 32);
  // never above tau: the pilot set itself always stays in (it matters when the caller's bound is too tight)
  const float thr = pilot_min ? fminf(__fsub_rn(alink_unorderable(*pilot_min), margin), tau) : __fsub_rn(tau, margin);
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
      if (keep) cand[basepos + __popc(m & ((1u << lane) - 1u))] = (uint32_t)n;
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
