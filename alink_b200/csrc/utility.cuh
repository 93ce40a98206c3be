// utility.cuh — K-learner reduction and utility measures (HBM-bound streaming kernel).
//
// Replaces Bagging.predict's np.sum(...)/K (/root/reference/code/committee.py:18),
// Committee.predict_proba's mean (code/learners.py:334-345), the C=1 widening
// (code/keras_wrapper.py:263-270), and the measures of code/uncertainty.py:63-130 plus modAL's
// vote_entropy / consensus_entropy / KL_max_disagreement (imported at code/learners.py:12).
// Arithmetic is include/alink_math.h, shared with the oracle -> bit-exact.
// Algorithmic bytes per pair: 4*K*C read + 4 written (+ 4*C consensus when requested).
#pragma once
#include "common.cuh"

template <int KT>  // KT > 0: compile-time K (registers); KT == 0: runtime K (local array)
__global__ void __launch_bounds__(256) utility_kernel(const float *__restrict__ probs, int64_t N, int K,
                                                      int C, int measure, float *__restrict__ util,
                                                      float *__restrict__ cons, const unsigned int *__restrict__ n_dev = nullptr) {
  const int KK = KT > 0 ? KT : K;
  if (n_dev) {  // the row count lives on the device (FP32_REFINE's candidate lists); N is then its capacity
    const int64_t nd = (int64_t)__ldg(n_dev);
    if (nd < N) N = nd;
  }
  for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < N;
       n += (int64_t)gridDim.x * blockDim.x) {
    float rows[2 * (KT > 0 ? KT : ALINK_MAX_K)];
    if (C == 2) {
      const float2 *src = reinterpret_cast<const float2 *>(probs) + (size_t)n * KK;
#pragma unroll
      for (int k = 0; k < KK; ++k) {
        const float2 v = __ldg(src + k);
        rows[2 * k] = v.x;
        rows[2 * k + 1] = v.y;
      }
    } else {
#pragma unroll
      for (int k = 0; k < KK; ++k) {
        const float p = __ldg(probs + (size_t)n * KK + k);
        rows[2 * k] = __fsub_rn(1.0f, p);  // keras_wrapper.py:268
        rows[2 * k + 1] = p;
      }
    }
    float c0, c1;
    alink_consensus2(rows, KK, &c0, &c1);
    if (cons) {
      if (C == 2)
        reinterpret_cast<float2 *>(cons)[n] = make_float2(c0, c1);
      else
        cons[n] = c1;
    }
    if (util) util[n] = alink_utility2(measure, rows, KK, c0, c1);
  }
}

static int launch_utility(alink_ctx *ctx, const float *probs, int64_t N, int K, int C, int measure,
                          float *util, float *cons, const unsigned int *n_dev = nullptr) {
  if (N == 0) return ALINK_OK;
  int64_t blocks = (N + 255) / 256;
  const int64_t cap = (int64_t)ctx->sm_count * 8;
  if (blocks > cap) blocks = cap;
  dim3 grid((unsigned)blocks);
  switch (K) {
    case 1: utility_kernel<1><<<grid, 256, 0, ctx->stream>>>(probs, N, K, C, measure, util, cons, n_dev); break;
    case 2: utility_kernel<2><<<grid, 256, 0, ctx->stream>>>(probs, N, K, C, measure, util, cons, n_dev); break;
    case 3: utility_kernel<3><<<grid, 256, 0, ctx->stream>>>(probs, N, K, C, measure, util, cons, n_dev); break;
    case 4: utility_kernel<4><<<grid, 256, 0, ctx->stream>>>(probs, N, K, C, measure, util, cons, n_dev); break;
    case 5: utility_kernel<5><<<grid, 256, 0, ctx->stream>>>(probs, N, K, C, measure, util, cons, n_dev); break;
    case 8: utility_kernel<8><<<grid, 256, 0, ctx->stream>>>(probs, N, K, C, measure, util, cons, n_dev); break;
    default: utility_kernel<0><<<grid, 256, 0, ctx->stream>>>(probs, N, K, C, measure, util, cons, n_dev); break;
  }
  ALINK_LAUNCHED(ctx);
  return ALINK_OK;
}
