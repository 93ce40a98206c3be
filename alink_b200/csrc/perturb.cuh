// perturb.cuh — the A2-LINK noise-augmented pool (BASELINE configs[3]), generated ON THE DEVICE.
//
// The reference perturbs the IMAGES of a batch with each noise type and re-featurises them
// (/root/reference/code/ALINK.py:163-170: bag.attackModel -> conversionModel.process -> disguisedFacesModel.predict per noise);
// its one-pixel attack moves 40 pixels (code/attack.py:91). The embedding-space pool of SURVEY 8d does the analogue on the
// features: e' = normalize(e + delta_t), delta_t = `ncoord` coordinates set to U(-vmax, vmax), drawn from a COUNTER-BASED
// stream keyed on (seed, pair, side, t, draw) so that the host oracle regenerates the same pool (its numpy restatement of
// these rows) and nothing crosses PCIe. One warp per output row; all arithmetic is single IEEE operations in a fixed
// order (no fma, a fixed reduction tree), so the rows are bit-identical to the numpy restatement.
// HBM-bound: D * 4 bytes read (the base row, usually an L2 hit) + D * 4 written per row.
#pragma once
#include "common.cuh"
#include "de.cuh"

// draw d of (pair, side, t): the stream of de_u01 with the key folded as (seed, pair, 2 t + side, d)
__device__ __forceinline__ double perturb_u01(uint64_t seed, uint64_t pair, uint64_t side, uint64_t t, uint64_t d) {
  return de_u01(seed, pair, 2ull * t + side, d, 0x5EEDull);
}

__global__ void __launch_bounds__(256) perturb_rows_kernel(const float *__restrict__ table, const int32_t *__restrict__ idx, int64_t N,
                                                           int D, int side, int t, uint64_t seed, uint64_t pair_base, int ncoord,
                                                           float vmax, int normalize, float *__restrict__ out) {
  extern __shared__ float sm[];  // [8 warps][D]
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  float *row = sm + (size_t)w * D;
  for (int64_t n = (int64_t)blockIdx.x * 8 + w; n < N; n += (int64_t)gridDim.x * 8) {
    const int64_t r = idx ? (int64_t)__ldg(idx + n) : n;
    const float4 *src = reinterpret_cast<const float4 *>(table + (size_t)r * D);
    for (int i = lane; i < D / 4; i += 32) reinterpret_cast<float4 *>(row)[i] = __ldg(src + i);
    __syncwarp();
    // delta: draw j sets coordinate c_j to base + v_j (ONE fp32 addition to the BASE value); later draws overwrite earlier
    // ones on the same coordinate. Lanes compute the draws in parallel (the hashes and the base-value loads are the cost);
    // the writes then go out in draw order.
    const uint64_t pair = pair_base + (uint64_t)n;
    for (int j0 = 0; j0 < ncoord; j0 += 32) {
      const int j = j0 + lane;
      int c = 0;
      float nv = 0.0f;
      if (j < ncoord) {
        c = (int)(perturb_u01(seed, pair, (uint64_t)side, (uint64_t)t, 2ull * j) * (double)D);
        if (c > D - 1) c = D - 1;
        const float v = (float)((2.0 * perturb_u01(seed, pair, (uint64_t)side, (uint64_t)t, 2ull * j + 1) - 1.0) * (double)vmax);
        nv = __fadd_rn(__ldg(table + (size_t)r * D + c), v);
      }
      const int cnt = ncoord - j0 < 32 ? ncoord - j0 : 32;
      for (int k = 0; k < cnt; ++k) {
        if (lane == k) row[c] = nv;
        __syncwarp();
      }
    }
    float ss = 0.0f;  // lane l: chunks l, l + 32, ... in order, x y z w in order; then the xor butterfly 16, 8, 4, 2, 1
    for (int i = lane; i < D / 4; i += 32) {
      const float4 v = reinterpret_cast<const float4 *>(row)[i];
      ss = __fadd_rn(ss, __fmul_rn(v.x, v.x));
      ss = __fadd_rn(ss, __fmul_rn(v.y, v.y));
      ss = __fadd_rn(ss, __fmul_rn(v.z, v.z));
      ss = __fadd_rn(ss, __fmul_rn(v.w, v.w));
    }
    for (int o = 16; o > 0; o >>= 1) ss = __fadd_rn(ss, __shfl_xor_sync(0xffffffffu, ss, o));
    const float inv = normalize ? __fdiv_rn(1.0f, __fsqrt_rn(ss)) : 1.0f;
    float4 *dst = reinterpret_cast<float4 *>(out + (size_t)n * D);
    for (int i = lane; i < D / 4; i += 32) {
      float4 v = reinterpret_cast<const float4 *>(row)[i];
      v.x = __fmul_rn(v.x, inv); v.y = __fmul_rn(v.y, inv); v.z = __fmul_rn(v.z, inv); v.w = __fmul_rn(v.w, inv);
      dst[i] = v;
    }
    __syncwarp();
  }
}

// column `col` of learner `learner` of the scores [N][K][C] -> out[N] (the p2[t] row of the A-LINK rule, ALINK.py:175)
__global__ void __launch_bounds__(256) extract_col_kernel(const float *__restrict__ probs, int64_t N, int K, int C, int learner, int col,
                                                          float *__restrict__ out) {
  for (int64_t n = (int64_t)blockIdx.x * 256 + threadIdx.x; n < N; n += (int64_t)gridDim.x * 256) {
    const float p = probs[((size_t)n * K + learner) * C + (C == 2 ? col : 0)];
    out[n] = (C == 1 && col == 0) ? __fsub_rn(1.0f, p) : p;
  }
}

// u[j] = min_t |p2[t][j] - p1[j]|: the smallest disparity over the noise types (config 4's second report)
__global__ void __launch_bounds__(256) min_disparity_kernel(const float *__restrict__ p2, const float *__restrict__ p1, int64_t N, int T,
                                                            float *__restrict__ u) {
  for (int64_t n = (int64_t)blockIdx.x * 256 + threadIdx.x; n < N; n += (int64_t)gridDim.x * 256) {
    const float c = __ldg(p1 + n);
    float m = INFINITY;
    for (int t = 0; t < T; ++t) m = fminf(m, fabsf(__fsub_rn(__ldg(p2 + (size_t)t * N + n), c)));
    u[n] = m;
  }
}
