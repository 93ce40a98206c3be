// roc.cuh — verification-curve counts on an all-pairs score matrix (HBM-bound streaming reduction).
//
// Replaces the O(N^2) Python loops of /root/reference/utilities/ROC_precompute.py:19-63 (same rule in
// ROC_protocolwise.py): over the strict upper triangle (j > i) of the score matrix, mask codes select the
// genuine / impostor pairs of a protocol (roc_case 1: 1 vs 3, 2: 2 vs 4, 3: {1,2} vs {3,4}, lines 26-41), and
// for every threshold t: TP[t] = #{genuine : score >= t}, FP[t] = #{impostor : score >= t} (lines 52-63).
//
// One pass over the matrix block: 8 B per pair (fp32 score + int32 mask code). The thresholds are sorted once
// on the host and staged in shared memory; each pair does one binary search (k = #thresholds <= score, compared
// in double like the reference, which reads the matrix back as float64) and one shared-memory histogram
// increment; CTAs merge their histograms with global atomics; the host turns bins into ">= t" counts with a
// suffix sum. Exact integer counts: bit-identical to the reference loops whatever the order.
// A block of probe rows [row0, row0 + P) of a larger matrix can be counted on its own (sharding by probe-row
// block: the counts add up).
#pragma once
#include "common.cuh"

constexpr int ROC_MAX_THR = 2047;  // bins 0..nthr; 2 classes x 2048 x 4 B + 2047 x 8 B of shared memory

// class of a mask code under a protocol: 0 genuine, 1 impostor, -1 ignored (ROC_precompute.py:26-41)
__device__ __forceinline__ int roc_class(int code, int roc_case) {
  if (roc_case == 3) return (code == 1 || code == 2) ? 0 : ((code == 3 || code == 4) ? 1 : -1);
  if (roc_case == 2) return code == 2 ? 0 : (code == 4 ? 1 : -1);
  return code == 1 ? 0 : (code == 3 ? 1 : -1);
}

// hist: [2][nthr + 1] global counters (zeroed by the caller)
__global__ void __launch_bounds__(256) roc_hist_kernel(const float *__restrict__ scores, const int32_t *__restrict__ mask,
                                                       int64_t P, int64_t G, int64_t row0, int roc_case,
                                                       const double *__restrict__ thr_sorted, int nthr,
                                                       unsigned long long *__restrict__ hist) {
  extern __shared__ __align__(16) unsigned char roc_smem[];
  double *thr = reinterpret_cast<double *>(roc_smem);
  unsigned int *sh = reinterpret_cast<unsigned int *>(roc_smem + sizeof(double) * (size_t)(nthr ? nthr : 1));
  const int nb = nthr + 1;
  for (int i = threadIdx.x; i < nthr; i += 256) thr[i] = thr_sorted[i];
  for (int i = threadIdx.x; i < 2 * nb; i += 256) sh[i] = 0;
  __syncthreads();
  const int64_t total = P * G;
  for (int64_t e = (int64_t)blockIdx.x * 256 + threadIdx.x; e < total; e += (int64_t)gridDim.x * 256) {
    const int64_t r = e / G, j = e - r * G;
    if (j <= row0 + r) continue;  // strict upper triangle: for j in range(i + 1, n)
    const int cls = roc_class(__ldg(mask + e), roc_case);
    if (cls < 0) continue;
    const double s = (double)__ldg(scores + e);
    int lo = 0, hi = nthr;  // k = number of thresholds <= s (NaN: every comparison false -> 0)
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (thr[mid] <= s) lo = mid + 1;
      else hi = mid;
    }
    atomicAdd(&sh[cls * nb + lo], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * nb; i += 256)
    if (sh[i]) atomicAdd(&hist[i], (unsigned long long)sh[i]);
}
