// de.cuh — batched differential evolution in embedding space: the adversarial pool generator.
//
// Replaces the generation loop of /root/reference/code/differential_evolution.py (a SciPy fork whose objective takes
// the WHOLE population in one call, :630-645, :692-714) as /root/reference/code/attack.py:59-103 drives it
// (strategy best1bin, mutation dithered in [0.5, 1), recombination 1, polish off), with the siamese head as the
// batched objective. The reference perturbs PIXELS and scores them through a conv encoder that is out of scope; the
// analogue here perturbs 40 COORDINATES of the left embedding of a pair: a member is `coord_count` (index, value)
// genes in [0, 1) — perturb_image (attack.py:5-29) assigns img[x, y] = rgb, this assigns e[int(index)] = value.
//
// One DE problem per pair, P problems in lockstep, n members x G genes each. Population arithmetic is float64 like
// numpy's (no contraction: the library is built with -fmad=false), energies are the fp32 `1 - P(target)` of the head, so
// with the reference's random stream (drawn on the host in the reference's order, alink_b200/attack.py) populations and
// energies are IDENTICAL to the reference solver's, generation by generation; with the device stream (a counter-based
// hash keyed on seed, problem, generation, member, gene) a generation needs no host round trip.
#pragma once
#include "common.cuh"

// splitmix64-style counter hash -> uniform double in [0, 1) (the device random stream)
__device__ __forceinline__ double de_u01(uint64_t seed, uint64_t a, uint64_t b, uint64_t c, uint64_t d) {
  uint64_t z = seed ^ (a * 0x9E3779B97F4A7C15ull) ^ (b * 0xBF58476D1CE4E5B9ull) ^ (c * 0x94D049BB133111EBull) ^ (d * 0xD6E8FEB86659FD93ull);
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return (double)(z >> 11) * (1.0 / 9007199254740992.0);
}

// best1bin (differential_evolution.py:782-826): bprime = pop[0] + scale * (pop[r0] - pop[r1]); trial = where(cross, bprime,
// pop[c]); genes outside [0, 1] are flagged for _ensure_constraint (:775-780). cross == nullptr: recombination = 1.
// device stream (r0 == nullptr): scale, the two samples and the crossover draws come from de_u01.
__global__ void __launch_bounds__(256) de_mutate_kernel(const double *__restrict__ pop, int P, int n, int G,
                                                        const double *__restrict__ scale, const int32_t *__restrict__ r0,
                                                        const int32_t *__restrict__ r1, const uint8_t *__restrict__ cross,
                                                        uint64_t seed, uint64_t gen, double dither_lo, double dither_hi,
                                                        double cr, double *__restrict__ trial, uint32_t *__restrict__ viol,
                                                        uint32_t *__restrict__ nviol) {
  const int64_t total = (int64_t)P * n * G;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int g = (int)(i % G);
    const int64_t pm = i / G;
    const int c = (int)(pm % n), p = (int)(pm / n);
    const double *pp = pop + (size_t)p * n * G;
    double sc;
    int a, b;
    bool take;
    if (r0) {
      sc = scale[p];
      a = r0[pm];
      b = r1[pm];
      take = cross ? cross[i] != 0 : true;
    } else {
      sc = de_u01(seed, (uint64_t)p, gen, 0xFFFFFFFFull, 0) * (dither_hi - dither_lo) + dither_lo;
      // two distinct members other than c
      a = (int)(de_u01(seed, (uint64_t)p, gen, (uint64_t)c, 1) * (n - 1));
      b = (int)(de_u01(seed, (uint64_t)p, gen, (uint64_t)c, 2) * (n - 2));
      if (a >= n - 1) a = n - 2;
      if (b >= n - 2) b = n - 3;
      if (b >= a) ++b;          // b in [0, n-1) \ {a}
      if (a >= c) ++a;          // skip c
      if (b >= c) ++b;
      const int fill = (int)(de_u01(seed, (uint64_t)p, gen, (uint64_t)c, 3) * G);
      take = g == fill || de_u01(seed, (uint64_t)p, gen, (uint64_t)c, 16 + (uint64_t)g) < cr;
    }
    double v = pp[(size_t)c * G + g];
    if (take) v = __dadd_rn(pp[g], __dmul_rn(sc, __dsub_rn(pp[(size_t)a * G + g], pp[(size_t)b * G + g])));
    const bool bad = (v < 0.0) | (v > 1.0);
    if (bad && !r0) v = de_u01(seed, (uint64_t)p, gen, (uint64_t)c, 0x10000 + (uint64_t)g);  // repaired in place
    trial[i] = v;
    if (viol) {
      viol[i] = (bad && r0) ? 1u : 0u;
      if (bad && r0) atomicAdd(nviol + p, 1u);
    }
  }
}

// _ensure_constraint with the reference's stream: the k-th flagged gene of problem p, in (member, gene) order, takes
// rnd[off[p] + k]. One CTA per problem (an ordered scan over n * G flags).
__global__ void __launch_bounds__(256) de_repair_kernel(double *__restrict__ trial, const uint32_t *__restrict__ viol, int n, int G,
                                                        const double *__restrict__ rnd, const int64_t *__restrict__ off) {
  __shared__ unsigned int part[256];
  __shared__ unsigned int carry;
  const int p = blockIdx.x, tid = threadIdx.x;
  const int64_t len = (int64_t)n * G;
  double *t = trial + (size_t)p * len;
  const uint32_t *f = viol + (size_t)p * len;
  if (tid == 0) carry = 0;
  __syncthreads();
  for (int64_t base = 0; base < len; base += 256) {
    const int64_t i = base + tid;
    const unsigned int mine = (i < len && f[i]) ? 1u : 0u;
    part[tid] = mine;
    __syncthreads();
    for (int o = 1; o < 256; o <<= 1) {
      const unsigned int v = tid >= o ? part[tid - o] : 0u;
      __syncthreads();
      part[tid] += v;
      __syncthreads();
    }
    if (mine) t[i] = rnd[off[p] + carry + part[tid] - 1];
    __syncthreads();
    if (tid == 255) carry += part[255];
    __syncthreads();
  }
}

// perturb_image in embedding space: row (p, m) = left row of pair p with e[int(index)] = (float) value for each gene pair,
// in gene order (a later assignment to the same coordinate wins, like the reference's loop). One warp per row.
__global__ void __launch_bounds__(256) de_perturb_kernel(const double *__restrict__ genes, int P, int n, int G, int64_t gstride,
                                                         const float *__restrict__ left, const int32_t *__restrict__ il,
                                                         const int32_t *__restrict__ ir, int D, const double *__restrict__ lo,
                                                         const double *__restrict__ hi, float *__restrict__ rows,
                                                         int32_t *__restrict__ ia, int32_t *__restrict__ ib) {
  const int lane = threadIdx.x & 31;
  const int64_t nrows = (int64_t)P * n;
  for (int64_t q = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); q < nrows; q += (int64_t)gridDim.x * 8) {
    const int p = (int)(q / n);
    const float4 *src = reinterpret_cast<const float4 *>(left + (size_t)__ldg(il + p) * D);
    float4 *dst = reinterpret_cast<float4 *>(rows + (size_t)q * D);
    for (int i = lane; i < D / 4; i += 32) dst[i] = __ldg(src + i);
    __syncwarp();
    if (lane == 0) {
      const double *gq = genes + (size_t)q * gstride;  // gstride = G, or n_members * G to read member 0 of every problem
      float *row = rows + (size_t)q * D;
      for (int j = 0; j + 1 < G; j += 2) {
        // _scale_parameters (:763-767): 0.5 (lo + hi) + (t - 0.5) |lo - hi|, float64
        const double xi = __dadd_rn(__dmul_rn(0.5, __dadd_rn(lo[j], hi[j])), __dmul_rn(__dsub_rn(gq[j], 0.5), fabs(__dsub_rn(lo[j], hi[j]))));
        const double xv = __dadd_rn(__dmul_rn(0.5, __dadd_rn(lo[j + 1], hi[j + 1])), __dmul_rn(__dsub_rn(gq[j + 1], 0.5), fabs(__dsub_rn(lo[j + 1], hi[j + 1]))));
        int idx = (int)xi;  // xs.astype(int) truncates (attack.py:17)
        if (idx < 0) idx = 0;
        if (idx > D - 1) idx = D - 1;
        row[idx] = (float)xv;
      }
      ia[q] = (int32_t)q;
      ib[q] = __ldg(ir + p);
    }
    __syncwarp();
  }
}

// energies = 1 - P(target class) of the K-learner mean (attack.py:36-41 with minimize = False), fp32 like numpy
__global__ void __launch_bounds__(256) de_energy_kernel(const float *__restrict__ probs, int64_t N, int n, int K, int C,
                                                        const int32_t *__restrict__ targets, float *__restrict__ e) {
  for (int64_t q = (int64_t)blockIdx.x * 256 + threadIdx.x; q < N; q += (int64_t)gridDim.x * 256) {
    const int target = __ldg(targets + q / n);
    float s = 0.0f;
    for (int k = 0; k < K; ++k) {
      float p;
      if (C == 2) p = probs[((size_t)q * K + k) * 2 + target];
      else p = target ? probs[(size_t)q * K + k] : __fsub_rn(1.0f, probs[(size_t)q * K + k]);
      s = __fadd_rn(s, p);
    }
    if (K > 1) s = __fdiv_rn(s, (float)K);
    e[q] = __fsub_rn(1.0f, s);
  }
}

// Selection of one generation (differential_evolution.py:704-714), one CTA per problem: member c takes its trial when
// the trial's energy is LOWER; then slot 0 takes the accepted trial of lowest energy if that beats slot 0 (the first such
// member in order on ties) — what the reference's sequential loop leaves behind.
__global__ void __launch_bounds__(256) de_select_kernel(double *__restrict__ pop, float *__restrict__ e_pop,
                                                        const double *__restrict__ trial, const float *__restrict__ e_trial, int n,
                                                        int G, const uint8_t *__restrict__ active) {
  __shared__ float best_e[256];
  __shared__ int best_c[256];
  const int p = blockIdx.x, tid = threadIdx.x;
  if (active && !active[p]) return;  // the problem's callback asked to stop (attack.py:83-84): its population is frozen
  double *pp = pop + (size_t)p * n * G;
  const double *tt = trial + (size_t)p * n * G;
  float *ep = e_pop + (size_t)p * n;
  const float *et = e_trial + (size_t)p * n;
  float be = INFINITY;
  int bc = -1;
  for (int c = tid; c < n; c += 256) {
    const float en = et[c];
    if (en < ep[c]) {
      ep[c] = en;
      for (int g = 0; g < G; ++g) pp[(size_t)c * G + g] = tt[(size_t)c * G + g];
      if (c > 0 && (en < be)) {  // ascending c within a thread: the first minimum stays
        be = en;
        bc = c;
      }
    }
  }
  best_e[tid] = be;
  best_c[tid] = bc;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (tid < o) {
      const float e2 = best_e[tid + o];
      const int c2 = best_c[tid + o];
      if (c2 >= 0 && (best_c[tid] < 0 || e2 < best_e[tid] || (e2 == best_e[tid] && c2 < best_c[tid]))) {
        best_e[tid] = e2;
        best_c[tid] = c2;
      }
    }
    __syncthreads();
  }
  const int c = best_c[0];
  if (c > 0 && best_e[0] < ep[0]) {  // ep[0] already holds member 0's own update
    __syncthreads();
    for (int g = tid; g < G; g += 256) pp[g] = tt[(size_t)c * G + g];
    if (tid == 0) ep[0] = best_e[0];
  }
}

// first evaluation (differential_evolution.py:630-668): the best member swaps into slot 0
__global__ void __launch_bounds__(256) de_best_first_kernel(double *__restrict__ pop, float *__restrict__ e_pop, int n, int G) {
  __shared__ float best_e[256];
  __shared__ int best_c[256];
  const int p = blockIdx.x, tid = threadIdx.x;
  double *pp = pop + (size_t)p * n * G;
  float *ep = e_pop + (size_t)p * n;
  float be = INFINITY;
  int bc = -1;
  for (int c = tid; c < n; c += 256)
    if (bc < 0 || ep[c] < be) {
      be = ep[c];
      bc = c;
    }
  best_e[tid] = be;
  best_c[tid] = bc;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (tid < o) {
      const float e2 = best_e[tid + o];
      const int c2 = best_c[tid + o];
      if (c2 >= 0 && (best_c[tid] < 0 || e2 < best_e[tid] || (e2 == best_e[tid] && c2 < best_c[tid]))) {  // np.argmin: first
        best_e[tid] = e2;
        best_c[tid] = c2;
      }
    }
    __syncthreads();
  }
  const int c = best_c[0];
  if (c > 0) {
    for (int g = tid; g < G; g += 256) {
      const double a = pp[g], b = pp[(size_t)c * G + g];
      pp[g] = b;
      pp[(size_t)c * G + g] = a;
    }
    if (tid == 0) {
      const float a = ep[0];
      ep[0] = ep[c];
      ep[c] = a;
    }
  }
}
