// disparity.cuh — A-LINK's own selector on column-extracted scores (HBM-bound streaming kernels).
//
// Replaces the Python double loop + sets.Set of /root/reference/code/ALINK.py:170-201
// (same rule in ALINK_arc.py:167-198, ALINK_MTP.py:178-209):
//   per noise t : d_j = -|p2[t][j] - p1[j]|; keep argsort(d)[:int(N*ratio)]   (ALINK.py:181-183)
//                 (blind strategy: keep j whose 0.5-decisions differ,          ALINK.py:177-179)
//   intersect over t                                                           (ALINK.py:186-189)
//   grey zone   : keep j with p1[j] <= .5-eps or >= .5+eps, each one query     (ALINK.py:195-198)
//   agreement   : keep j with (p1[j] >= .5) == (y[j] >= .5)                    (ALINK.py:196-200)
// The per-noise argsort prefix is a global order statistic: radix-select (topk.cuh) gives the
// threshold key of |p2-p1| with the lowest-index tie rule, membership is key >= threshold.
// Output = ascending pair indices via a 3-kernel ordered compaction (count / scan / scatter).
#pragma once
#include "common.cuh"
#include "topk.cuh"

__global__ void __launch_bounds__(256) absdiff_vec_kernel(const float *__restrict__ q, const float *__restrict__ p1,
                                                          int64_t N, float *__restrict__ d) {
  for (int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x; j < N; j += (int64_t)gridDim.x * 256)
    d[j] = fabsf(__fsub_rn(__ldg(q + j), __ldg(p1 + j)));
}

__global__ void __launch_bounds__(256) disparity_mark_kernel(const float *__restrict__ d, int64_t N,
                                                             const SelectState *st, int has_thr,
                                                             int *__restrict__ cnt) {
  const unsigned long long thr = st->prefix;
  for (int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x; j < N; j += (int64_t)gridDim.x * 256)
    if (has_thr && alink_pack_key(d[j], (uint64_t)j) >= thr) cnt[j] += 1;
}

// sharded pools (alink_b200/dist.py): membership against an explicit GLOBAL threshold key, keys carry base + j
__global__ void __launch_bounds__(256) disparity_mark_thr_kernel(const float *__restrict__ q, const float *__restrict__ p1,
                                                                 int64_t N, uint64_t base, unsigned long long thr,
                                                                 int *__restrict__ cnt, const unsigned long long *thr_dev = nullptr) {
  if (thr_dev) thr = *thr_dev;  // the threshold was selected on the device (alink_disparity_select_comm)
  if (thr == 0ull) return;      // 0 = nothing passes
  for (int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x; j < N; j += (int64_t)gridDim.x * 256)
    if (alink_pack_key(fabsf(__fsub_rn(__ldg(q + j), __ldg(p1 + j))), base + (uint64_t)j) >= thr) cnt[j] += 1;
}

// one radix-select histogram pass over this shard's keys (|q - p1|, base + j): the ranks sum the
// histograms (all-reduce) and pick the bin on the host, so the threshold is the global order statistic
__global__ void __launch_bounds__(256) disparity_hist_kernel(const float *__restrict__ q, const float *__restrict__ p1,
                                                             int64_t N, uint64_t base, unsigned long long prefix, int shift,
                                                             int bits, int first, unsigned int *__restrict__ hist,
                                                             const SelectState *st = nullptr) {
  __shared__ unsigned int sh[2048];
  if (st) prefix = st->prefix;  // device-resident select state (the bins are picked on the device after the all-reduce)
  const int nb = 1 << bits;
  const unsigned mask = (unsigned)nb - 1u;
  for (int i = threadIdx.x; i < nb; i += 256) sh[i] = 0;
  __syncthreads();
  const int hi = shift + bits;
  for (int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x; j < N; j += (int64_t)gridDim.x * 256) {
    const uint64_t key = alink_pack_key(fabsf(__fsub_rn(__ldg(q + j), __ldg(p1 + j))), base + (uint64_t)j);
    if (first || (key >> hi) == prefix) atomicAdd(&sh[(unsigned)(key >> shift) & mask], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nb; i += 256)
    if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

__global__ void __launch_bounds__(256) disparity_blind_kernel(const float *__restrict__ q,
                                                              const float *__restrict__ p1, int64_t N,
                                                              int *__restrict__ cnt) {
  for (int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x; j < N; j += (int64_t)gridDim.x * 256)
    if ((__ldg(q + j) >= 0.5f) != (__ldg(p1 + j) >= 0.5f)) cnt[j] += 1;
}

struct DispFilter {
  const int *cnt;
  const float *p1;
  const float *y;
  int T;
  float lo, hi;
};

// flags: bit0 = queried (counts toward ACTIVE_COUNT), bit1 = selected
__device__ __forceinline__ int disparity_flags(const DispFilter &f, int64_t j) {
  if (f.cnt[j] != f.T) return 0;
  const float p = __ldg(f.p1 + j);
  if (!(p <= f.lo || p >= f.hi)) return 0;
  int r = 1;
  if (!f.y || ((p >= 0.5f) == (__ldg(f.y + j) >= 0.5f))) r |= 2;
  return r;
}

constexpr int DISP_ITEMS = 4;  // items per thread; one CTA owns 1024 consecutive pairs

__global__ void __launch_bounds__(256) disparity_count_kernel(DispFilter f, int64_t N, unsigned int *blk_sel,
                                                              unsigned long long *active) {
  __shared__ unsigned int s_sel, s_act;
  if (threadIdx.x == 0) s_sel = s_act = 0;
  __syncthreads();
  const int64_t j0 = ((int64_t)blockIdx.x * 256 + threadIdx.x) * DISP_ITEMS;
  unsigned int sel = 0, act = 0;
  for (int i = 0; i < DISP_ITEMS; ++i)
    if (j0 + i < N) {
      const int fl = disparity_flags(f, j0 + i);
      act += fl & 1;
      sel += (fl >> 1) & 1;
    }
  if (sel) atomicAdd(&s_sel, sel);
  if (act) atomicAdd(&s_act, act);
  __syncthreads();
  if (threadIdx.x == 0) {
    blk_sel[blockIdx.x] = s_sel;
    if (s_act) atomicAdd(active, (unsigned long long)s_act);
  }
}

// single CTA: exclusive scan of the per-block counts (in place); total -> *n_sel
__global__ void __launch_bounds__(1024) disparity_scan_kernel(unsigned int *blk_sel, int nblk,
                                                              unsigned long long *n_sel) {
  __shared__ unsigned long long carry;
  __shared__ unsigned int part[1024];
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < nblk; base += 1024) {
    const int i = base + threadIdx.x;
    const unsigned int v = i < nblk ? blk_sel[i] : 0u;
    part[threadIdx.x] = v;
    __syncthreads();
    for (int off = 1; off < 1024; off <<= 1) {
      unsigned int t = threadIdx.x >= off ? part[threadIdx.x - off] : 0u;
      __syncthreads();
      part[threadIdx.x] += t;
      __syncthreads();
    }
    if (i < nblk) blk_sel[i] = (unsigned int)(carry + part[threadIdx.x] - v);
    __syncthreads();
    if (threadIdx.x == 1023) carry += part[1023];
    __syncthreads();
  }
  if (threadIdx.x == 0) *n_sel = carry;
}

__global__ void __launch_bounds__(256) disparity_scatter_kernel(DispFilter f, int64_t N,
                                                                const unsigned int *blk_off, int64_t *idx_out) {
  __shared__ unsigned int part[256];
  const int64_t j0 = ((int64_t)blockIdx.x * 256 + threadIdx.x) * DISP_ITEMS;
  int fl[DISP_ITEMS];
  unsigned int mine = 0;
  for (int i = 0; i < DISP_ITEMS; ++i) {
    fl[i] = (j0 + i < N) ? ((disparity_flags(f, j0 + i) >> 1) & 1) : 0;
    mine += fl[i];
  }
  part[threadIdx.x] = mine;
  __syncthreads();
  for (int off = 1; off < 256; off <<= 1) {
    unsigned int t = threadIdx.x >= off ? part[threadIdx.x - off] : 0u;
    __syncthreads();
    part[threadIdx.x] += t;
    __syncthreads();
  }
  unsigned int pos = blk_off[blockIdx.x] + part[threadIdx.x] - mine;
  for (int i = 0; i < DISP_ITEMS; ++i)
    if (fl[i]) idx_out[pos++] = j0 + i;
}

// idx[i] += base for the first *n entries (shard-local -> global pair indices)
__global__ void __launch_bounds__(256) offset_i64_kernel(int64_t *idx, int64_t cap, const unsigned long long *n, int64_t base) {
  const int64_t m = (int64_t)*n < cap ? (int64_t)*n : cap;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < m; i += (int64_t)gridDim.x * 256) idx[i] += base;
}
