// topk.cuh — deterministic top-B selection on packed 64-bit keys (HBM/L2-bound streaming passes).
//
// Replaces modAL's multi_argmax = np.argpartition(-v, n-1)[:n] (/root/reference/code/uncertainty.py:155,
// 183,213) with the north_star tie rule: order = (utility descending, pair index ascending). A key is
// (orderable(utility) << 32) | (0xFFFFFFFF - index) (include/alink_math.h), so all keys are distinct
// and plain descending u64 order IS the rule. Algorithm:
//   1. radix-select the exact B-th largest key: 6 histogram passes (11/11/10 bits of the utility,
//      11/11/10 bits of the index part). Each pass = shared-memory histogram per CTA -> global
//      histogram -> the last CTA to finish (ticket) scans it and narrows (prefix, remaining).
//   2. compact every key >= that threshold (exactly B keys) with one atomic slot counter.
//   3. bitonic-sort the B keys descending (one CTA in shared memory up to 4096, global passes above).
//   4. unpack (index, utility).
// The key source is either a utility vector + base index (keys built on the fly: 4 B/pair per pass)
// or raw keys (the cross-shard merge after the all-gather).
#pragma once
#include "common.cuh"
#include <cooperative_groups.h>

struct KeySrc {
  const float *util;
  const uint64_t *keys;
  uint64_t base;
  const uint32_t *map;  // optional: index reported for element i is base + map[i] (refine.cuh)
};

struct SelectState {
  unsigned long long prefix;     // high bits decided so far (value of key >> hi_shift)
  unsigned long long remaining;  // rank still to find inside the prefix bucket (1-based)
  unsigned int ticket;
  unsigned int out_count;
  unsigned int hist[2048];
};

__device__ __forceinline__ uint64_t load_key(const KeySrc &s, int64_t i) {
  if (s.keys) return s.keys[i];
  return alink_pack_key(__ldg(s.util + i), s.base + (s.map ? (uint64_t)__ldg(s.map + i) : (uint64_t)i));
}

__global__ void select_init_kernel(SelectState *st, unsigned long long B) {
  for (int i = threadIdx.x; i < 2048; i += blockDim.x) st->hist[i] = 0;
  if (threadIdx.x == 0) {
    st->prefix = 0;
    st->remaining = B;
    st->ticket = 0;
    st->out_count = 0;
  }
}

// One radix pass over `bits` bits at `shift`. first != 0: no prefix filter.
__global__ void __launch_bounds__(256) radix_pass_kernel(KeySrc src, int64_t n, int shift, int bits, int first,
                                                         SelectState *st) {
  __shared__ unsigned int sh[2048];
  __shared__ unsigned int part[256];
  __shared__ bool is_last;
  const int nb = 1 << bits;
  const unsigned mask = (unsigned)nb - 1u;
  for (int i = threadIdx.x; i < nb; i += 256) sh[i] = 0;
  __syncthreads();
  const unsigned long long prefix = st->prefix;
  const int hi = shift + bits;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
    const uint64_t key = load_key(src, i);
    if (first || (key >> hi) == prefix) atomicAdd(&sh[(unsigned)(key >> shift) & mask], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nb; i += 256)
    if (sh[i]) atomicAdd(&st->hist[i], sh[i]);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = (atomicAdd(&st->ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  // ---- last CTA: descending scan of the global histogram
  for (int i = threadIdx.x; i < nb; i += 256) sh[i] = __ldcg(&st->hist[i]);
  __syncthreads();
  const int per = nb / 256;  // 8 or 4 bins per thread; thread t owns bins [nb-1-per*t-(per-1), nb-1-per*t]
  unsigned int mine = 0;
  for (int j = 0; j < per; ++j) mine += sh[nb - 1 - per * threadIdx.x - j];
  part[threadIdx.x] = mine;
  const unsigned long long remaining = st->remaining;  // read before anyone may overwrite it below
  __syncthreads();
  for (int off = 1; off < 256; off <<= 1) {  // inclusive Hillis-Steele scan
    unsigned int v = threadIdx.x >= off ? part[threadIdx.x - off] : 0u;
    __syncthreads();
    part[threadIdx.x] += v;
    __syncthreads();
  }
  const unsigned long long incl = part[threadIdx.x];
  const unsigned long long before = incl - mine;
  if (before < remaining && remaining <= incl) {
    unsigned long long cum = before;
    for (int j = 0; j < per; ++j) {
      const int bin = nb - 1 - per * threadIdx.x - j;
      const unsigned int c = sh[bin];
      if (cum + c >= remaining) {
        st->prefix = first ? (unsigned long long)bin : ((prefix << bits) | (unsigned long long)bin);
        st->remaining = remaining - cum;
        break;
      }
      cum += c;
    }
  }
  for (int i = threadIdx.x; i < 2048; i += 256) st->hist[i] = 0;
  if (threadIdx.x == 0) st->ticket = 0;
}

// The bin-picking tail of radix_pass_kernel on its own: for histograms that were summed ACROSS RANKS first (ncclAllReduce;
// the sharded A-LINK rule, disparity.cuh). One CTA; narrows (prefix, remaining) in `st` and clears the histogram.
__global__ void __launch_bounds__(256) select_pick_kernel(unsigned int *__restrict__ hist, int bits, int first, SelectState *st) {
  __shared__ unsigned int sh[2048];
  __shared__ unsigned int part[256];
  const int nb = 1 << bits;
  for (int i = threadIdx.x; i < nb; i += 256) sh[i] = hist[i];
  __syncthreads();
  const int per = nb / 256;
  unsigned int mine = 0;
  for (int j = 0; j < per; ++j) mine += sh[nb - 1 - per * threadIdx.x - j];
  part[threadIdx.x] = mine;
  const unsigned long long remaining = st->remaining, prefix = st->prefix;
  __syncthreads();
  for (int off = 1; off < 256; off <<= 1) {
    unsigned int v = threadIdx.x >= off ? part[threadIdx.x - off] : 0u;
    __syncthreads();
    part[threadIdx.x] += v;
    __syncthreads();
  }
  const unsigned long long incl = part[threadIdx.x], before = incl - mine;
  if (before < remaining && remaining <= incl) {
    unsigned long long cum = before;
    for (int j = 0; j < per; ++j) {
      const int bin = nb - 1 - per * threadIdx.x - j;
      const unsigned int c = sh[bin];
      if (cum + c >= remaining) {
        st->prefix = first ? (unsigned long long)bin : ((prefix << bits) | (unsigned long long)bin);
        st->remaining = remaining - cum;
        break;
      }
      cum += c;
    }
  }
  for (int i = threadIdx.x; i < 2048; i += 256) hist[i] = 0;
}

__device__ __forceinline__ void stage1_pick(const unsigned int *__restrict__ ghist, int nb, unsigned long long &remaining,
                                            unsigned int &bin_out, unsigned int *sh, unsigned int *part) {
  // descending scan of a global histogram by the whole CTA (256 threads); same walk as radix_pass_kernel's last CTA
  const int tid = threadIdx.x;
  for (int i = tid; i < nb; i += 256) sh[i] = __ldcg(ghist + i);
  __syncthreads();
  const int per = nb / 256;
  unsigned int mine = 0;
  for (int j = 0; j < per; ++j) mine += sh[nb - 1 - per * tid - j];
  part[tid] = mine;
  __syncthreads();
  for (int off = 1; off < 256; off <<= 1) {
    const unsigned int v = tid >= off ? part[tid - off] : 0u;
    __syncthreads();
    part[tid] += v;
    __syncthreads();
  }
  __shared__ unsigned int s_bin;
  __shared__ unsigned long long s_rem;
  const unsigned long long incl = part[tid], before = incl - mine;
  if (before < remaining && remaining <= incl) {
    unsigned long long cum = before;
    for (int j = 0; j < per; ++j) {
      const int bin = nb - 1 - per * tid - j;
      const unsigned int c = sh[bin];
      if (cum + c >= remaining) {
        s_bin = (unsigned int)bin;
        s_rem = remaining - cum;
        break;
      }
      cum += c;
    }
  }
  __syncthreads();
  bin_out = s_bin;
  remaining = s_rem;
  __syncthreads();
}


// The whole select of a LARGE source in ONE cooperative launch: six histogram passes over the packed keys with a grid-wide
// barrier after each (every CTA scans the global histogram redundantly: nothing to broadcast), then the compaction of the
// keys >= the B-th largest. Replaces select_init + 6 radix passes + compact (8 launches, ~10 us of launch + tail latency
// each for a 4 MB source). hist: 6 x 2048 zeroed counters; st->out_count zeroed.
__global__ void __launch_bounds__(256) topk_select_coop_kernel(KeySrc src, int64_t n, unsigned long long B, unsigned int *__restrict__ hist,
                                                               SelectState *st, uint64_t *__restrict__ out, unsigned int cap) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  __shared__ unsigned int sh[2048];
  __shared__ unsigned int part[256];
  const int tid = threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * 256;
  const int shifts[6] = {53, 42, 32, 21, 10, 0}, nbits[6] = {11, 11, 10, 11, 11, 10};
  unsigned long long remaining = B, prefix = 0;
#pragma unroll 1
  for (int p = 0; p < 6; ++p) {
    const int nb = 1 << nbits[p], hi = shifts[p] + nbits[p];
    for (int i = tid; i < nb; i += 256) sh[i] = 0;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * 256 + tid; i < n; i += stride) {
      const uint64_t key = load_key(src, i);
      if (p == 0 || (key >> hi) == prefix) atomicAdd(&sh[(unsigned)(key >> shifts[p]) & (unsigned)(nb - 1)], 1u);
    }
    __syncthreads();
    unsigned int *gh = hist + p * 2048;
    for (int i = tid; i < nb; i += 256)
      if (sh[i]) atomicAdd(gh + i, sh[i]);
    __threadfence();
    grid.sync();
    unsigned int bin;
    stage1_pick(gh, nb, remaining, bin, sh, part);
    prefix = p == 0 ? (unsigned long long)bin : ((prefix << nbits[p]) | (unsigned long long)bin);
  }
  for (int64_t i = (int64_t)blockIdx.x * 256 + tid; i < n; i += stride) {
    const uint64_t key = load_key(src, i);
    if (key >= prefix) {
      const unsigned int pos = atomicAdd(&st->out_count, 1u);
      if (pos < cap) out[pos] = key;
    }
  }
  if (blockIdx.x == 0 && tid == 0) st->prefix = prefix;
}

// Small sources (the exact re-score's candidates, the cross-shard merge of B keys per rank): the same passes
// inside ONE CTA — no global histogram, no ticket, one launch instead of seven. Each multi-CTA pass costs
// ~10 us of launch + tail-scan latency whatever n is, which is all a 10 k-key select consists of.
__device__ __forceinline__ void radix_select_cta(const KeySrc &src, int n, unsigned long long B, int passes,
                                                 unsigned int *sh /*2048*/, unsigned int *part /*256*/,
                                                 unsigned long long &s_prefix, unsigned long long &s_remaining) {
  const int tid = threadIdx.x;
  if (tid == 0) {
    s_prefix = 0;
    s_remaining = B;
  }
  const int shifts[6] = {53, 42, 32, 21, 10, 0};
  const int nbits[6] = {11, 11, 10, 11, 11, 10};
  for (int p = 0; p < passes; ++p) {
    const int shift = shifts[p], bits = nbits[p], nb = 1 << bits, hi = shift + bits;
    const unsigned mask = (unsigned)nb - 1u;
    for (int i = tid; i < nb; i += 1024) sh[i] = 0;
    __syncthreads();
    const unsigned long long prefix = s_prefix, remaining = s_remaining;
    for (int i = tid; i < n; i += 1024) {
      const uint64_t key = load_key(src, i);
      if (p == 0 || (key >> hi) == prefix) atomicAdd(&sh[(unsigned)(key >> shift) & mask], 1u);
    }
    __syncthreads();
    const int per = nb / 256;  // thread t < 256 owns bins [nb-1-per*t-(per-1), nb-1-per*t], descending
    unsigned int mine = 0;
    if (tid < 256) {
      for (int j = 0; j < per; ++j) mine += sh[nb - 1 - per * tid - j];
      part[tid] = mine;
    }
    __syncthreads();
    for (int off = 1; off < 256; off <<= 1) {
      unsigned int v = (tid < 256 && tid >= off) ? part[tid - off] : 0u;
      __syncthreads();
      if (tid < 256) part[tid] += v;
      __syncthreads();
    }
    if (tid < 256) {
      const unsigned long long incl = part[tid], before = incl - mine;
      if (before < remaining && remaining <= incl) {
        unsigned long long cum = before;
        for (int j = 0; j < per; ++j) {
          const int bin = nb - 1 - per * tid - j;
          const unsigned int c = sh[bin];
          if (cum + c >= remaining) {
            s_prefix = p == 0 ? (unsigned long long)bin : ((prefix << bits) | (unsigned long long)bin);
            s_remaining = remaining - cum;
            break;
          }
          cum += c;
        }
      }
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(1024) radix_select_small_kernel(KeySrc src, int n, unsigned long long B, int passes,
                                                                  SelectState *st) {
  __shared__ unsigned int sh[2048];
  __shared__ unsigned int part[256];
  __shared__ unsigned long long s_prefix, s_remaining;
  radix_select_cta(src, n, B, passes, sh, part, s_prefix, s_remaining);
  if (threadIdx.x == 0) {
    st->prefix = s_prefix;
    st->remaining = s_remaining;
    st->ticket = 0;
    st->out_count = 0;
  }
}

// Whole top-B of a small source in one launch: select the B-th key, gather the B keys >= it into shared
// memory, bitonic-sort them descending, write them out (B <= npad <= 4096).
// n_dev: the source length lives on the device (FP32_REFINE's candidate count); n is then its capacity
// (the body as a device function: the peer-push merge of comm.cuh runs it in the tail of its own kernel; 1024 threads)
__device__ __forceinline__ void topk_small_body(const KeySrc &src, int n, int B, int npad, uint64_t *__restrict__ out) {
  __shared__ unsigned int sh[2048];
  __shared__ unsigned int part[256];
  __shared__ unsigned long long s_prefix, s_remaining;
  __shared__ unsigned int s_count;
  __shared__ uint64_t sk[4096];
  const int tid = threadIdx.x;
  if (tid == 0) s_count = 0;
  radix_select_cta(src, n, (unsigned long long)B, 6, sh, part, s_prefix, s_remaining);
  const unsigned long long thr = s_prefix;
  for (int i = tid; i < npad; i += 1024) sk[i] = 0ull;
  __syncthreads();
  for (int i = tid; i < n; i += 1024) {
    const uint64_t key = load_key(src, i);
    if (key >= thr) {
      const unsigned int pos = atomicAdd(&s_count, 1u);
      if (pos < (unsigned)B) sk[pos] = key;
    }
  }
  __syncthreads();
  for (int k = 2; k <= npad; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < npad; i += 1024) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const uint64_t a = sk[i], b = sk[ixj];
          const bool desc = (i & k) == 0;
          if (desc ? (a < b) : (a > b)) {
            sk[i] = b;
            sk[ixj] = a;
          }
        }
      }
      __syncthreads();
    }
  }
  for (int i = tid; i < B; i += 1024) out[i] = sk[i];
}

__global__ void __launch_bounds__(1024) topk_small_kernel(KeySrc src, int n, int B, int npad, uint64_t *__restrict__ out,
                                                          const unsigned int *__restrict__ n_dev = nullptr) {
  if (n_dev) {
    const unsigned int nd = *n_dev;
    if (nd < (unsigned int)n) n = (int)nd;
  }
  topk_small_body(src, n, B, npad, out);
}

// keys >= threshold (st->prefix after the last pass) -> out (exactly B of them, unordered)
__global__ void __launch_bounds__(256) compact_kernel(KeySrc src, int64_t n, SelectState *st, uint64_t *out,
                                                      unsigned int cap) {
  const unsigned long long thr = st->prefix;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
    const uint64_t key = load_key(src, i);
    if (key >= thr) {
      const unsigned int pos = atomicAdd(&st->out_count, 1u);
      if (pos < cap) out[pos] = key;
    }
  }
}

// one-CTA bitonic sort, descending; npad = power of two <= 4096; entries >= n are padded with 0
// (0 is below every real key: index 0xFFFFFFFF is excluded by the ABI).
__global__ void __launch_bounds__(1024) bitonic_smem_kernel(uint64_t *keys, int n, int npad) {
  extern __shared__ uint64_t sk[];
  for (int i = threadIdx.x; i < npad; i += blockDim.x) sk[i] = i < n ? keys[i] : 0ull;
  __syncthreads();
  for (int k = 2; k <= npad; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < npad; i += blockDim.x) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const uint64_t a = sk[i], b = sk[ixj];
          const bool desc = (i & k) == 0;
          if (desc ? (a < b) : (a > b)) {
            sk[i] = b;
            sk[ixj] = a;
          }
        }
      }
      __syncthreads();
    }
  }
  for (int i = threadIdx.x; i < n; i += blockDim.x) keys[i] = sk[i];
}

__global__ void pad_zero_kernel(uint64_t *keys, int64_t n, int64_t npad) {
  for (int64_t i = n + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < npad;
       i += (int64_t)gridDim.x * blockDim.x)
    keys[i] = 0ull;
}

__global__ void bitonic_global_step_kernel(uint64_t *keys, int64_t npad, int64_t k, int64_t j) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < npad;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t ixj = i ^ j;
    if (ixj > i) {
      const uint64_t a = keys[i], b = keys[ixj];
      const bool desc = (i & k) == 0;
      if (desc ? (a < b) : (a > b)) {
        keys[i] = b;
        keys[ixj] = a;
      }
    }
  }
}

__global__ void unpack_keys_kernel(const uint64_t *keys, int64_t B, int64_t *idx, float *util) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < B; i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t k = keys[i];
    if (idx) idx[i] = (int64_t)alink_key_index(k);
    if (util) util[i] = alink_key_utility(k);
  }
}

static inline int64_t next_pow2(int64_t v) {
  int64_t p = 1;
  while (p < v) p <<= 1;
  return p;
}

static const int64_t SELECT_SMALL_N = 32768;  // below this one CTA runs every pass (radix_select_small_kernel)

// Radix-select the B-th largest key; on return st->prefix holds it (device). passes = 3 stops after the 32
// utility bits: st->prefix is then key >> 32 of the B-th largest key (its utility, ties unresolved) — all the
// refine filter needs (refine.cuh), at half the streaming passes.
static int select_threshold(alink_ctx *ctx, const KeySrc &src, int64_t n, int64_t B, SelectState *st, int passes = 6) {
  if (n <= SELECT_SMALL_N) {
    radix_select_small_kernel<<<1, 1024, 0, ctx->stream>>>(src, (int)n, (unsigned long long)B, passes, st);
    ALINK_LAUNCHED(ctx);
    return ALINK_OK;
  }
  select_init_kernel<<<1, 256, 0, ctx->stream>>>(st, (unsigned long long)B);
  ALINK_LAUNCHED(ctx);
  int64_t blocks = (n + 256 * 8 - 1) / (256 * 8);
  const int64_t cap = (int64_t)ctx->sm_count * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  static const int shifts[6] = {53, 42, 32, 21, 10, 0};
  static const int nbits[6] = {11, 11, 10, 11, 11, 10};
  for (int p = 0; p < passes; ++p) {
    radix_pass_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(src, n, shifts[p], nbits[p], p == 0, st);
    ALINK_LAUNCHED(ctx);
  }
  return ALINK_OK;
}

// Top-B keys of the source, sorted descending, into ctx->sel_keys (device). B >= 1, B <= n.
static int topk_device(alink_ctx *ctx, const KeySrc &src, int64_t n, int64_t B) {
  ALINK_TRY(dev_reserve(ctx, ctx->sel_state, sizeof(SelectState)));
  const int64_t npad = next_pow2(B);
  ALINK_TRY(dev_reserve(ctx, ctx->sel_keys, sizeof(uint64_t) * (size_t)npad));
  SelectState *st = (SelectState *)ctx->sel_state.p;
  uint64_t *out = (uint64_t *)ctx->sel_keys.p;
  if (n <= SELECT_SMALL_N && npad <= 4096) {
    topk_small_kernel<<<1, 1024, 0, ctx->stream>>>(src, (int)n, (int)B, (int)(npad < 2 ? 2 : npad), out);
    ALINK_LAUNCHED(ctx);
    return ALINK_OK;
  }
  int64_t blocks = (n + 256 * 8 - 1) / (256 * 8);
  const int64_t cap = (int64_t)ctx->sm_count * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  static const int coop_env = [] { const char *e = getenv("ALINK_TOPK_COOP"); return e ? atoi(e) : 1; }();
  static int coop_per_sm = -1;
  if (coop_per_sm < 0) {
    int occ = 0, coop = 0;
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, ctx->device);
    coop_per_sm = (coop && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, topk_select_coop_kernel, 256, 0) == cudaSuccess && occ > 0)
                      ? (occ < 4 ? occ : 4) : 0;
  }
  if (coop_env && coop_per_sm > 0) {
    ALINK_TRY(dev_reserve(ctx, ctx->sel_hist, sizeof(unsigned int) * 6 * 2048));
    ALINK_CUDA(ctx, cudaMemsetAsync(ctx->sel_hist.p, 0, sizeof(unsigned int) * 6 * 2048, ctx->stream));
    ALINK_CUDA(ctx, cudaMemsetAsync(&st->out_count, 0, sizeof(unsigned int), ctx->stream));
    int64_t cb = blocks < (int64_t)ctx->sm_count * coop_per_sm ? blocks : (int64_t)ctx->sm_count * coop_per_sm;
    unsigned long long Bu = (unsigned long long)B;
    unsigned int *hist = (unsigned int *)ctx->sel_hist.p;
    unsigned int capu = (unsigned int)B;
    KeySrc srcv = src;
    int64_t nv = n;
    void *kargs[] = {(void *)&srcv, (void *)&nv, (void *)&Bu, (void *)&hist, (void *)&st, (void *)&out, (void *)&capu};
    ALINK_CUDA(ctx, cudaLaunchCooperativeKernel((const void *)topk_select_coop_kernel, dim3((unsigned)cb), dim3(256), kargs, 0, ctx->stream));
    ALINK_LAUNCHED(ctx);
  } else {
    ALINK_TRY(select_threshold(ctx, src, n, B, st));
    compact_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(src, n, st, out, (unsigned int)B);
    ALINK_LAUNCHED(ctx);
  }
  if (npad <= 4096) {
    const int threads = npad >= 2048 ? 1024 : (npad >= 64 ? (int)(npad / 2) : 32);
    bitonic_smem_kernel<<<1, threads, sizeof(uint64_t) * (size_t)npad, ctx->stream>>>(out, (int)B, (int)npad);
    ALINK_LAUNCHED(ctx);
  } else {
    if (npad > B) {
      pad_zero_kernel<<<64, 256, 0, ctx->stream>>>(out, B, npad);
      ALINK_LAUNCHED(ctx);
    }
    int64_t gb = npad / 256;
    if (gb > cap) gb = cap;
    for (int64_t k = 2; k <= npad; k <<= 1)
      for (int64_t j = k >> 1; j > 0; j >>= 1) {
        bitonic_global_step_kernel<<<(unsigned)gb, 256, 0, ctx->stream>>>(out, npad, k, j);
        ALINK_LAUNCHED(ctx);
      }
  }
  return ALINK_OK;
}
