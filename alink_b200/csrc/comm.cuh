// comm.cuh — the cross-shard exchange behind the C ABI: NCCL over NVLink, bound at run time.
//
// The reference is single-process; the pool shards by pair index over the GPUs of one box and the only exchange of the
// query path is one all-gather of B packed keys per rank. Round 1 did it from Python (torch.distributed); a host that
// binds include/alink_b200.h directly had no merge. libnccl.so.2 is dlopen'ed on first use — by soname, so a process that
// already loaded NCCL (PyTorch) shares its copy — and a library without NCCL still serves one GPU.
#pragma once
#include "common.cuh"
#include <dlfcn.h>

typedef struct { char internal[128]; } alink_nccl_uid;  // ncclUniqueId
typedef void *alink_nccl_comm;                          // ncclComm_t

struct NcclApi {
  void *lib = nullptr;
  int (*GetUniqueId)(alink_nccl_uid *) = nullptr;
  int (*CommInitRank)(alink_nccl_comm *, int, alink_nccl_uid, int) = nullptr;
  int (*CommDestroy)(alink_nccl_comm) = nullptr;
  int (*AllGather)(const void *, void *, size_t, int, alink_nccl_comm, cudaStream_t) = nullptr;
  int (*AllReduce)(const void *, void *, size_t, int, int, alink_nccl_comm, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(int) = nullptr;
  int (*GetVersion)(int *) = nullptr;
};
enum { ALINK_NCCL_UINT32 = 3, ALINK_NCCL_UINT64 = 5, ALINK_NCCL_SUM = 0 };  // ncclDataType_t / ncclRedOp_t values (nccl.h)

static NcclApi *nccl_api() {
  static NcclApi api;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (h) {
      api.lib = h;
      api.GetUniqueId = (int (*)(alink_nccl_uid *))dlsym(h, "ncclGetUniqueId");
      api.CommInitRank = (int (*)(alink_nccl_comm *, int, alink_nccl_uid, int))dlsym(h, "ncclCommInitRank");
      api.CommDestroy = (int (*)(alink_nccl_comm))dlsym(h, "ncclCommDestroy");
      api.AllGather = (int (*)(const void *, void *, size_t, int, alink_nccl_comm, cudaStream_t))dlsym(h, "ncclAllGather");
      api.AllReduce = (int (*)(const void *, void *, size_t, int, int, alink_nccl_comm, cudaStream_t))dlsym(h, "ncclAllReduce");
      api.GetErrorString = (const char *(*)(int))dlsym(h, "ncclGetErrorString");
      api.GetVersion = (int (*)(int *))dlsym(h, "ncclGetVersion");
      if (!api.GetUniqueId || !api.CommInitRank || !api.CommDestroy || !api.AllGather || !api.AllReduce) api.lib = nullptr;
    }
  }
  return api.lib ? &api : nullptr;
}

#define ALINK_NCCL(ctx, api, expr)                                                                          \
  do {                                                                                                      \
    int r__ = (expr);                                                                                       \
    if (r__ != 0)                                                                                           \
      return alink_fail((ctx), ALINK_E_CUDA, "%s:%d %s -> NCCL error %d (%s)", __FILE__, __LINE__, #expr, r__, \
                        (api)->GetErrorString ? (api)->GetErrorString(r__) : "?");                          \
  } while (0)

// ------------------------------------------------------------------------------- the verdict that rides with the keys
// A sharded FP32_REFINE step does not stop for its verification (refine_verify_kernel leaves the verdict on the device):
// word B of every rank's all-gather payload carries it (2 = ok, 1 = not ok), every rank reduces the W words to the same
// all_ok after the gather, and zeroes them so that the merge sees key 0 ("nothing") there.
__global__ void comm_flag_kernel(uint64_t *flag, const unsigned int *ok_dev) {
  if (threadIdx.x == 0) *flag = ok_dev ? (uint64_t)(*ok_dev & 1u) + 1ull : 2ull;
}
__global__ void comm_flags_reduce_kernel(uint64_t *all, int W, int stride, int B, unsigned int *status) {
  bool ok = true;
  for (int r = threadIdx.x; r < W; r += 32) {  // (one warp)
    ok = ok && all[(size_t)r * stride + B] == 2ull;
    all[(size_t)r * stride + B] = 0ull;
  }
  const bool all_ok = __all_sync(0xffffffffu, ok);
  if (threadIdx.x == 0) status[1] = all_ok ? 1u : 0u;
}

// ---------------------------------------------------------------------------------------------------- peer push
// The key exchange WITHOUT a collective call ("one-shot peer-mapped push"): every rank owns a small window
// that its peers map through CUDA IPC at alink_comm_init; ONE single-CTA kernel per step stores this rank's B sorted keys
// into slot [rank] of every peer's window over NVLink, releases a flag per peer, waits for the W flags of its own window
// and merges the W x B keys to the global top-B in its tail (topk_small_body) — exchange and merge in one launch, no
// proxy thread, no second kernel. Windows are double-buffered by the step's parity: a peer can only write parity p again
// after it has merged the step in between, which needs this rank's push of that step, which is stream-ordered after this
// rank's merge of p. A peer that never arrives trips a 2 s timeout (status word), not a hang.
#include "topk.cuh"
constexpr int PUSH_MAX_W = 8, PUSH_MAX_B = 4096;
constexpr size_t PUSH_KEYS_BYTES = 2ull * PUSH_MAX_W * PUSH_MAX_B * 8;        // [parity][W x B keys, slots contiguous at the step's B]
constexpr size_t PUSH_WIN_BYTES = PUSH_KEYS_BYTES + 2ull * PUSH_MAX_W * 8;     // + [parity][W] flags: (epoch << 1) | ok

struct PushArgs {
  uint8_t *win[PUSH_MAX_W];  // win[r]: rank r's window as mapped here (win[rank] = own)
  int rank, W, B, npad;
  unsigned long long epoch;  // >= 1, the same on every rank
  const uint64_t *mine;      // this rank's B keys, sorted (device)
  uint64_t *gathered;        // W x B staging on this GPU
  uint64_t *out;             // merged top-B
  unsigned int *status;      // [0] = 1: a peer timed out | [1] = AND of the ranks' ok bits
  const unsigned int *ok_dev;  // this rank's ok bit (nullptr = 1): FP32_REFINE's verification, evaluated on the device
};

__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v) {
  asm volatile("st.global.release.sys.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.global.acquire.sys.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__global__ void __launch_bounds__(1024) push_merge_kernel(const PushArgs a) {
  __shared__ unsigned int s_bad, s_ok;
  const int tid = threadIdx.x, par = (int)(a.epoch & 1ull);
  if (tid == 0) {
    s_bad = 0;
    s_ok = 1;
  }
  const size_t slot_off = ((size_t)par * PUSH_MAX_W * PUSH_MAX_B + (size_t)a.rank * a.B) * 8;
  for (int p = 0; p < a.W; ++p) {
    uint64_t *dst = reinterpret_cast<uint64_t *>(a.win[p] + slot_off);
    for (int i = tid; i < a.B; i += 1024) dst[i] = a.mine[i];
  }
  __syncthreads();  // (one system-scope fence per flag writer below, not one per thread: 1024 membar.sys cost ~0.1 ms)
  if (tid < a.W) {
    __threadfence_system();
    const unsigned long long okbit = a.ok_dev ? (unsigned long long)(*a.ok_dev & 1u) : 1ull;
    unsigned long long *flag = reinterpret_cast<unsigned long long *>(a.win[tid] + PUSH_KEYS_BYTES) + par * PUSH_MAX_W + a.rank;
    st_release_sys(flag, (a.epoch << 1) | okbit);
    // ... and wait for peer `tid`'s flag in the own window
    const unsigned long long *mineflag =
        reinterpret_cast<const unsigned long long *>(a.win[a.rank] + PUSH_KEYS_BYTES) + par * PUSH_MAX_W + tid;
    const unsigned long long t0 = global_timer_ns();
    unsigned long long v;
    for (;;) {
      v = ld_acquire_sys(mineflag);
      if ((v >> 1) == a.epoch) break;
      if (global_timer_ns() - t0 > 2000000000ull) {
        atomicExch(&s_bad, 1u);
        break;
      }
      __nanosleep(200);
    }
    if (!(v & 1ull)) atomicExch(&s_ok, 0u);
  }
  __syncthreads();
  if (s_bad) {
    if (tid == 0) a.status[0] = 1u;
    return;
  }
  // the W x B keys, read past L1 (peers wrote them), staged on this GPU, merged
  const uint64_t *src = reinterpret_cast<const uint64_t *>(a.win[a.rank] + (size_t)par * PUSH_MAX_W * PUSH_MAX_B * 8);
  const int n = a.W * a.B;
  for (int i = tid; i < n; i += 1024) a.gathered[i] = __ldcg(src + i);
  __syncthreads();
  if (tid == 0) a.status[1] = s_ok;
  KeySrc ks{nullptr, a.gathered, 0, nullptr};
  topk_small_body(ks, n, a.B, a.npad, a.out);
}
