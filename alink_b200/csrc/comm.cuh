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
