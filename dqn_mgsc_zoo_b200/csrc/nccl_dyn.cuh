// NCCL through dlopen: libdqnb200.so must load (and the CPU test suite must run) on machines without NCCL, and
// inside a PyTorch process it must use the very libnccl.so.2 torch already mapped (two NCCL copies in one
// process fight over the same shared-memory / IPC resources).  Only the six entry points the data-parallel
// learner needs; types restated from nccl.h 2.27 (stable since 2.x: ncclFloat32 = 7, ncclSum = 0, 128-byte id).
#pragma once
#include <dlfcn.h>
#include <stdlib.h>

#include "common.cuh"

namespace dqnb {

struct NcclUniqueId { char internal[128]; };
typedef struct ncclComm* NcclComm;

struct NcclApi {
  void* lib = nullptr;
  int (*GetVersion)(int*) = nullptr;
  int (*GetUniqueId)(NcclUniqueId*) = nullptr;
  int (*CommInitRank)(NcclComm*, int, NcclUniqueId, int) = nullptr;
  int (*CommDestroy)(NcclComm) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};

// nullptr + error message when NCCL cannot be found.
inline const NcclApi* nccl_api() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return api.lib ? &api : nullptr;
  tried = true;
  const char* env = getenv("DQNB_NCCL_LIB");
  void* lib = env ? dlopen(env, RTLD_NOW | RTLD_GLOBAL) : nullptr;
  if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);     // the copy already in the process (torch's)
  if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) { set_error("NCCL not found (libnccl.so.2): %s; set DQNB_NCCL_LIB", dlerror()); return nullptr; }
  api.GetVersion = reinterpret_cast<decltype(api.GetVersion)>(dlsym(lib, "ncclGetVersion"));
  api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(dlsym(lib, "ncclGetUniqueId"));
  api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(dlsym(lib, "ncclCommInitRank"));
  api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(dlsym(lib, "ncclCommDestroy"));
  api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(dlsym(lib, "ncclAllReduce"));
  api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(dlsym(lib, "ncclGetErrorString"));
  if (!api.GetVersion || !api.GetUniqueId || !api.CommInitRank || !api.CommDestroy || !api.AllReduce ||
      !api.GetErrorString) {
    set_error("libnccl is missing an entry point the data-parallel learner needs");
    return nullptr;
  }
  api.lib = lib;
  return &api;
}

constexpr int kNcclFloat32 = 7, kNcclSum = 0;

#define DQNB_NCCL(api, expr)                                                                          \
  do {                                                                                                \
    int _r = (expr);                                                                                  \
    if (_r != 0) {                                                                                    \
      dqnb::set_error("%s:%d %s -> NCCL error %d: %s", __FILE__, __LINE__, #expr, _r, (api)->GetErrorString(_r)); \
      return DQNB_ECUDA;                                                                              \
    }                                                                                                 \
  } while (0)

}  // namespace dqnb
