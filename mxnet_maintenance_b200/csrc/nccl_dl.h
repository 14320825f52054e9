// NCCL is bound at run time with dlopen so that libb200kv.so has no link-time dependency on a
// particular libnccl: inside a PyTorch process it resolves to the already-loaded torch-bundled
// NCCL, in a plain C/C++ host to the system libnccl.so.2.
#pragma once
#include <dlfcn.h>
#include <nccl.h>

#include <mutex>

#include "common.h"

namespace b200kv {

struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*);
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
  ncclResult_t (*CommDestroy)(ncclComm_t);
  ncclResult_t (*ReduceScatter)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                                cudaStream_t);
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                            cudaStream_t);
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t,
                            cudaStream_t);
  ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t,
                         cudaStream_t);
  ncclResult_t (*GroupStart)();
  ncclResult_t (*GroupEnd)();
  const char* (*GetErrorString)(ncclResult_t);
  ncclResult_t (*GetVersion)(int*);

  static NcclApi& get() {
    static NcclApi api;
    static std::once_flag once;
    static std::string err;
    std::call_once(once, [&]() {
      void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
      if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
      if (!h) {
        err = std::string("cannot load libnccl.so.2: ") + dlerror();
        return;
      }
#define LOAD(field, sym)                                                  \
  api.field = reinterpret_cast<decltype(api.field)>(dlsym(h, sym));       \
  if (!api.field) err = std::string("libnccl lacks symbol ") + sym;
      LOAD(GetUniqueId, "ncclGetUniqueId")
      LOAD(CommInitRank, "ncclCommInitRank")
      LOAD(CommDestroy, "ncclCommDestroy")
      LOAD(ReduceScatter, "ncclReduceScatter")
      LOAD(AllGather, "ncclAllGather")
      LOAD(AllReduce, "ncclAllReduce")
      LOAD(Broadcast, "ncclBroadcast")
      LOAD(Reduce, "ncclReduce")
      LOAD(GroupStart, "ncclGroupStart")
      LOAD(GroupEnd, "ncclGroupEnd")
      LOAD(GetErrorString, "ncclGetErrorString")
      LOAD(GetVersion, "ncclGetVersion")
#undef LOAD
    });
    KV_CHECK(err.empty(), err);
    return api;
  }
};

#define NCCL_OK(expr)                                                                         \
  do {                                                                                        \
    ncclResult_t r_ = (expr);                                                                 \
    KV_CHECK(r_ == ncclSuccess, "NCCL error: " << ::b200kv::NcclApi::get().GetErrorString(r_) \
                                               << " in " #expr);                              \
  } while (0)

inline ncclDataType_t nccl_type(int dtype) { return dtype == kF16 ? ncclFloat16 : ncclFloat32; }

}  // namespace b200kv
