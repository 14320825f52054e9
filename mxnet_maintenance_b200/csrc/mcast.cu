#include "mcast.h"

#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cstring>

#include "common.h"

namespace b200kv {

// driver entry points, resolved once
struct McArena::Api {
#define B200KV_DRV(X)                  \
  X(cuDeviceGet)                       \
  X(cuDeviceGetAttribute)              \
  X(cuMulticastCreate)                 \
  X(cuMulticastAddDevice)              \
  X(cuMulticastBindMem)                \
  X(cuMulticastGetGranularity)         \
  X(cuMemCreate)                       \
  X(cuMemRelease)                      \
  X(cuMemAddressReserve)               \
  X(cuMemAddressFree)                  \
  X(cuMemMap)                          \
  X(cuMemUnmap)                        \
  X(cuMemSetAccess)                    \
  X(cuMemGetAllocationGranularity)
#define DECL(name) decltype(&::name) name = nullptr;
  B200KV_DRV(DECL)
#undef DECL
  bool ok = false;

  static const Api& get() {
    static const Api a = [] {
      Api x;
      bool all = true;
#define LOAD(name)                                                                                          \
  {                                                                                                         \
    void* f = nullptr;                                                                                      \
    cudaDriverEntryPointQueryResult st;                                                                     \
    if (cudaGetDriverEntryPoint(#name, &f, cudaEnableDefault, &st) != cudaSuccess || f == nullptr ||        \
        st != cudaDriverEntryPointSuccess) {                                                                \
      (void)cudaGetLastError();                                                                             \
      all = false;                                                                                          \
    }                                                                                                       \
    x.name = reinterpret_cast<decltype(x.name)>(f);                                                         \
  }
      B200KV_DRV(LOAD)
#undef LOAD
      x.ok = all;
      return x;
    }();
    return a;
  }
};

struct McArena::Set {
  size_t bytes = 0;        // what the members asked for
  size_t mapped = 0;       // rounded up to the granularity
  CUmemGenericAllocationHandle mc = 0;
  std::vector<CUmemGenericAllocationHandle> phys;
  std::vector<CUdeviceptr> uc;
  std::vector<bool> uc_mapped;
  CUdeviceptr mcva = 0;
  bool mc_mapped = false;
};

std::unique_ptr<McArena> McArena::create(const std::vector<int>& devices) {
  const Api& api = Api::get();
  if (!api.ok || devices.size() < 2) return nullptr;
  for (size_t i = 0; i < devices.size(); ++i) {
    if (std::find(devices.begin(), devices.begin() + i, devices[i]) != devices.begin() + i) return nullptr;   // virtual ranks
    CUdevice d;
    int yes = 0;
    if (api.cuDeviceGet(&d, devices[i]) != CUDA_SUCCESS) return nullptr;
    if (api.cuDeviceGetAttribute(&yes, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d) != CUDA_SUCCESS || !yes) return nullptr;
  }
  std::unique_ptr<McArena> a(new McArena(devices));
  a->api_ = &api;
  return a;
}

McArena::McArena(const std::vector<int>& devices) : dev_(devices), next_(devices.size(), 0) {}

void McArena::destroy(Set& s) {
  const Api& api = *api_;
  if (s.mc_mapped) api.cuMemUnmap(s.mcva, s.mapped);
  if (s.mcva) api.cuMemAddressFree(s.mcva, s.mapped);
  for (size_t k = 0; k < s.uc.size(); ++k)
    if (s.uc[k]) {
      if (s.uc_mapped[k]) api.cuMemUnmap(s.uc[k], s.mapped);
      api.cuMemAddressFree(s.uc[k], s.mapped);
    }
  for (auto h : s.phys)
    if (h) api.cuMemRelease(h);
  if (s.mc) api.cuMemRelease(s.mc);
}

McArena::~McArena() {
  for (auto& sp : sets_) destroy(*sp);
}

McArena::Set* McArena::make(size_t bytes) {
  const Api& api = *api_;
  const int n = static_cast<int>(dev_.size());
  std::unique_ptr<Set> sp(new Set());
  Set& s = *sp;
  s.bytes = bytes;
  s.phys.assign(n, 0);
  s.uc.assign(n, 0);
  s.uc_mapped.assign(n, false);
  struct Undo {   // a failure below releases what has been built so far
    McArena* a;
    Set* s;
    bool good = false;
    ~Undo() {
      if (!good) a->destroy(*s);
    }
  } undo{this, &s};

  CUmulticastObjectProp mp;
  std::memset(&mp, 0, sizeof(mp));
  mp.numDevices = n;
  mp.size = bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_NONE;
  size_t gran = 0, g_alloc = 0;
  if (api.cuMulticastGetGranularity(&gran, &mp, CU_MULTICAST_GRANULARITY_MINIMUM) != CUDA_SUCCESS) return nullptr;
  std::vector<CUmemAccessDesc> access(n);
  for (int k = 0; k < n; ++k) {
    std::memset(&access[k], 0, sizeof(access[k]));
    access[k].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    access[k].location.id = dev_[k];
    access[k].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  }
  CUmemAllocationProp ap;
  std::memset(&ap, 0, sizeof(ap));
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = dev_[0];
  if (api.cuMemGetAllocationGranularity(&g_alloc, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS) return nullptr;
  gran = std::max(gran, g_alloc);   // both are powers of two
  s.mapped = (bytes + gran - 1) / gran * gran;
  mp.size = s.mapped;
  if (api.cuMulticastCreate(&s.mc, &mp) != CUDA_SUCCESS) return nullptr;
  for (int k = 0; k < n; ++k) {
    CUdevice d;
    if (api.cuDeviceGet(&d, dev_[k]) != CUDA_SUCCESS) return nullptr;
    if (api.cuMulticastAddDevice(s.mc, d) != CUDA_SUCCESS) return nullptr;
  }
  for (int k = 0; k < n; ++k) {
    ap.location.id = dev_[k];
    if (api.cuMemCreate(&s.phys[k], s.mapped, &ap, 0) != CUDA_SUCCESS) return nullptr;
    if (api.cuMulticastBindMem(s.mc, 0, s.phys[k], 0, s.mapped, 0) != CUDA_SUCCESS) return nullptr;
    CUdeviceptr va = 0;
    if (api.cuMemAddressReserve(&va, s.mapped, gran, 0, 0) != CUDA_SUCCESS) return nullptr;
    s.uc[k] = va;
    if (api.cuMemMap(va, s.mapped, 0, s.phys[k], 0) != CUDA_SUCCESS) return nullptr;
    s.uc_mapped[k] = true;
    if (api.cuMemSetAccess(va, s.mapped, access.data(), n) != CUDA_SUCCESS) return nullptr;   // every device of the group
  }
  if (api.cuMemAddressReserve(&s.mcva, s.mapped, gran, 0, 0) != CUDA_SUCCESS) return nullptr;
  if (api.cuMemMap(s.mcva, s.mapped, 0, s.mc, 0) != CUDA_SUCCESS) return nullptr;
  s.mc_mapped = true;
  if (api.cuMemSetAccess(s.mcva, s.mapped, access.data(), n) != CUDA_SUCCESS) return nullptr;
  undo.good = true;
  sets_.push_back(std::move(sp));
  return sets_.back().get();
}

bool McArena::request(int member, size_t bytes, void** local, void** peers, void** mc) {
  const int n = static_cast<int>(dev_.size());
  if (dead_ || member < 0 || member >= n || bytes == 0) return false;
  const size_t j = next_[member];
  if (j == sets_.size()) {           // this member is the first to ask for buffer j
    if (make(bytes) == nullptr) {
      dead_ = true;                  // its siblings must not get multicast memory for buffer j either
      return false;
    }
  }
  Set& s = *sets_[j];
  if (s.bytes != bytes) {            // the members are not asking for the same things: give up for good
    dead_ = true;
    return false;
  }
  ++next_[member];
  *local = reinterpret_cast<void*>(s.uc[member]);
  for (int k = 0; k < n; ++k) peers[k] = reinterpret_cast<void*>(s.uc[k]);
  *mc = reinterpret_cast<void*>(s.mcva);
  return true;
}

}  // namespace b200kv
