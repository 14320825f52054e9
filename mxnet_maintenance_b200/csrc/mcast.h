// NVSwitch multicast memory for a Group (one process, several GPUs): one physical allocation per device,
// every one mapped for all devices of the group, plus a multicast object over them, so that the exchange
// kernel can use multimem.ld_reduce / multimem.st exactly as in the one-process-per-GPU mode (where the
// host's symmetric allocator provides such memory).  Driver API only (cuMulticast*, cuMem*), reached
// through cudaGetDriverEntryPoint so that the library keeps no link-time dependency on libcuda.
#pragma once
#include <cstddef>
#include <memory>
#include <vector>

namespace b200kv {

class McArena {
 public:
  // nullptr when the driver or one of the devices has no multicast support
  static std::unique_ptr<McArena> create(const std::vector<int>& devices);
  ~McArena();

  // The j-th request of every member names the same buffer: the first member to ask creates it for all.
  // Returns false (nothing changed) when the request does not match what its siblings asked for or the
  // driver refuses; the caller then falls back to plain device memory and the peer load/store exchange.
  bool request(int member, size_t bytes, void** local, void** peers, void** mc);

 private:
  struct Set;
  struct Api;
  explicit McArena(const std::vector<int>& devices);
  Set* make(size_t bytes);
  void destroy(Set& s);

  std::vector<int> dev_;
  std::vector<size_t> next_;                 // per member: index of its next request
  std::vector<std::unique_ptr<Set>> sets_;
  const Api* api_ = nullptr;
  bool dead_ = false;                        // a request failed: every later one is refused, for all members alike
};

}  // namespace b200kv
