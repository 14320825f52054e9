// Host-side plumbing shared by the store and the operator entry points.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/b200kv.h"
#include "mta.h"

namespace b200kv {

struct Error : public std::runtime_error {
  explicit Error(const std::string& m) : std::runtime_error(m) {}
};

#define KV_CHECK(cond, msg)                                                      \
  do {                                                                           \
    if (!(cond)) {                                                               \
      std::ostringstream os_;                                                    \
      os_ << msg << " [" << __FILE__ << ":" << __LINE__ << "]";                  \
      throw ::b200kv::Error(os_.str());                                          \
    }                                                                            \
  } while (0)

#define CUDA_OK(expr)                                                            \
  do {                                                                           \
    cudaError_t e_ = (expr);                                                     \
    if (e_ != cudaSuccess) {                                                     \
      std::ostringstream os_;                                                    \
      os_ << "CUDA error: " << cudaGetErrorString(e_) << " in " #expr " ["       \
          << __FILE__ << ":" << __LINE__ << "]";                                 \
      throw ::b200kv::Error(os_.str());                                          \
    }                                                                            \
  } while (0)

void set_last_error(const std::string& msg);

// What a b200kv_handle points to: a Store (one GPU per process) or a Group (one process, several GPUs)
struct KVHandle {
  bool is_group = false;
  virtual ~KVHandle() = default;
};

// mshadow type flags (3rdparty/mshadow/mshadow/base.h): the dtype codes MXNet itself uses
enum DType { kF32 = 0, kF16 = 2, kI32 = 4, kI64 = 6 };
inline size_t dtype_size(int dt) { return dt == kF32 || dt == kI32 ? 4 : (dt == kF16 ? 2 : 8); }

// A validated view of a borrowed DLPack tensor
struct TensorView {
  void* data = nullptr;
  int64_t numel = 0;
  int dtype = kF32;
  int ndim = 0;
  int64_t shape[8] = {0};
};
TensorView view_of(const DLManagedTensor* t, int device_id, const char* what);

// grow-only device buffer
struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
  bool owned = true;   // false: memory provided by the host side (symmetric allocator), never freed here
  DevBuf() = default;
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  ~DevBuf() { release(); }
  void release() {
    if (p && owned) cudaFree(p);
    p = nullptr;
    bytes = 0;
    owned = true;
  }
  void adopt(void* ptr, size_t n) {
    release();
    p = ptr;
    bytes = n;
    owned = false;
  }
  // returns true when the buffer was (re)allocated; contents are NOT preserved
  bool reserve(size_t n) {
    if (n <= bytes) return false;
    release();
    CUDA_OK(cudaMalloc(&p, n));
    bytes = n;
    return true;
  }
  template <typename T> T* as() const { return static_cast<T*>(p); }
};

// Ring of pinned-host + device staging slots for the per-call tensor tables, so a call never
// overwrites a table an earlier, still-running launch is reading.
class TableRing {
 public:
  static constexpr int kSlots = 8;
  TableRing() = default;
  ~TableRing();
  // returns a host pointer to fill (bytes) for the next slot
  uint64_t* begin(size_t bytes);
  // uploads the slot on `stream`; returns the device address of the table
  const uint64_t* commit(cudaStream_t stream);
  // call after the launches that read the committed table have been enqueued
  void end(cudaStream_t stream);
  // the device copy of the last table was modified in place: never reuse it
  void invalidate() { last_ = -1; }
  // make every slot hold at least `bytes` now, so that begin() never allocates later (an allocation
  // between two launches keeps them from running concurrently — fatal when they wait for each other)
  void reserve_all(size_t bytes);

 private:
  struct Slot {
    void* host = nullptr;
    void* dev = nullptr;
    size_t cap = 0;
    cudaEvent_t done = nullptr;
    bool busy = false;
  };
  Slot slots_[kSlots];
  int cur_ = -1;
  size_t cur_bytes_ = 0;
  // the slot whose device table was uploaded last: a call whose table is byte-identical (same
  // arrays, same scalars — the steady state of SGD with a constant learning rate) reuses it and
  // skips the upload, so consecutive steps are back-to-back kernel launches
  int last_ = -1;
  size_t last_bytes_ = 0;
  bool reused_ = false;
};

// A chunk list + per-tensor chunk ranges resident on the device
struct ChunkPlan {
  DevBuf chunks;   // Chunk[nchunks]
  DevBuf first;    // int[nt + 1]
  int nchunks = 0;
  int nt = 0;
  // ranges: per tensor [begin, end) element range to cover (relative to tensor element 0);
  // chunk_elems <= kChunk, a multiple of 1024
  void build(const std::vector<int64_t>& begin, const std::vector<int64_t>& end, int chunk_elems = kChunk);
};

int sm_count(int device);

// CUDA loads a kernel lazily at its first launch, and that load can wait for the device to drain.
// Inside a Group a member's kernel may be spinning for a sibling whose launch would then never be
// issued, so a Group loads every kernel of the library up front on each of its devices.
bool register_kernel(const void* host_stub);
void preload_kernels();   // on the current device

}  // namespace b200kv
