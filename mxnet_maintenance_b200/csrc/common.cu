#include "common.h"

#include <algorithm>
#include <cstring>
#include <mutex>

namespace b200kv {

static thread_local std::string g_last_error;
void set_last_error(const std::string& msg) { g_last_error = msg; }
const char* get_last_error() { return g_last_error.c_str(); }

TensorView view_of(const DLManagedTensor* t, int device_id, const char* what) {
  KV_CHECK(t != nullptr, what << ": null DLManagedTensor");
  const DLTensor& d = t->dl_tensor;
  KV_CHECK(d.device.device_type == kDLCUDA,
           what << ": tensor must live in CUDA device memory (DLPack device_type "
                << d.device.device_type << ")");
  KV_CHECK(d.device.device_id == device_id, what << ": tensor is on device " << d.device.device_id
                                                 << " but the store drives device " << device_id);
  KV_CHECK(d.dtype.lanes == 1, what << ": vector lanes unsupported");
  KV_CHECK(d.ndim >= 0 && d.ndim <= 8, what << ": unsupported ndim " << d.ndim);
  TensorView v;
  if (d.dtype.code == kDLFloat && d.dtype.bits == 32) v.dtype = kF32;
  else if (d.dtype.code == kDLFloat && d.dtype.bits == 16) v.dtype = kF16;
  else if (d.dtype.code == kDLInt && d.dtype.bits == 64) v.dtype = kI64;
  else if (d.dtype.code == kDLInt && d.dtype.bits == 32) v.dtype = kI32;
  else KV_CHECK(false, what << ": unsupported dtype code " << int(d.dtype.code) << " bits "
                            << int(d.dtype.bits));
  v.ndim = d.ndim;
  v.numel = 1;
  for (int i = 0; i < d.ndim; ++i) {
    v.shape[i] = d.shape[i];
    v.numel *= d.shape[i];
  }
  if (d.strides != nullptr && v.numel > 0) {  // MXNet only accepts compact row-major (tensor_blob.h:116-137)
    int64_t expect = 1;
    for (int i = d.ndim - 1; i >= 0; --i) {
      KV_CHECK(d.shape[i] <= 1 || d.strides[i] == expect, what << ": tensor is not compact row-major");
      expect *= d.shape[i];
    }
  }
  v.data = static_cast<char*>(d.data) + d.byte_offset;
  return v;
}

// ---- TableRing ---------------------------------------------------------------------------------
TableRing::~TableRing() {
  for (auto& s : slots_) {
    if (s.host) cudaFreeHost(s.host);
    if (s.dev) cudaFree(s.dev);
    if (s.done) cudaEventDestroy(s.done);
  }
}

static void grow_slot(void** host, void** dev, size_t* cap, size_t bytes) {
  if (bytes <= *cap) return;
  if (*host) cudaFreeHost(*host);
  if (*dev) cudaFree(*dev);
  *host = *dev = nullptr;
  const size_t c = std::max<size_t>((bytes + 4095) & ~size_t(4095), size_t(256) << 10);
  CUDA_OK(cudaMallocHost(host, c));
  CUDA_OK(cudaMalloc(dev, c));
  *cap = c;
}

void TableRing::reserve_all(size_t bytes) {
  for (auto& s : slots_) {
    if (bytes > s.cap) {
      if (s.busy) {
        CUDA_OK(cudaEventSynchronize(s.done));
        s.busy = false;
      }
      if (&s == &slots_[last_ >= 0 ? last_ : 0] && last_ >= 0) last_ = -1;   // its device table goes away
      grow_slot(&s.host, &s.dev, &s.cap, bytes);
    }
    if (!s.done) CUDA_OK(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
  }
}

uint64_t* TableRing::begin(size_t bytes) {
  cur_ = (cur_ + 1) % kSlots;
  if (cur_ == last_) cur_ = (cur_ + 1) % kSlots;   // never stage over the table that may be reused
  Slot& s = slots_[cur_];
  if (s.busy) {  // the launches that read this slot's previous table must have finished
    CUDA_OK(cudaEventSynchronize(s.done));
    s.busy = false;
  }
  grow_slot(&s.host, &s.dev, &s.cap, bytes);
  if (!s.done) CUDA_OK(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
  cur_bytes_ = bytes;
  return static_cast<uint64_t*>(s.host);
}

const uint64_t* TableRing::commit(cudaStream_t stream) {
  Slot& s = slots_[cur_];
  reused_ = last_ >= 0 && last_bytes_ == cur_bytes_ &&
            std::memcmp(slots_[last_].host, s.host, cur_bytes_) == 0;
  if (reused_) {
    cur_ = cur_ == 0 ? kSlots - 1 : cur_ - 1;   // hand the untouched staging slot back
    if (cur_ == last_) cur_ = cur_ == 0 ? kSlots - 1 : cur_ - 1;
    return static_cast<const uint64_t*>(slots_[last_].dev);
  }
  CUDA_OK(cudaMemcpyAsync(s.dev, s.host, cur_bytes_, cudaMemcpyHostToDevice, stream));
  last_ = cur_;
  last_bytes_ = cur_bytes_;
  return static_cast<const uint64_t*>(s.dev);
}

void TableRing::end(cudaStream_t stream) {
  // the event guards the device table these launches read: the reused one, or the new upload
  Slot& s = slots_[last_ >= 0 && reused_ ? last_ : cur_];
  CUDA_OK(cudaEventRecord(s.done, stream));
  s.busy = true;
  reused_ = false;
}

// ---- ChunkPlan ---------------------------------------------------------------------------------
void ChunkPlan::build(const std::vector<int64_t>& begin, const std::vector<int64_t>& end, int chunk_elems) {
  KV_CHECK(chunk_elems > 0 && chunk_elems <= kChunk && chunk_elems % 1024 == 0, "bad chunk size");
  nt = static_cast<int>(begin.size());
  std::vector<Chunk> ch;
  std::vector<int> fst(nt + 1, 0);
  for (int t = 0; t < nt; ++t) {
    fst[t] = static_cast<int>(ch.size());
    for (int64_t o = begin[t]; o < end[t]; o += chunk_elems) {
      Chunk c;
      c.tensor = t;
      c.off = o;
      c.len = static_cast<int32_t>(end[t] - o < chunk_elems ? end[t] - o : chunk_elems);
      ch.push_back(c);
    }
  }
  fst[nt] = static_cast<int>(ch.size());
  nchunks = static_cast<int>(ch.size());
  chunks.reserve(sizeof(Chunk) * (ch.empty() ? 1 : ch.size()));
  first.reserve(sizeof(int) * fst.size());
  if (!ch.empty())
    CUDA_OK(cudaMemcpy(chunks.p, ch.data(), sizeof(Chunk) * ch.size(), cudaMemcpyHostToDevice));
  CUDA_OK(cudaMemcpy(first.p, fst.data(), sizeof(int) * fst.size(), cudaMemcpyHostToDevice));
}

static std::vector<const void*>& kernel_registry() {
  static std::vector<const void*> v;
  return v;
}
static std::mutex& kernel_registry_mu() {
  static std::mutex m;
  return m;
}

bool register_kernel(const void* host_stub) {
  std::lock_guard<std::mutex> lk(kernel_registry_mu());
  kernel_registry().push_back(host_stub);
  return true;
}

void preload_kernels() {
  std::lock_guard<std::mutex> lk(kernel_registry_mu());
  for (const void* k : kernel_registry()) {
    cudaFuncAttributes a;
    CUDA_OK(cudaFuncGetAttributes(&a, k));
  }
}

int sm_count(int device) {
  static std::mutex mu;
  static int cache[64] = {0};
  std::lock_guard<std::mutex> lk(mu);
  if (device < 0 || device >= 64) return 148;
  if (cache[device] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || n <= 0)
      n = 148;
    cache[device] = n;
  }
  return cache[device];
}

}  // namespace b200kv
