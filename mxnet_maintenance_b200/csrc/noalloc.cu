// b200kv_multi_sum_sq_static: the multi_sum_sq entry point for callers that must not allocate.
//
// The regular b200kv_multi_sum_sq keeps its tensor table, chunk list and per-chunk partial sums in
// buffers it obtains with cudaMalloc / cudaMallocHost and grows on demand.  Some hosts of this
// library cannot afford that: MXNet's pooled storage manager hands (almost) all device memory to
// its own pool, so a late cudaMalloc from a plugin can fail in the middle of training; cudaMalloc
// also synchronises the whole device, which stalls every other stream of a latency-sensitive
// trainer; and processes that run under a fixed memory budget want the footprint of a library to
// be known when it is loaded.  For them this variant works entirely inside ONE fixed array in
// device memory that is part of the library image: every GPU that loads libb200kv.so holds these
// kStaticBytes (about 335 KiB) until the process exits, and nothing else is ever allocated here —
// no cudaMalloc, cudaFree, cudaMallocHost or any other allocating call, on the device or through
// a framework allocator.
#include <cstring>
#include <mutex>

#include "optim.cuh"

#include "common.h"

namespace b200kv {
const char* get_last_error();
}

namespace {

using namespace b200kv;

constexpr int kMaxArrays = 256;      // B200KV_STATIC_MAX_ARRAYS
constexpr int kMaxChunks = 16384;    // B200KV_STATIC_MAX_CHUNKS (4096 elements per chunk)

constexpr size_t align256(size_t x) { return (x + 255) & ~size_t(255); }
constexpr size_t kBlobBytes = align256(sizeof(uint64_t) * 7 * kMaxArrays);   // grad | 6 sections
constexpr size_t kChunkBytes = align256(sizeof(Chunk) * kMaxChunks);
constexpr size_t kFirstBytes = align256(sizeof(int) * (kMaxArrays + 1));
constexpr size_t kPartialBytes = align256(sizeof(float) * kMaxChunks);
constexpr size_t kOffBlob = 0, kOffChunks = kOffBlob + kBlobBytes, kOffFirst = kOffChunks + kChunkBytes,
                 kOffPartials = kOffFirst + kFirstBytes, kStaticBytes = kOffPartials + kPartialBytes;

// The scratch itself.  In host code this name is only a stand-in for the symbol — it is NOT a
// device address; the address each GPU gave the array is looked up with cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];

// Host-side staging for the three small tables (plain static storage: nothing is allocated).
struct HostStage {
  uint64_t blob[7 * kMaxArrays];
  Chunk chunks[kMaxChunks];
  int first[kMaxArrays + 1];
};
HostStage g_stage;
std::mutex g_mu;                       // one call at a time fills the staging area
unsigned char* g_dev_ptr[64] = {nullptr};   // per-device address of g_static_scratch

// Once per device, with that device current: ask the runtime where this GPU placed the array.
unsigned char* scratch_on(int device) {
  KV_CHECK(device >= 0 && device < 64, "device ordinal out of range");
  if (g_dev_ptr[device] == nullptr) {
    unsigned char* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_scratch);
    KV_CHECK(e == cudaSuccess && p != nullptr,
             "cudaGetSymbolAddress(static scratch) failed: " << cudaGetErrorString(e));
    g_dev_ptr[device] = p;
  }
  return g_dev_ptr[device];
}

// SM count x resident CTAs per SM of this instantiation, capped by the chunk count: the rule
// launch_impl.cuh's grid_for applies to the allocating entry point.
template <class K>
int persistent_grid(int device, int nchunks, K kernel) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kThreads, 0) != cudaSuccess || per_sm < 1)
    per_sm = 2;
  const int g = sm_count(device) * per_sm;
  return nchunks < g ? nchunks : g;
}

}  // namespace

extern "C" int b200kv_multi_sum_sq_static(int num, int dtype, const void* const* arrays,
                                          const int64_t* sizes, float* out, void* stream) {
  try {
    if (num <= 0) return 0;
    KV_CHECK(dtype == kF32 || dtype == kF16, "dtype must be 0 (float32) or 2 (float16)");
    KV_CHECK(num <= kMaxArrays, "b200kv_multi_sum_sq_static handles at most " << kMaxArrays
                                    << " arrays per call, got " << num
                                    << " (use b200kv_multi_sum_sq, which allocates, or split the list)");
    int64_t nchunks = 0;
    for (int i = 0; i < num; ++i) {
      KV_CHECK(sizes[i] >= 0, "negative size for array " << i);
      nchunks += (sizes[i] + kChunk - 1) / kChunk;
    }
    KV_CHECK(nchunks <= kMaxChunks, "b200kv_multi_sum_sq_static handles at most " << kMaxChunks
                                        << " chunks of " << kChunk << " elements per call ("
                                        << static_cast<int64_t>(kMaxChunks) * kChunk
                                        << " elements when every array is a multiple of " << kChunk
                                        << "); this call needs " << nchunks);
    int device = 0;
    CUDA_OK(cudaGetDevice(&device));
    std::lock_guard<std::mutex> lk(g_mu);
    unsigned char* base = scratch_on(device);   // first call on this device resolves the address
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    // tensor table (one gradient section + the six fixed sections, see mta.h) and chunk list
    std::memset(g_stage.blob, 0, sizeof(uint64_t) * 7 * static_cast<size_t>(num));
    int c = 0;
    for (int t = 0; t < num; ++t) {
      g_stage.blob[t] = reinterpret_cast<uint64_t>(arrays[t]);
      g_stage.first[t] = c;
      for (int64_t o = 0; o < sizes[t]; o += kChunk) {
        g_stage.chunks[c].tensor = t;
        g_stage.chunks[c].off = o;
        g_stage.chunks[c].len = static_cast<int32_t>(sizes[t] - o < kChunk ? sizes[t] - o : kChunk);
        ++c;
      }
    }
    g_stage.first[num] = c;
    // Pageable-source async copies return once the source has been staged by the driver, so the
    // static staging area may be refilled by the next call right away.
    CUDA_OK(cudaMemcpyAsync(base + kOffBlob, g_stage.blob, sizeof(uint64_t) * 7 * static_cast<size_t>(num),
                            cudaMemcpyHostToDevice, st));
    if (c > 0)
      CUDA_OK(cudaMemcpyAsync(base + kOffChunks, g_stage.chunks, sizeof(Chunk) * static_cast<size_t>(c),
                              cudaMemcpyHostToDevice, st));
    CUDA_OK(cudaMemcpyAsync(base + kOffFirst, g_stage.first, sizeof(int) * static_cast<size_t>(num + 1),
                            cudaMemcpyHostToDevice, st));

    MtaArgs a;
    a.blob = reinterpret_cast<const uint64_t*>(base + kOffBlob);
    a.chunks = reinterpret_cast<const Chunk*>(base + kOffChunks);
    a.nchunks = c;
    a.nt = num;
    a.ngrad = 1;
    a.nout = 0;
    a.sync = nullptr;
    a.epoch = 0;
    a.sync_mode = 0;
    float* partials = reinterpret_cast<float*>(base + kOffPartials);
    // Nothing in the scratch has to start at zero: stage 1 writes partials[0, c) and stage 2 reads
    // exactly that range, so whatever an earlier call left behind is never observed.
    if (c > 0) {
      // Same persistent grid as b200kv_multi_sum_sq (launch_impl.cuh grid_for): the kernel carries
      // register partials from chunk c to chunk c + gridDim of the same tensor, so the fp32
      // summation order — and with it the last bit of the result — depends on the grid size.
      // The occupancy query allocates nothing.
      SumSqOp op;
      if (dtype == kF32) {
        const int grid = persistent_grid(device, c, mta_kernel<SumSqOp, float, float, 1>);
        mta_kernel<SumSqOp, float, float, 1><<<grid, kThreads, 0, st>>>(a, op, nullptr, partials);
      } else {
        const int grid = persistent_grid(device, c, mta_kernel<SumSqOp, __half, __half, 1>);
        mta_kernel<SumSqOp, __half, __half, 1><<<grid, kThreads, 0, st>>>(a, op, nullptr, partials);
      }
      CUDA_OK(cudaGetLastError());
    }
    const int warps_per_block = 8;
    reduce_partials_kernel<<<(num + warps_per_block - 1) / warps_per_block, warps_per_block * 32, 0, st>>>(
        partials, reinterpret_cast<const int*>(base + kOffFirst), num, 1, out);
    CUDA_OK(cudaGetLastError());
  } catch (const std::exception& e) {
    b200kv::set_last_error(e.what());
    return -1;
  } catch (...) {
    b200kv::set_last_error("unknown C++ exception");
    return -1;
  }
  return 0;
}
