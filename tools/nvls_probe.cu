// What the NVSwitch multicast path (multimem.ld_reduce / multimem.st) can reach on this box: one
// process, N GPUs, all running the same kernel at the same time over one multicast object.
//   reduce : each GPU switch-reduces ITS 1/N of the buffer into local memory   (reduce-scatter half)
//   bcast  : each GPU multicasts ITS 1/N of a local buffer to every GPU         (all-gather half)
//   both   : r = ld_reduce(shard); multimem.st(shard) = r                        (what the fused kernel does)
// S = 102 MB (the ResNet-50 bucket).  Prints ms and the two byte rates DESIGN.md §6 uses: algorithmic
// ring bytes 2·S·(N-1)/N and the bytes that really cross a link direction, S·(1+1/N).
// Evidence for "what limits the exchange kernel", not part of the library.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/nvls_probe tools/nvls_probe.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define OK(x)                                                                         \
  do {                                                                                \
    cudaError_t e_ = (x);                                                             \
    if (e_ != cudaSuccess) {                                                          \
      std::fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); \
      std::exit(1);                                                                   \
    }                                                                                 \
  } while (0)
#define DRV(x)                                                              \
  do {                                                                      \
    CUresult r_ = (x);                                                      \
    if (r_ != CUDA_SUCCESS) {                                               \
      const char* s_ = nullptr;                                             \
      cuGetErrorString(r_, &s_);                                            \
      std::fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, s_ ? s_ : "?"); \
      std::exit(1);                                                         \
    }                                                                       \
  } while (0)

__device__ __forceinline__ float4 mc_ld(const float4* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void mc_st(float4* p, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// mode 0 reduce, 1 bcast, 2 both
template <int U, int MODE>
__global__ void probe_k(const float4* __restrict__ mc_in, float4* __restrict__ mc_out,
                        const float4* __restrict__ local_in, float4* __restrict__ local_out, size_t n) {
  const size_t stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride * U) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i + u * stride < n) v[u] = MODE == 1 ? local_in[i + u * stride] : mc_ld(mc_in + i + u * stride);
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i + u * stride < n) {
        if (MODE == 0) local_out[i + u * stride] = v[u];
        else mc_st(mc_out + i + u * stride, v[u]);
      }
  }
}

struct Dev {
  int id;
  CUdeviceptr uc_in = 0, uc_out = 0, mc_in = 0, mc_out = 0;
  float4 *scratch_in = nullptr, *scratch_out = nullptr;
  cudaStream_t s;
  cudaEvent_t e0, e1;
};

static size_t round_up(size_t a, size_t g) { return (a + g - 1) / g * g; }

// one buffer of `bytes` on every device + one multicast object over all of them
static void make_mc(std::vector<Dev>& d, size_t bytes, bool in) {
  const int n = static_cast<int>(d.size());
  CUmulticastObjectProp mp = {};
  mp.numDevices = n;
  mp.size = bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_NONE;
  size_t gran = 0, g_alloc = 0;
  DRV(cuMulticastGetGranularity(&gran, &mp, CU_MULTICAST_GRANULARITY_MINIMUM));
  {
    CUmemAllocationProp ap = {};
    ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ap.location.id = d[0].id;
    DRV(cuMemGetAllocationGranularity(&g_alloc, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  }
  gran = std::max(gran, g_alloc);   // both are powers of two
  mp.size = bytes = round_up(bytes, gran);
  CUmemGenericAllocationHandle mc;
  DRV(cuMulticastCreate(&mc, &mp));
  for (int k = 0; k < n; ++k) DRV(cuMulticastAddDevice(mc, d[k].id));
  for (int k = 0; k < n; ++k) {
    CUmemAllocationProp ap = {};
    ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ap.location.id = d[k].id;
    CUmemGenericAllocationHandle ph;
    DRV(cuMemCreate(&ph, bytes, &ap, 0));
    DRV(cuMulticastBindMem(mc, 0, ph, 0, bytes, 0));
    CUmemAccessDesc ad = {};
    ad.location = ap.location;
    ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    CUdeviceptr uc = 0, mcp = 0;
    DRV(cuMemAddressReserve(&uc, bytes, gran, 0, 0));
    DRV(cuMemMap(uc, bytes, 0, ph, 0));
    DRV(cuMemSetAccess(uc, bytes, &ad, 1));
    DRV(cuMemAddressReserve(&mcp, bytes, gran, 0, 0));
    DRV(cuMemMap(mcp, bytes, 0, mc, 0));
    DRV(cuMemSetAccess(mcp, bytes, &ad, 1));
    OK(cudaSetDevice(d[k].id));
    OK(cudaMemset(reinterpret_cast<void*>(uc), 0, bytes));
    OK(cudaDeviceSynchronize());
    (in ? d[k].uc_in : d[k].uc_out) = uc;
    (in ? d[k].mc_in : d[k].mc_out) = mcp;
  }
}

template <class F>
static double run(std::vector<Dev>& d, int iters, F launch) {
  for (auto& x : d) {
    OK(cudaSetDevice(x.id));
    launch(x);
    launch(x);
  }
  for (auto& x : d) {
    OK(cudaSetDevice(x.id));
    OK(cudaStreamSynchronize(x.s));
  }
  for (auto& x : d) {
    OK(cudaSetDevice(x.id));
    OK(cudaEventRecord(x.e0, x.s));
  }
  for (int it = 0; it < iters; ++it)
    for (auto& x : d) {
      OK(cudaSetDevice(x.id));
      launch(x);
    }
  for (auto& x : d) {
    OK(cudaSetDevice(x.id));
    OK(cudaEventRecord(x.e1, x.s));
  }
  double worst = 0;
  for (auto& x : d) {
    OK(cudaSetDevice(x.id));
    OK(cudaEventSynchronize(x.e1));
    float ms;
    OK(cudaEventElapsedTime(&ms, x.e0, x.e1));
    if (ms > worst) worst = ms;
  }
  return worst / iters;
}

template <int U>
static void sweep(std::vector<Dev>& d, size_t shard, int sms) {
  const int n = static_cast<int>(d.size());
  const double S = shard * 16.0 * n / 1e6;   // MB in the whole bucket
  for (int per_sm : {1, 2, 4, 8}) {
    const int grid = sms * per_sm;
    auto in = [&](Dev& x) { return reinterpret_cast<const float4*>(x.mc_in) + size_t(x.id) * shard; };
    auto out = [&](Dev& x) { return reinterpret_cast<float4*>(x.mc_out) + size_t(x.id) * shard; };
    double t0 = run(d, 20, [&](Dev& x) { probe_k<U, 0><<<grid, 256, 0, x.s>>>(in(x), out(x), x.scratch_in, x.scratch_out, shard); });
    double t1 = run(d, 20, [&](Dev& x) { probe_k<U, 1><<<grid, 256, 0, x.s>>>(in(x), out(x), x.scratch_in, x.scratch_out, shard); });
    double t2 = run(d, 20, [&](Dev& x) { probe_k<U, 2><<<grid, 256, 0, x.s>>>(in(x), out(x), x.scratch_in, x.scratch_out, shard); });
    const double ring = 2.0 * S * (n - 1) / n, cross = S * (1.0 + 1.0 / n);
    std::printf("N=%d U=%d grid=%4d  reduce %.4f ms | bcast %.4f ms | both %.4f ms = %6.1f GB/s ring-equivalent, %6.1f GB/s crossing each link direction\n",
                n, U, grid, t0, t1, t2, ring / t2, cross / t2);
  }
}

int main(int argc, char** argv) {
  int ndev = 0;
  OK(cudaGetDeviceCount(&ndev));
  if (argc > 1) ndev = std::min(ndev, std::atoi(argv[1]));
  if (ndev < 2) {
    std::printf("needs 2 or more GPUs\n");
    return 0;
  }
  DRV(cuInit(0));
  const size_t total = 25557032 / 4;                      // float4 count of the ResNet-50 bucket
  const size_t shard = (total / ndev + 63) / 64 * 64;     // per GPU
  std::vector<Dev> d(ndev);
  for (int k = 0; k < ndev; ++k) {
    d[k].id = k;
    OK(cudaSetDevice(k));
    OK(cudaFree(nullptr));
    int mc_ok = 0;
    DRV(cuDeviceGetAttribute(&mc_ok, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, k));
    if (!mc_ok) {
      std::printf("device %d: no multicast support\n", k);
      return 0;
    }
    for (int j = 0; j < ndev; ++j)
      if (j != k) OK(cudaDeviceEnablePeerAccess(j, 0));
    OK(cudaMalloc(&d[k].scratch_in, shard * 16));
    OK(cudaMalloc(&d[k].scratch_out, shard * 16));
    OK(cudaMemset(d[k].scratch_in, 0, shard * 16));
    OK(cudaStreamCreateWithFlags(&d[k].s, cudaStreamNonBlocking));
    OK(cudaEventCreate(&d[k].e0));
    OK(cudaEventCreate(&d[k].e1));
  }
  make_mc(d, shard * 16 * ndev, true);
  make_mc(d, shard * 16 * ndev, false);
  int sms = 0;
  OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  sweep<1>(d, shard, sms);
  sweep<2>(d, shard, sms);
  sweep<4>(d, shard, sms);
  sweep<8>(d, shard, sms);
  return 0;
}
