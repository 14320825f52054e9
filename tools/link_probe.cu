// What an SM-driven exchange over NVLink can reach on this box: one process, two GPUs, both running
// the same kernel at the same time.  Three access patterns over a 51 MB shard (the C2 shard at N=2):
//   pull : read the peer's buffer, write a local one            (reduce-scatter half of the exchange)
//   push : read a local buffer, write the peer's                (all-gather half)
//   both : out = peer_in + local_in, stored locally AND on the peer (what the fused kernel does)
// Prints GB/s per link direction for several grid sizes / loads in flight.  Evidence for DESIGN.md §6
// ("what limits the exchange kernel"), not part of the library.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/link_probe tools/link_probe.cu
#include <cuda_runtime.h>

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

template <int U>
__global__ void pull_k(const float4* __restrict__ peer_in, float4* __restrict__ local_out, size_t n) {
  const size_t stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride * U) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i + u * stride < n) v[u] = peer_in[i + u * stride];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i + u * stride < n) local_out[i + u * stride] = v[u];
  }
}

template <int U>
__global__ void both_k(const float4* __restrict__ peer_in, const float4* __restrict__ local_in,
                       float4* __restrict__ local_out, float4* __restrict__ peer_out, size_t n) {
  const size_t stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride * U) {
    float4 a[U], b[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i + u * stride < n) {
        a[u] = peer_in[i + u * stride];
        b[u] = local_in[i + u * stride];
      }
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i + u * stride < n) {
        float4 r = make_float4(a[u].x + b[u].x, a[u].y + b[u].y, a[u].z + b[u].z, a[u].w + b[u].w);
        local_out[i + u * stride] = r;
        peer_out[i + u * stride] = r;
      }
  }
}

struct Dev {
  int id;
  float4 *in, *out, *scratch;
  cudaStream_t s;
  cudaEvent_t e0, e1;
};

template <class F>
static double run(Dev d[2], int iters, F launch) {
  for (int k = 0; k < 2; ++k) {
    OK(cudaSetDevice(d[k].id));
    launch(k);
    launch(k);
  }
  for (int k = 0; k < 2; ++k) {
    OK(cudaSetDevice(d[k].id));
    OK(cudaStreamSynchronize(d[k].s));
  }
  for (int k = 0; k < 2; ++k) {
    OK(cudaSetDevice(d[k].id));
    OK(cudaEventRecord(d[k].e0, d[k].s));
  }
  for (int it = 0; it < iters; ++it)
    for (int k = 0; k < 2; ++k) {
      OK(cudaSetDevice(d[k].id));
      launch(k);
    }
  double worst = 0;
  for (int k = 0; k < 2; ++k) {
    OK(cudaSetDevice(d[k].id));
    OK(cudaEventRecord(d[k].e1, d[k].s));
  }
  for (int k = 0; k < 2; ++k) {
    OK(cudaSetDevice(d[k].id));
    OK(cudaEventSynchronize(d[k].e1));
    float ms;
    OK(cudaEventElapsedTime(&ms, d[k].e0, d[k].e1));
    if (ms > worst) worst = ms;
  }
  return worst / iters;
}

template <int U>
static void sweep(Dev d[2], size_t n, int sms) {
  const double mb = n * 16.0 / 1e6;
  for (int per_sm : {1, 2, 4, 8}) {
    const int grid = sms * per_sm;
    double t_pull = run(d, 20, [&](int k) { pull_k<U><<<grid, 256, 0, d[k].s>>>(d[1 - k].in, d[k].scratch, n); });
    double t_push = run(d, 20, [&](int k) { pull_k<U><<<grid, 256, 0, d[k].s>>>(d[k].in, d[1 - k].out, n); });
    double t_both = run(d, 20, [&](int k) {
      both_k<U><<<grid, 256, 0, d[k].s>>>(d[1 - k].in, d[k].in, d[k].scratch, d[1 - k].out, n);
    });
    std::printf("U=%d grid=%4d  pull %.4f ms %6.1f GB/s | push %.4f ms %6.1f GB/s | both %.4f ms %6.1f GB/s per direction "
                "(%.1f MB in + %.1f MB out per GPU)\n",
                U, grid, t_pull, mb / t_pull, t_push, mb / t_push, t_both, 2 * mb / t_both, mb, mb);
  }
}

int main(int argc, char** argv) {
  int ndev = 0;
  OK(cudaGetDeviceCount(&ndev));
  if (ndev < 2) {
    std::printf("needs 2 GPUs\n");
    return 0;
  }
  const size_t n = (argc > 1 ? std::atol(argv[1]) : 12778516) / 4;   // float4 count: half of ResNet-50
  Dev d[2];
  for (int k = 0; k < 2; ++k) {
    d[k].id = k;
    OK(cudaSetDevice(k));
    OK(cudaDeviceEnablePeerAccess(1 - k, 0));
    OK(cudaMalloc(&d[k].in, n * 16));
    OK(cudaMalloc(&d[k].out, n * 16));
    OK(cudaMalloc(&d[k].scratch, n * 16));
    OK(cudaMemset(d[k].in, 0, n * 16));
    OK(cudaStreamCreateWithFlags(&d[k].s, cudaStreamNonBlocking));
    OK(cudaEventCreate(&d[k].e0));
    OK(cudaEventCreate(&d[k].e1));
  }
  int sms = 0;
  OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  // copy-engine reference: both GPUs pull 51 MB from each other at once
  double t_ce = run(d, 20, [&](int k) { OK(cudaMemcpyPeerAsync(d[k].scratch, k, d[1 - k].in, 1 - k, n * 16, d[k].s)); });
  std::printf("copy engine, both GPUs pulling at once: %.4f ms  %.1f GB/s per direction\n", t_ce, n * 16.0 / 1e6 / t_ce);
  sweep<1>(d, n, sms);
  sweep<2>(d, n, sms);
  sweep<4>(d, n, sms);
  sweep<8>(d, n, sms);
  return 0;
}
