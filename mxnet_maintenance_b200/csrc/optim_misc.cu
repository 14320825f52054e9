// Host launchers of the small helper kernels (reductions, barriers, casts).
#include "launch_impl.cuh"

namespace b200kv {

int launch_all_finite_final(const float* sums, int nt, int init_output, float* out, cudaStream_t stream) {
  all_finite_final_kernel<<<1, 256, 0, stream>>>(sums, nt, init_output, out);
  CUDA_OK(cudaGetLastError());
  return 1;
}

int launch_multi_lars(int n, const float* lrs, const float* w2, const float* g2, const float* wds, float eta,
                      float eps, float rescale, float* out, cudaStream_t stream) {
  if (n <= 0) return 0;
  multi_lars_kernel<<<(n + 255) / 256, 256, 0, stream>>>(n, lrs, w2, g2, wds, eta, eps, rescale, out);
  CUDA_OK(cudaGetLastError());
  return 1;
}

int launch_fill_lrwd(const MtaCall& c, const float* lrs_dev, const float* wds_dev) {
  uint64_t* section = const_cast<uint64_t*>(c.blob) + static_cast<size_t>(c.ngrad + c.nout + kSecLrWd) * c.nt;
  fill_lrwd_kernel<<<(c.nt + 255) / 256, 256, 0, c.stream>>>(section, lrs_dev, wds_dev, c.nt);
  CUDA_OK(cudaGetLastError());
  return 1;
}

int launch_p2p_push(const void* src, void* const* dst, int64_t slice_bytes, int world, int rank, int device,
                    cudaStream_t stream) {
  KV_CHECK(slice_bytes % 16 == 0, "slices must be multiples of 16 bytes");
  PushArgs a;
  for (int k = 0; k < kMaxCopies; ++k) a.dst[k] = k < world ? static_cast<uint4*>(dst[k]) : nullptr;
  a.src = static_cast<const uint4*>(src);
  a.slice_vecs = slice_bytes / 16;
  a.world = world;
  a.rank = rank;
  const int64_t ntiles = (a.slice_vecs + 1023) / 1024 * (world - 1);
  const int g = sm_count(device) * 4;
  const int grid = static_cast<int>(ntiles < g ? (ntiles > 0 ? ntiles : 1) : g);
  p2p_push_kernel<<<grid, 256, 0, stream>>>(a);
  CUDA_OK(cudaGetLastError());
  return 1;
}

int launch_p2p_barrier(const P2PFlags& f, int which, cudaStream_t stream) {
  p2p_barrier_kernel<<<1, 32, 0, stream>>>(f, which);
  CUDA_OK(cudaGetLastError());
  return 1;
}

static const bool g_misc_registered =
    register_kernel(reinterpret_cast<const void*>(&xsum_kernel)) &&
    register_kernel(reinterpret_cast<const void*>(&reduce_partials_kernel)) &&
    register_kernel(reinterpret_cast<const void*>(&lamb_ratio_kernel)) &&
    register_kernel(reinterpret_cast<const void*>(&global_norm_kernel)) &&
    register_kernel(reinterpret_cast<const void*>(&p2p_barrier_kernel)) &&
    register_kernel(reinterpret_cast<const void*>(&p2p_push_kernel));

int launch_xsum(float* sums, int n, const P2PFlags& f, cudaStream_t stream) {
  KV_CHECK(n >= 0 && n <= kXsumCap, "too many partial sums for the peer-memory exchange (" << n << " > " << kXsumCap << ")");
  if (n == 0) return 0;
  xsum_kernel<<<1, 256, 0, stream>>>(sums, n, f);
  CUDA_OK(cudaGetLastError());
  return 1;
}

int launch_reduce_partials(const float* partials, const ChunkPlan& plan, int nred, float* sums,
                           cudaStream_t stream) {
  if (plan.nt == 0) return 0;
  const int warps_per_block = 8;
  const int blocks = (plan.nt + warps_per_block - 1) / warps_per_block;
  reduce_partials_kernel<<<blocks, warps_per_block * 32, 0, stream>>>(partials, plan.first.as<int>(),
                                                                      plan.nt, nred, sums);
  CUDA_OK(cudaGetLastError());
  return 1;
}

int launch_lamb_ratio(const float* sums, int nt, float lower, float upper, float* ratio,
                      cudaStream_t stream) {
  if (nt == 0) return 0;
  lamb_ratio_kernel<<<(nt + 255) / 256, 256, 0, stream>>>(sums, nt, lower, upper, ratio);
  CUDA_OK(cudaGetLastError());
  return 1;
}

int launch_global_norm(const float* sums, int nt, float max_norm, float* out2, cudaStream_t stream) {
  global_norm_kernel<<<1, kThreads, 0, stream>>>(sums, nt, max_norm, out2);
  CUDA_OK(cudaGetLastError());
  return 1;
}

}  // namespace b200kv

namespace b200kv {

template <typename S, typename D>
__global__ void cast_kernel(const S* __restrict__ src, D* __restrict__ dst, int64_t n) {
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x)
    st1(dst + i, ld1(src + i));
}

// whole-array fp16 <-> fp32 conversion (master creation / materialising the fp16 stored value)
int launch_cast(const void* src, int src_dtype, void* dst, int dst_dtype, int64_t n,
                cudaStream_t stream) {
  if (n <= 0) return 0;
  const int grid = static_cast<int>(std::min<int64_t>((n + 255) / 256, 148 * 8));
  if (src_dtype == kF16 && dst_dtype == kF32)
    cast_kernel<<<grid, 256, 0, stream>>>(static_cast<const __half*>(src), static_cast<float*>(dst), n);
  else if (src_dtype == kF32 && dst_dtype == kF16)
    cast_kernel<<<grid, 256, 0, stream>>>(static_cast<const float*>(src), static_cast<__half*>(dst), n);
  else
    KV_CHECK(false, "unsupported cast " << src_dtype << " -> " << dst_dtype);
  CUDA_OK(cudaGetLastError());
  return 1;
}

}  // namespace b200kv
