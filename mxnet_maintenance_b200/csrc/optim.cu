// Host launchers for the multi-tensor kernels in optim.cuh.
#include "optim.cuh"

#include "launch.h"

#include <algorithm>

namespace b200kv {

namespace {

// persistent grid: SM count x resident CTAs per SM for this instantiation (occupancy API, cached)
template <class K>
int grid_for(const MtaCall& c, K kernel) {
  static int per_sm = 0;
  if (per_sm == 0) {
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kThreads, 0) != cudaSuccess || n < 1) n = 2;
    per_sm = n;
  }
  const int g = sm_count(c.device) * per_sm;
  return c.plan->nchunks < g ? (c.plan->nchunks > 0 ? c.plan->nchunks : 1) : g;
}

MtaArgs args_of(const MtaCall& c) {
  MtaArgs a;
  a.blob = c.blob;
  a.chunks = c.plan->chunks.as<Chunk>();
  a.nchunks = c.plan->nchunks;
  a.nt = c.nt;
  a.ngrad = c.ngrad;
  a.nout = c.nout;
  a.sync = c.sync;
  a.epoch = c.epoch;
  a.sync_mode = c.sync_mode;
  return a;
}

template <class Op, typename T, typename WT>
int launch_ng(const MtaCall& c, const Op& op, const float* gscale, float* partials) {
  if (c.plan->nchunks == 0) return 0;
  KV_CHECK(c.ngrad >= 1 && c.ngrad <= kMaxCopies, "unsupported number of gradient copies " << c.ngrad);
  KV_CHECK(c.nout >= 0 && c.nout <= kMaxCopies, "unsupported number of output copies " << c.nout);
  const MtaArgs a = args_of(c);
#define GO(NG)                                                                                     \
  mta_kernel<Op, T, WT, NG><<<grid_for(c, mta_kernel<Op, T, WT, NG>), kThreads, 0, c.stream>>>(a, op, gscale, \
                                                                                              partials)
  if (!Op::F.read_g || c.ngrad == 1) GO(1);
  else if (c.ngrad == 2) GO(2);
  else if (c.ngrad == 4) GO(4);
  else if (c.ngrad == 8) GO(8);
  else GO(0);
#undef GO
  CUDA_OK(cudaGetLastError());
  return 1;
}

// update kernels: W is always fp32
template <class Op>
int launch_update(const MtaCall& c, const Op& op, const float* gscale, float* partials) {
  if (c.dtype == kF32) return launch_ng<Op, float, float>(c, op, gscale, partials);
  KV_CHECK(c.dtype == kF16, "unsupported dtype " << c.dtype);
  return launch_ng<Op, __half, float>(c, op, gscale, partials);
}

}  // namespace

int launch_sgd(const MtaCall& c, bool has_mom, const OptScalars& s, const float* gscale) {
  if (c.staged) {
    const int r = launch_sgd_staged(c, has_mom, s, gscale);
    if (r >= 0) return r;
  }
  if (has_mom) {
    SgdOp<true> op{s.momentum, s.rescale, s.clip};
    return launch_update(c, op, gscale, nullptr);
  }
  SgdOp<false> op{0.f, s.rescale, s.clip};
  return launch_update(c, op, gscale, nullptr);
}

int launch_adam(const MtaCall& c, const OptScalars& s, const float* gscale) {
  if (c.staged) {
    const int r = launch_adam_staged(c, s, gscale);
    if (r >= 0) return r;
  }
  AdamOp op{s.beta1, s.beta2, s.eps, s.rescale, s.clip};
  return launch_update(c, op, gscale, nullptr);
}

int launch_adamw(const MtaCall& c, const OptScalars& s, const float* rescale_dev) {
  AdamWOp op{s.beta1, s.beta2, s.eps, s.clip, rescale_dev};
  return launch_update(c, op, nullptr, nullptr);
}

int launch_nonfinite(const MtaCall& c, float* partials) {
  NonFiniteOp op;
  if (c.dtype == kF32) return launch_ng<NonFiniteOp, float, float>(c, op, nullptr, partials);
  KV_CHECK(c.dtype == kF16, "unsupported dtype " << c.dtype);
  return launch_ng<NonFiniteOp, __half, __half>(c, op, nullptr, partials);
}

int launch_all_finite_final(const float* sums, int nt, int init_output, float* out, cudaStream_t stream) {
  all_finite_final_kernel<<<1, 256, 0, stream>>>(sums, nt, init_output, out);
  CUDA_OK(cudaGetLastError());
  return 1;
}

int launch_zero(const MtaCall& c) {
  ZeroOp op;
  if (c.dtype == kF32) return launch_ng<ZeroOp, float, float>(c, op, nullptr, nullptr);
  KV_CHECK(c.dtype == kF16, "unsupported dtype " << c.dtype);
  return launch_ng<ZeroOp, __half, __half>(c, op, nullptr, nullptr);
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

int launch_lamb1(const MtaCall& c, const OptScalars& s, const float* gscale, float* partials) {
  Lamb1Op op{s.beta1, s.beta2, s.eps, s.rescale, s.clip, s.bias_correction};
  return launch_update(c, op, gscale, partials);
}

int launch_lamb2(const MtaCall& c, const OptScalars& s, const float* ratio) {
  Lamb2Op op{s.eps, s.bias_correction, ratio};
  return launch_update(c, op, nullptr, nullptr);
}

int launch_sumsq(const MtaCall& c, float* partials) {
  SumSqOp op;
  if (c.dtype == kF32) return launch_ng<SumSqOp, float, float>(c, op, nullptr, partials);
  KV_CHECK(c.dtype == kF16, "unsupported dtype " << c.dtype);
  return launch_ng<SumSqOp, __half, __half>(c, op, nullptr, partials);
}

int launch_copy(const MtaCall& c, bool to_w, bool to_out, bool w_is_f32) {
  if (c.staged && to_w && !to_out && !w_is_f32) {
    const int r = launch_copy_staged(c);
    if (r >= 0) return r;
  }
#define GO(TW, TO)                                                                          \
  do {                                                                                      \
    CopyOp<TW, TO> op;                                                                      \
    if (c.dtype == kF32) return launch_ng<CopyOp<TW, TO>, float, float>(c, op, nullptr, nullptr); \
    KV_CHECK(c.dtype == kF16, "unsupported dtype " << c.dtype);                             \
    if (w_is_f32) return launch_ng<CopyOp<TW, TO>, __half, float>(c, op, nullptr, nullptr);  \
    return launch_ng<CopyOp<TW, TO>, __half, __half>(c, op, nullptr, nullptr);               \
  } while (0)
  if (to_w && to_out) GO(true, true);
  if (to_w) GO(true, false);
  KV_CHECK(to_out, "launch_copy: nothing to write");
  GO(false, true);
#undef GO
}

int launch_scale(const MtaCall& c, const float* gscale) {
  CopyOp<false, true> op;
  if (c.dtype == kF32) return launch_ng<CopyOp<false, true>, float, float>(c, op, gscale, nullptr);
  KV_CHECK(c.dtype == kF16, "unsupported dtype " << c.dtype);
  return launch_ng<CopyOp<false, true>, __half, __half>(c, op, gscale, nullptr);
}

int launch_master_to_out(const MtaCall& c) {
  MasterToOutOp op;
  KV_CHECK(c.dtype == kF16, "master_to_out is for fp16 outputs");
  return launch_ng<MasterToOutOp, __half, float>(c, op, nullptr, nullptr);
}

int launch_p2p_barrier(const P2PFlags& f, int which, cudaStream_t stream) {
  p2p_barrier_kernel<<<1, 32, 0, stream>>>(f, which);
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
