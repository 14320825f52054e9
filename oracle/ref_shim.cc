// TEST INFRASTRUCTURE ONLY — never imported, linked or executed by the product path.
//
// oracle/_ref/libmxref.so: the reference's OWN CPU functors for the gradient-sync + update
// path, compiled from the sources where they lie under /root/reference (nothing is copied
// into this repository).  This file only contains the glue that a `Kernel<OP,cpu>::Launch`
// (src/operator/mxnet_op.h:945-965) would provide: a plain loop `for i in [0,N): OP::Map(i,…)`.
//
// Reference code reached from here:
//   MultiSGDKernel            src/operator/optimizer_op-inl.h:225-258
//   SGDMomKernel              src/operator/optimizer_op-inl.h:601-620
//   MP_SGDMomKernel           src/operator/optimizer_op-inl.h:700-724
//   AdamUpdateKernel          src/operator/optimizer_op-inl.h:1293-1315
//   SGDMomDnsRspDnsKernel<cpu> src/operator/optimizer_op-inl.h:747-774
//   AdamDnsRspDnsKernel<cpu>  src/operator/optimizer_op-inl.h:1352-1382
//   MultiLAMBKernelStep1/2    src/operator/contrib/multi_lamb.cc:33-118
//   CalcSumSq                 src/operator/contrib/multi_sum_sq.cc:68-84
//   SparseRetainRspThreadKernel src/operator/tensor/sparse_retain-inl.h:121-150
//   MultiMPAdamWKernel        src/operator/contrib/adamw-inl.h:346-377
//   MultiLARSKernel           src/operator/contrib/multi_lars-inl.h:60-72
//   AdagradDnsRspDnsKernel<cpu> src/operator/optimizer_op-inl.h:2514-2539
//   mshadow half_t / Tensor<cpu,1,DType> expressions for the CommCPU reduce
//                             (restated driver: src/kvstore/comm.h:357-410)
#include <cstdint>
#include <cstring>
#include <vector>
#include <algorithm>
#include <omp.h>

#include "operator/optimizer_op-inl.h"
#include "operator/tensor/sparse_retain-inl.h"
#include "operator/contrib/adamw-inl.h"
#include "operator/contrib/multi_lars-inl.h"
// The LAMB step functors and CalcSumSq live in .cc files; pull them in as-is.
#include "operator/contrib/multi_sum_sq.cc"
#include "operator/contrib/multi_lamb.cc"
// The standard (non-lazy) sparse update functors are specialised in optimizer_op.cc (:108-229).
#include "operator/optimizer_op.cc"

namespace mxnet {
// The only engine symbol the included TUs reference but never call on our path.
void* Resource::get_space_internal(size_t) const { return nullptr; }
void* Resource::get_host_space_internal(size_t) const { return nullptr; }
}  // namespace mxnet

using namespace mxnet;
using namespace mxnet::op;
using mshadow::half::half_t;

namespace {

// ---- CommCPU::ReduceSumCPU, comm.h:357-392 (same mshadow expressions => same fp16 rounding)
template <typename DType>
void ReduceSumCPU(const std::vector<DType*>& dptr, size_t offset, index_t size) {
  using namespace mshadow;  // NOLINT
  Tensor<cpu, 1, DType> in_0(dptr[0] + offset, Shape1(size));
  for (size_t i = 1; i < dptr.size(); i += 4) {
    switch (dptr.size() - i) {
      case 1: {
        Tensor<cpu, 1, DType> in_1(dptr[i] + offset, Shape1(size));
        in_0 += in_1;
        break;
      }
      case 2: {
        Tensor<cpu, 1, DType> in_1(dptr[i] + offset, Shape1(size));
        Tensor<cpu, 1, DType> in_2(dptr[i + 1] + offset, Shape1(size));
        in_0 += in_1 + in_2;
        break;
      }
      case 3: {
        Tensor<cpu, 1, DType> in_1(dptr[i] + offset, Shape1(size));
        Tensor<cpu, 1, DType> in_2(dptr[i + 1] + offset, Shape1(size));
        Tensor<cpu, 1, DType> in_3(dptr[i + 2] + offset, Shape1(size));
        in_0 += in_1 + in_2 + in_3;
        break;
      }
      default: {
        Tensor<cpu, 1, DType> in_1(dptr[i] + offset, Shape1(size));
        Tensor<cpu, 1, DType> in_2(dptr[i + 1] + offset, Shape1(size));
        Tensor<cpu, 1, DType> in_3(dptr[i + 2] + offset, Shape1(size));
        Tensor<cpu, 1, DType> in_4(dptr[i + 3] + offset, Shape1(size));
        in_0 += in_1 + in_2 + in_3 + in_4;
        break;
      }
    }
  }
}

// comm.h:394-410
template <typename DType>
void ReduceSumCPUImpl(std::vector<DType*> dptr, size_t total, size_t bigarray_bound, int nthread) {
  const size_t step = std::min(bigarray_bound, static_cast<size_t>(4 << 10));
  long ntask = (total + step - 1) / step;  // NOLINT
  if (total < bigarray_bound || nthread <= 1) {
    ReduceSumCPU(dptr, 0, total);
  } else {
#pragma omp parallel for schedule(static) num_threads(nthread)
    for (long j = 0; j < ntask; ++j) {  // NOLINT
      size_t k = static_cast<size_t>(j);
      size_t begin = std::min(k * step, total);
      size_t end = std::min((k + 1) * step, total);
      ReduceSumCPU(dptr, begin, static_cast<index_t>(end - begin));
    }
  }
}

template <typename DType, typename MPDType, bool has_mom, bool has_mp>
void RunMultiSGD(int count, void** weights, void** grads, void** moms, void** w32, void** outs,
                 const int64_t* sizes, const float* lrs, const float* wds, float momentum,
                 float rescale, float clip, int nthread) {
  // optimizer_op-inl.h:207-223 limits one launch to N=60 tensors; batch like the op would.
  for (int base = 0; base < count; base += MultiSGDKernelParam<DType, MPDType>::N) {
    MultiSGDKernelParam<DType, MPDType> p;
    p.count = std::min(count - base, MultiSGDKernelParam<DType, MPDType>::N);
    p.max_size = 0;
    p.clip_gradient = clip;
    p.rescale_grad = rescale;
    p.momentum = momentum;
    for (int i = 0; i < p.count; ++i) {
      p.sizes[i] = sizes[base + i];
      p.max_size = std::max(p.max_size, p.sizes[i]);
      p.weights[i] = static_cast<DType*>(weights[base + i]);
      p.grads[i] = static_cast<DType*>(grads[base + i]);
      p.mom[i] = has_mom ? static_cast<MPDType*>(moms[base + i]) : nullptr;
      p.weights32[i] = has_mp ? static_cast<MPDType*>(w32[base + i]) : nullptr;
      p.out_data[i] = static_cast<DType*>(outs[base + i]);
      p.lrs[i] = lrs[base + i];
      p.wds[i] = wds[base + i];
    }
    const index_t N = p.max_size;
#pragma omp parallel for num_threads(nthread)
    for (index_t i = 0; i < N; ++i) {
      MultiSGDKernel<MPDType, has_mom, has_mp>::Map(i, p, kWriteTo);
    }
  }
}

template <typename DType, typename MPDType, bool has_mp>
void RunMultiLAMB(int count, void** weights, void** grads, void** means, void** vars, void** w32,
                  void** outs, const int64_t* sizes, const float* lrs, const float* wds,
                  const int* steps, float beta1, float beta2, float eps, float rescale, float clip,
                  float lower, float upper, int bias_correction, int nthread,
                  float* r1_out, float* r2_out) {
  using KP = MultiLAMBKernelParam<DType, MPDType>;
  for (int base = 0; base < count; base += KP::N) {  // multi_lamb-inl.h:193 (N=45), check :119
    KP kp;
    kp.ntensors = std::min(count - base, static_cast<int>(KP::N));
    kp.total_size = 0;
    kp.max_size = 0;
    std::vector<TBlob> wblobs, gblobs;
    for (size_t i = 0; i < kp.ntensors; ++i) {
      kp.sizes[i] = sizes[base + i];
      kp.tensor2temp_g[i] = kp.total_size;
      kp.total_size += kp.sizes[i];
      kp.max_size = std::max(kp.max_size, kp.sizes[i]);
      kp.weights[i] = static_cast<DType*>(weights[base + i]);
      kp.grads[i] = static_cast<DType*>(grads[base + i]);
      kp.mean[i] = static_cast<MPDType*>(means[base + i]);
      kp.var[i] = static_cast<MPDType*>(vars[base + i]);
      kp.weights32[i] = has_mp ? static_cast<MPDType*>(w32[base + i]) : nullptr;
      kp.out_data[i] = static_cast<DType*>(outs[base + i]);
      kp.step_count[i] = steps[base + i];
      kp.learning_rates[i] = lrs[base + i];
      kp.wds[i] = wds[base + i];
    }
    std::vector<float> temp_g(kp.total_size), r1(kp.ntensors), r2(kp.ntensors);
    for (size_t i = 0; i < kp.ntensors; ++i) {
      // r1 from the op's `weights` inputs (DType, i.e. fp16 in MP mode): multi_lamb-inl.h:286-289,330
      wblobs.emplace_back(kp.weights[i], mxnet::TShape(mshadow::Shape1(kp.sizes[i])), int(cpu::kDevMask));
      gblobs.emplace_back(temp_g.data() + kp.tensor2temp_g[i],
                          mxnet::TShape(mshadow::Shape1(kp.sizes[i])), int(cpu::kDevMask));
    }
    omp_set_num_threads(nthread);
    CalcSumSq<DType>(wblobs, kp.ntensors, r1.data(), nullptr);
    const int N = static_cast<int>(kp.max_size);
#pragma omp parallel for num_threads(nthread)
    for (int i = 0; i < N; ++i) {
      MultiLAMBKernelStep1<MPDType, has_mp>::Map(i, kp, beta1, beta2, eps, clip,
                                                 bias_correction != 0, rescale, temp_g.data());
    }
    CalcSumSq<float>(gblobs, kp.ntensors, r2.data(), nullptr);
#pragma omp parallel for num_threads(nthread)
    for (int i = 0; i < N; ++i) {
      MultiLAMBKernelStep2<MPDType, has_mp>::Map(i, kp, r1.data(), r2.data(), temp_g.data(), lower,
                                                 upper, kWriteTo);
    }
    if (r1_out) std::memcpy(r1_out + base, r1.data(), kp.ntensors * sizeof(float));
    if (r2_out) std::memcpy(r2_out + base, r2.data(), kp.ntensors * sizeof(float));
  }
}

// _multi_adamw_update / _multi_mp_adamw_update (contrib/adamw-inl.h:326-377; N = 50 tensors per launch)
template <typename DType, typename MPDType, bool has_mp>
static void RunMultiAdamW(int count, void** weights, void** grads, void** means, void** vars, void** w32,
                          void** outs, const int64_t* sizes, const float* lrs, const float* wds,
                          const float* etas, float beta1, float beta2, float eps, float rescale,
                          float clip, int nthread) {
  using KP = MultiAdamKernelParam<DType, MPDType>;
  for (int base = 0; base < count; base += KP::N) {
    KP p;
    p.count = std::min(count - base, static_cast<int>(KP::N));
    p.max_size = 0;
    p.clip_gradient = clip; p.beta1 = beta1; p.beta2 = beta2; p.epsilon = eps;
    for (int i = 0; i < p.count; ++i) {
      p.sizes[i] = sizes[base + i];
      p.max_size = std::max(p.max_size, p.sizes[i]);
      p.weights[i] = static_cast<DType*>(weights[base + i]);
      p.grad_data[i] = static_cast<DType*>(grads[base + i]);
      p.mean_data[i] = static_cast<MPDType*>(means[base + i]);
      p.var_data[i] = static_cast<MPDType*>(vars[base + i]);
      p.weights32[i] = has_mp ? static_cast<MPDType*>(w32[base + i]) : nullptr;
      p.out_data[i] = static_cast<DType*>(outs[base + i]);
      p.etas[i] = etas[base + i]; p.lrs[i] = lrs[base + i]; p.wds[i] = wds[base + i];
    }
    const int N = static_cast<int>(p.max_size);
#pragma omp parallel for num_threads(nthread)
    for (int i = 0; i < N; ++i) MultiMPAdamWKernel<MPDType, has_mp>::Map(i, p, kWriteTo, rescale);
  }
}

}  // namespace

extern "C" {

// dtype: 0 = float32, 2 = float16 (mshadow type flags, 3rdparty/mshadow/mshadow/base.h)
int mxref_reduce_sum(int dtype, int n, void** ptrs, int64_t size, int64_t bigarray_bound,
                     int nthread) {
  if (dtype == 0) {
    std::vector<float*> d(n);
    for (int i = 0; i < n; ++i) d[i] = static_cast<float*>(ptrs[i]);
    ReduceSumCPUImpl<float>(d, size, bigarray_bound, nthread);
  } else if (dtype == 2) {
    std::vector<half_t*> d(n);
    for (int i = 0; i < n; ++i) d[i] = static_cast<half_t*>(ptrs[i]);
    ReduceSumCPUImpl<half_t>(d, size, bigarray_bound, nthread);
  } else {
    return -1;
  }
  return 0;
}

int mxref_multi_sgd(int dtype, int has_mom, int has_mp, int count, void** weights, void** grads,
                    void** moms, void** w32, void** outs, const int64_t* sizes, const float* lrs,
                    const float* wds, float momentum, float rescale, float clip, int nthread) {
#define GO(D, M, HM, HP) \
  RunMultiSGD<D, M, HM, HP>(count, weights, grads, moms, w32, outs, sizes, lrs, wds, momentum, \
                            rescale, clip, nthread)
  if (dtype == 0 && !has_mp) {
    if (has_mom) GO(float, float, true, false); else GO(float, float, false, false);
  } else if (dtype == 2 && has_mp) {
    if (has_mom) GO(half_t, float, true, true); else GO(half_t, float, false, true);
  } else if (dtype == 2 && !has_mp) {
    if (has_mom) GO(half_t, half_t, true, false); else GO(half_t, half_t, false, false);
  } else {
    return -1;
  }
#undef GO
  return 0;
}

// sgd_update / mp_sgd_update: the reference's SGDKernel / MP_SGDKernel (optimizer_op-inl.h:382-399,656-679)
int mxref_sgd_update(int64_t n, void* out, const void* w, const void* g, float* w32, float clip, float lr,
                     float wd, float rescale) {
  for (int64_t i = 0; i < n; ++i) {
    if (w32)
      MP_SGDKernel::Map<half_t>(i, static_cast<half_t*>(out), static_cast<const half_t*>(w),
                                static_cast<const half_t*>(g), w32, clip, lr, wd, rescale, kWriteTo);
    else
      SGDKernel::Map<float>(i, static_cast<float*>(out), static_cast<const float*>(w),
                            static_cast<const float*>(g), clip, lr, wd, rescale, kWriteTo);
  }
  return 0;
}

// lamb_update_phase1 / phase2: LambUpdatePhaseOneKernel / LambUpdatePhaseTwoKernel (optimizer_op-inl.h:1622-1730)
int mxref_lamb_phase1(int64_t n, float* out, float* mean, float* var, const float* w, const float* g,
                      float clip, float rescale, float beta1, float beta2, float wd, float eps, int t,
                      int bias_correction) {
  const float beta1_t = std::pow(beta1, t), beta2_t = std::pow(beta2, t);   // :1661-1662 (float, int -> double pow)
  for (int64_t i = 0; i < n; ++i)
    LambUpdatePhaseOneKernel::Map<float>(i, out, mean, var, w, g, clip, rescale, beta1, beta1_t, beta2, beta2_t,
                                         wd, eps, t, bias_correction != 0, kWriteTo);
  return 0;
}

int mxref_lamb_phase2(int64_t n, float* out, const float* w, const float* g, float r1, float r2, float lr,
                      float lower, float upper) {
  for (int64_t i = 0; i < n; ++i)
    LambUpdatePhaseTwoKernel::Map<float>(i, out, w, g, &r1, &r2, lr, lower, upper, kWriteTo);
  return 0;
}

int mxref_sgd_mom(int64_t n, float* out, float* mom, const float* w, const float* g, float clip,
                  float momentum, float lr, float wd, float rescale) {
  for (int64_t i = 0; i < n; ++i)
    SGDMomKernel::Map<float>(i, out, mom, w, g, clip, momentum, lr, wd, rescale, kWriteTo);
  return 0;
}

int mxref_mp_sgd_mom(int64_t n, void* out16, float* mom, const void* w16, const void* g16,
                     float* w32, float clip, float momentum, float lr, float wd, float rescale) {
  for (int64_t i = 0; i < n; ++i)
    MP_SGDMomKernel::Map<half_t>(i, static_cast<half_t*>(out16), mom,
                                 static_cast<const half_t*>(w16), static_cast<const half_t*>(g16),
                                 w32, clip, momentum, lr, wd, rescale, kWriteTo);
  return 0;
}

int mxref_adam(int64_t n, float* out, float* mean, float* var, const float* w, const float* g,
               float clip, float rescale, float beta1, float beta2, float lr, float wd, float eps,
               int nthread) {
#pragma omp parallel for num_threads(nthread)
  for (int64_t i = 0; i < n; ++i)
    AdamUpdateKernel::Map<float>(i, out, mean, var, w, g, clip, rescale, beta1, beta2, lr, wd, eps,
                                 kWriteTo);
  return 0;
}

int mxref_multi_lamb(int dtype, int has_mp, int count, void** weights, void** grads, void** means,
                     void** vars, void** w32, void** outs, const int64_t* sizes, const float* lrs,
                     const float* wds, const int* steps, float beta1, float beta2, float eps,
                     float rescale, float clip, float lower, float upper, int bias_correction,
                     int nthread, float* r1_out, float* r2_out) {
  if (dtype == 0 && !has_mp)
    RunMultiLAMB<float, float, false>(count, weights, grads, means, vars, w32, outs, sizes, lrs, wds,
                                      steps, beta1, beta2, eps, rescale, clip, lower, upper,
                                      bias_correction, nthread, r1_out, r2_out);
  else if (dtype == 2 && has_mp)
    RunMultiLAMB<half_t, float, true>(count, weights, grads, means, vars, w32, outs, sizes, lrs, wds,
                                      steps, beta1, beta2, eps, rescale, clip, lower, upper,
                                      bias_correction, nthread, r1_out, r2_out);
  else
    return -1;
  return 0;
}

int mxref_multi_adamw(int dtype, int has_mp, int count, void** weights, void** grads, void** means,
                      void** vars, void** w32, void** outs, const int64_t* sizes, const float* lrs,
                      const float* wds, const float* etas, float beta1, float beta2, float eps,
                      float rescale, float clip, int nthread) {
  // MPUpdate (adamw-inl.h:447-460): a non-finite or zero rescale_grad skips the whole update
  if (!std::isfinite(rescale) || rescale == 0) return 0;
  if (dtype == 0 && !has_mp)
    RunMultiAdamW<float, float, false>(count, weights, grads, means, vars, w32, outs, sizes, lrs, wds, etas,
                                       beta1, beta2, eps, rescale, clip, nthread);
  else if (dtype == 2 && has_mp)
    RunMultiAdamW<half_t, float, true>(count, weights, grads, means, vars, w32, outs, sizes, lrs, wds, etas,
                                       beta1, beta2, eps, rescale, clip, nthread);
  else
    return -1;
  return 0;
}

int mxref_multi_lars(int n, float* out, const float* lrs, const float* w2, const float* g2,
                     const float* wds, float eta, float eps, float rescale) {
  for (int i = 0; i < n; ++i) MultiLARSKernel::Map(i, out, lrs, w2, g2, wds, eta, eps, rescale, kWriteTo);
  return 0;
}

int mxref_multi_sum_sq(int dtype, int count, void** arrays, const int64_t* sizes, float* out,
                       int nthread) {
  std::vector<TBlob> blobs;
  for (int i = 0; i < count; ++i) {
    if (dtype == 0)
      blobs.emplace_back(static_cast<float*>(arrays[i]), mxnet::TShape(mshadow::Shape1(sizes[i])),
                         int(cpu::kDevMask));
    else
      blobs.emplace_back(static_cast<half_t*>(arrays[i]), mxnet::TShape(mshadow::Shape1(sizes[i])),
                         int(cpu::kDevMask));
  }
  omp_set_num_threads(nthread);
  if (dtype == 0) CalcSumSq<float>(blobs, count, out, nullptr);
  else if (dtype == 2) CalcSumSq<half_t>(blobs, count, out, nullptr);
  else return -1;
  return 0;
}

// lazy row-sparse SGD-momentum: one Map call per gradient row (optimizer_op-inl.h:747-774)
int mxref_sgd_mom_rsp(int64_t nnr, int64_t row_len, float* out, float* mom, const float* w,
                      const int64_t* idx, const float* g, float clip, float momentum, float lr,
                      float wd, float rescale, int nthread) {
#pragma omp parallel for num_threads(nthread)
  for (int64_t i = 0; i < nnr; ++i)
    SGDMomDnsRspDnsKernel<kWriteTo, cpu>::Map<float, int64_t>(i, row_len, out, mom, w, idx, g, clip,
                                                            momentum, lr, wd, rescale);
  return 0;
}

// standard (non-lazy) updates: the reference functors of optimizer_op.cc take the inclusive prefix sum of
// the row flags, as SGDMomStdUpdateDnsRspDnsImpl<cpu> builds it (optimizer_op.cc:170-190)
static std::vector<nnvm::dim_t> row_prefix(int64_t num_rows, int64_t nnr, const int64_t* idx) {
  std::vector<nnvm::dim_t> p(num_rows, 0);
  for (int64_t i = 0; i < nnr; ++i) p[idx[i]] = 1;
  for (int64_t i = 1; i < num_rows; ++i) p[i] += p[i - 1];
  return p;
}

int mxref_sgd_mom_std_rsp(int64_t num_rows, int64_t nnr, int64_t row_len, float* out, float* mom, const float* w,
                          const int64_t* idx, const float* g, float clip, float momentum, float lr, float wd,
                          float rescale, int nthread) {
  const std::vector<nnvm::dim_t> p = row_prefix(num_rows, nnr, idx);
#pragma omp parallel for num_threads(nthread)
  for (int64_t i = 0; i < num_rows; ++i)
    SGDMomStdDnsRspDnsKernel<kWriteTo, cpu>::Map<float, int64_t, nnvm::dim_t>(
        static_cast<int>(i), row_len, out, mom, w, idx, g, p.data(), clip, momentum, lr, wd, rescale);
  return 0;
}

int mxref_adam_std_rsp(int64_t num_rows, int64_t nnr, int64_t row_len, float* out, float* mean, float* var,
                       const float* w, const int64_t* idx, const float* g, float clip, float beta1, float beta2,
                       float lr, float wd, float eps, float rescale, int nthread) {
  const std::vector<nnvm::dim_t> p = row_prefix(num_rows, nnr, idx);
#pragma omp parallel for num_threads(nthread)
  for (int64_t i = 0; i < num_rows; ++i)
    AdamStdDnsRspDnsKernel<kWriteTo, cpu>::Map<float, int64_t, nnvm::dim_t>(
        static_cast<int>(i), row_len, out, mean, var, w, idx, g, p.data(), clip, beta1, beta2, lr, wd, eps, rescale);
  return 0;
}

// lazy row-sparse SGD without momentum (optimizer_op-inl.h:452-475)
int mxref_sgd_rsp(int64_t nnr, int64_t row_len, float* out, const float* w, const int64_t* idx,
                  const float* g, float clip, float lr, float wd, float rescale, int nthread) {
#pragma omp parallel for num_threads(nthread)
  for (int64_t i = 0; i < nnr; ++i)
    SGDDnsRspKernel<kWriteTo, cpu>::Map<float, int64_t>(i, row_len, out, w, idx, g, clip, lr, wd, rescale);
  return 0;
}

int mxref_adam_rsp(int64_t nnr, int64_t row_len, float* out, float* mean, float* var,
                   const float* w, const int64_t* idx, const float* g, float clip, float beta1,
                   float beta2, float lr, float wd, float eps, float rescale, int nthread) {
#pragma omp parallel for num_threads(nthread)
  for (int64_t i = 0; i < nnr; ++i)
    AdamDnsRspDnsKernel<kWriteTo, cpu>::Map<float, int64_t>(i, row_len, out, mean, var, w, idx, g,
                                                          clip, beta1, beta2, lr, wd, eps, rescale);
  return 0;
}

// _sparse_adagrad_update, lazy rows (optimizer_op-inl.h:2514-2539; requires wd == 0, :2497-2502)
int mxref_adagrad_rsp(int64_t nnr, int64_t row_len, float* out, float* hist, const float* w,
                      const int64_t* idx, const float* g, float clip, float eps, float lr, float rescale,
                      int nthread) {
#pragma omp parallel for num_threads(nthread)
  for (int64_t i = 0; i < nnr; ++i)
    AdagradDnsRspDnsKernel<cpu>::Map<float, int64_t>(i, row_len, out, hist, w, idx, g, clip, eps, lr, rescale);
  return 0;
}

// out_data must be zero-filled by the caller (sparse_retain-inl.h:296-300 does the same)
int mxref_sparse_retain(int64_t nids, float* out_data, int64_t* out_idx, const float* in_data,
                        const int64_t* in_idx, const int64_t* ids, int64_t nnr, int64_t row_len) {
  for (int64_t i = 0; i < nids; ++i)
    SparseRetainRspThreadKernel::Map(static_cast<int>(i), out_data, out_idx, in_data, in_idx, ids,
                                     static_cast<size_t>(nnr), static_cast<size_t>(row_len));
  return 0;
}

// fp32 <-> fp16 through mshadow::half::half_t (3rdparty/mshadow/mshadow/half.h)
void mxref_f32_to_f16(int64_t n, const float* in, void* out) {
  half_t* o = static_cast<half_t*>(out);
  for (int64_t i = 0; i < n; ++i) o[i] = half_t(in[i]);
}

}  // extern "C"
