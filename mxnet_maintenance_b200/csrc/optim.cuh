// Multi-tensor update kernels for sm_100a: one launch per optimizer step over a device-side
// pointer/size table (mta.h), 128-bit coalesced loads/stores, chunk-per-block.
//
// Arithmetic is written with explicit round-to-nearest intrinsics (__fmul_rn/__fadd_rn/...) in the
// reference functors' evaluation order, so nvcc never contracts into FMA and fp32 results are
// bit-identical with the reference's CPU kernels:
//   SGD   MultiSGDKernel::Map        src/operator/optimizer_op-inl.h:225-258
//   Adam  AdamUpdateKernel::Map      src/operator/optimizer_op-inl.h:1293-1315
//   LAMB  MultiLAMBKernelStep1/2     src/operator/contrib/multi_lamb.cc:33-118
//   sumsq CalcSumSq                  src/operator/contrib/multi_sum_sq.cc:68-84 (values; order differs)
//   local copy sum  CommCPU::ReduceSumCPU  src/kvstore/comm.h:357-392 (same grouping => same bits)
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "mta.h"

namespace b200kv {

constexpr int kU = 4;  // 128-bit vectors per thread per chunk


// ------------------------------------------------------------------------------------------------
// element-type helpers
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void ld4(const float* p, float (&v)[4]) {
  const float4 x = *reinterpret_cast<const float4*>(p);
  v[0] = x.x; v[1] = x.y; v[2] = x.z; v[3] = x.w;
}
__device__ __forceinline__ void ld4(const __half* p, float (&v)[4]) {
  const uint2 raw = *reinterpret_cast<const uint2*>(p);
  const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&raw.x));
  const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
__device__ __forceinline__ void st4(float* p, const float (&v)[4]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
__device__ __forceinline__ void st4(__half* p, const float (&v)[4]) {
  // fp32 -> fp16 is round-to-nearest-even, like F16C _cvtss_sh(x, 0) in mshadow/half.h:340-349
  const __half2 a = __floats2half2_rn(v[0], v[1]);
  const __half2 b = __floats2half2_rn(v[2], v[3]);
  uint2 raw;
  raw.x = *reinterpret_cast<const uint32_t*>(&a);
  raw.y = *reinterpret_cast<const uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = raw;
}
__device__ __forceinline__ float ld1(const float* p) { return *p; }
__device__ __forceinline__ float ld1(const __half* p) { return __half2float(*p); }
__device__ __forceinline__ void st1(float* p, float v) { *p = v; }
__device__ __forceinline__ void st1(__half* p, float v) { *p = __float2half_rn(v); }

// A 4-element vector as it comes out of memory.  The kernel issues ALL loads of an iteration into
// these before the first conversion or addition touches one, so every load of a thread is in flight
// at once (an fp16 load followed directly by its half->float conversion would stall the next load
// behind a full DRAM round trip).
template <typename T> struct Raw;
template <> struct Raw<float> { float4 v; };
template <> struct Raw<__half> { uint2 v; };
// (Measured, round 2: telling the compiler these pointers are global — __builtin_assume(__isGlobal(p)),
// which turns the generic LD.E.128 / ST.E.128 into LDG / STG — made the fp32 SGD kernel 30% SLOWER on
// B200 (0.109 vs 0.083 ms on ResNet-50), so the accesses stay generic.)
__device__ __forceinline__ void ldraw(const float* p, Raw<float>& r) {
  r.v = *reinterpret_cast<const float4*>(p);
}
__device__ __forceinline__ void ldraw(const __half* p, Raw<__half>& r) {
  r.v = *reinterpret_cast<const uint2*>(p);
}
__device__ __forceinline__ void cvt(const Raw<float>& r, float (&v)[4]) {
  v[0] = r.v.x; v[1] = r.v.y; v[2] = r.v.z; v[3] = r.v.w;
}
__device__ __forceinline__ void cvt(const Raw<__half>& r, float (&v)[4]) {
  const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&r.v.x));
  const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&r.v.y));
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}

// NVLS (NVLink SHARP): a load from a MULTICAST address returns the sum of that element over every
// GPU bound to the multicast object, reduced inside the NVSwitch; a store to it lands on every GPU.
// One such load replaces W-1 peer loads (the reduce-scatter half), one such store W-1 peer stores
// (the all-gather half): NVLink bytes per GPU and direction drop from 2(W-1)/W*S to S.  The order
// of the in-switch additions is not specified, so this path matches the reference within the fp32
// tolerance rather than bit for bit.
__device__ __forceinline__ void mc_ld4(const float* p, float (&v)[4]) {
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]) : "l"(p) : "memory");
}
__device__ __forceinline__ void mc_st4(float* p, const float (&v)[4]) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v[0]), "f"(v[1]),
               "f"(v[2]), "f"(v[3]) : "memory");
}
__device__ __forceinline__ float mc_ld1(const float* p) {
  float v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void mc_st1(float* p, float v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
// fp16 buckets: the switch adds the W half values in fp32 and rounds ONCE (.acc::f32), where the
// reference rounds to half after every add — within the fp16 tolerance (rtol 2e-3), not bit for bit.
// Only vector forms exist (the store pads multicast chunks to whole 4-element vectors).
__device__ __forceinline__ void mc_ld_raw(const __half* p, Raw<__half>& r) {
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v2.f16x2 {%0,%1}, [%2];"
               : "=r"(r.v.x), "=r"(r.v.y) : "l"(p) : "memory");
}
__device__ __forceinline__ void mc_st4(__half* p, const float (&v)[4]) {
  const __half2 a = __floats2half2_rn(v[0], v[1]);
  const __half2 b = __floats2half2_rn(v[2], v[3]);
  asm volatile("multimem.st.relaxed.sys.global.v2.f16x2 [%0], {%1,%2};" ::"l"(p),
               "r"(*reinterpret_cast<const uint32_t*>(&a)), "r"(*reinterpret_cast<const uint32_t*>(&b)) : "memory");
}

// value of x after a round trip through the tensor's element type
template <typename T> __device__ __forceinline__ float round_to(float x);
template <> __device__ __forceinline__ float round_to<float>(float x) { return x; }
template <> __device__ __forceinline__ float round_to<__half>(float x) {
  return __half2float(__float2half_rn(x));
}
// DType add: fp32 add, rounded to the element type after every add (half.h:64-85)
template <typename T> __device__ __forceinline__ float add_as(float a, float b) {
  return round_to<T>(__fadd_rn(a, b));
}

__device__ __forceinline__ float clipf(float x, float b) {  // mshadow_op::clip, mshadow_op.h:1022-1032
  return x > b ? b : (x < -b ? -b : x);
}

__device__ __forceinline__ const uint64_t* sec_grad(const MtaArgs& a, int k) {
  return a.blob + static_cast<size_t>(k) * a.nt;
}
__device__ __forceinline__ const uint64_t* sec_out(const MtaArgs& a, int k) {
  return a.blob + static_cast<size_t>(a.ngrad + k) * a.nt;
}
__device__ __forceinline__ const uint64_t* sec(const MtaArgs& a, int s) {
  return a.blob + static_cast<size_t>(a.ngrad + a.nout + s) * a.nt;
}

// ------------------------------------------------------------------------------------------------
// gradient fetch: NG == 1 is the hot path; NG == 0 means "a.ngrad copies, summed in CommCPU order"
//   acc = g0;  acc += g1 (+ g2 (+ g3 (+ g4)));  acc += g5 + ...       (comm.h:362-391)
// ------------------------------------------------------------------------------------------------
// vectors per thread issued before any use: fewer when many copies are summed, so that all peer
// loads of a vector are in flight together without blowing the register budget
template <int NG> struct Unroll { static constexpr int value = NG <= 2 ? 4 : (NG <= 4 ? 2 : 1); };

// phase 1: issue the loads of every gradient copy of one vector (nothing is converted or added yet)
// MC = false compiles the multicast branch out (an instantiation that never runs over the switch keeps
// its loads free of the `asm volatile` ordering point: fp16 Adam lost 3 % of its bandwidth to it)
template <typename T, int NG, bool MC = true>
__device__ __forceinline__ void grad_load(const MtaArgs& a, const T* const* gp, int64_t e,
                                          Raw<T> (&r)[NG > 0 ? NG : 1]) {
  if constexpr (NG == 1 && !MC) {
    ldraw(gp[0] + e, r[0]);
  } else if constexpr (NG == 1) {
    if (a.sync_mode & kSyncMcGrad) {   // NVLS: the switch returns the sum over all ranks
      if constexpr (sizeof(T) == 4) {
        float v[4];
        mc_ld4(gp[0] + e, v);
        r[0].v = make_float4(v[0], v[1], v[2], v[3]);
      } else {
        mc_ld_raw(gp[0] + e, r[0]);
      }
      return;
    }
    ldraw(gp[0] + e, r[0]);
  } else if constexpr (NG > 1) {
    // all NG loads (local or NVLink peer memory) are issued back to back
#pragma unroll
    for (int k = 0; k < NG; ++k) ldraw(gp[k] + e, r[k]);
  }
}
// phase 2: convert and sum in CommCPU order
template <typename T, int NG>
__device__ __forceinline__ void grad_finish(const MtaArgs& a, int t, int64_t e,
                                            const Raw<T> (&r)[NG > 0 ? NG : 1], float (&g)[4]) {
  if constexpr (NG == 1) {
    cvt(r[0], g);
  } else if constexpr (NG == 0) {
    ld4(reinterpret_cast<const T*>(sec_grad(a, 0)[t]) + e, g);
    for (int i = 1; i < a.ngrad; i += 4) {
      float s[4];
      ld4(reinterpret_cast<const T*>(sec_grad(a, i)[t]) + e, s);
      for (int k = 1; k < 4 && i + k < a.ngrad; ++k) {
        float x[4];
        ld4(reinterpret_cast<const T*>(sec_grad(a, i + k)[t]) + e, x);
#pragma unroll
        for (int j = 0; j < 4; ++j) s[j] = add_as<T>(s[j], x[j]);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) g[j] = add_as<T>(g[j], s[j]);
    }
  } else {
    float x[NG][4];
#pragma unroll
    for (int k = 0; k < NG; ++k) cvt(r[k], x[k]);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float acc = x[0][j];
#pragma unroll
      for (int i = 1; i < NG; i += 4) {
        float sum = x[i][j];
#pragma unroll
        for (int k = 1; k < 4; ++k)
          if (i + k < NG) sum = add_as<T>(sum, x[i + k][j]);
        acc = add_as<T>(acc, sum);
      }
      g[j] = acc;
    }
  }
}
template <typename T, int NG>
__device__ __forceinline__ float grad1(const MtaArgs& a, int t, int64_t e) {
  if constexpr (NG == 1 && sizeof(T) == 4) {
    if (a.sync_mode & kSyncMcGrad) return mc_ld1(reinterpret_cast<const T*>(sec_grad(a, 0)[t]) + e);
  } else if constexpr (NG == 1) {
    if (a.sync_mode & kSyncMcGrad) __trap();   // no scalar fp16 multimem form; multicast chunks are whole vectors
  }
  float g = ld1(reinterpret_cast<const T*>(sec_grad(a, 0)[t]) + e);
  if (NG != 1) {
    for (int i = 1; i < a.ngrad; i += 4) {
      float s = ld1(reinterpret_cast<const T*>(sec_grad(a, i)[t]) + e);
      for (int k = 1; k < 4 && i + k < a.ngrad; ++k)
        s = add_as<T>(s, ld1(reinterpret_cast<const T*>(sec_grad(a, i + k)[t]) + e));
      g = add_as<T>(g, s);
    }
  }
  return g;
}

// ------------------------------------------------------------------------------------------------
// element operators.  Flags tell the generic kernel which arrays to touch.
// ------------------------------------------------------------------------------------------------
struct OpFlags {
  bool read_g, read_w, write_w, write_out, read_s0, write_s0, read_s1, write_s1;
  int nred;
  // optional fp32 scratch array X (table section kSecX): the element is handed to / taken from
  // elem() through red[kXSlot]
  bool read_x = false, write_x = false;
};
constexpr int kXSlot = 2;

// multi_sgd_update / multi_sgd_mom_update / multi_mp_sgd(_mom)_update
template <bool HAS_MOM>
struct SgdOp {
  static constexpr OpFlags F = {true, true, true, true, HAS_MOM, HAS_MOM, false, false, 0};
  float momentum, rescale, clip;
  struct PT { float lr, lrwd, lrrs; };
  __device__ __forceinline__ PT prep(float lr, float wd, float, float) const {
    return {lr, __fmul_rn(lr, wd), __fmul_rn(lr, rescale)};
  }
  template <typename T>
  __device__ __forceinline__ void elem(const PT& p, float& w, float g, float& s0, float&,
                                       float*) const {
    float mom = HAS_MOM ? s0 : 0.0f;
    const float a = __fmul_rn(momentum, mom);
    const float b = __fmul_rn(p.lrwd, w);
    const float c = clip >= 0.0f ? __fmul_rn(p.lr, clipf(__fmul_rn(rescale, g), clip))
                                 : __fmul_rn(p.lrrs, g);
    mom = __fsub_rn(__fsub_rn(a, b), c);
    s0 = mom;
    w = __fadd_rn(w, mom);
  }
};

// adam_update (bias correction already folded into lr by the host, optimizer.py:1616-1619);
// with fp16 weights/grads this is the fused form of Optimizer.update_multi_precision
// (optimizer.py:298-304): cast grad -> fp32, update master, cast master -> fp16 weight.
struct AdamOp {
  static constexpr OpFlags F = {true, true, true, true, true, true, true, true, 0};
  float beta1, beta2, eps, rescale, clip;
  struct PT { float lr, wd, omb1, omb2; };
  __device__ __forceinline__ PT prep(float lr, float wd, float, float) const {
    return {lr, wd, __fsub_rn(1.f, beta1), __fsub_rn(1.f, beta2)};
  }
  template <typename T>
  __device__ __forceinline__ void elem(const PT& p, float& w, float g, float& m, float& v,
                                       float*) const {
    float gr = __fadd_rn(__fmul_rn(g, rescale), __fmul_rn(w, p.wd));
    if (clip >= 0.f) gr = clipf(gr, clip);
    m = __fadd_rn(__fmul_rn(beta1, m), __fmul_rn(p.omb1, gr));
    v = __fadd_rn(__fmul_rn(beta2, v), __fmul_rn(__fmul_rn(p.omb2, gr), gr));
    w = __fsub_rn(w, __fdiv_rn(__fmul_rn(p.lr, m), __fadd_rn(__fsqrt_rn(v), eps)));
  }
};

// _multi_adamw_update / _multi_mp_adamw_update: MultiMPAdamWKernel::Map
// (src/operator/contrib/adamw-inl.h:346-377).  aux0 = eta.  rescale_grad is a DEVICE scalar; a
// non-finite or zero value skips the update (MPUpdate, adamw-inl.h:447-460) — decided on the
// device, where the reference copies the scalar to the host and synchronises (adamw.cu:32-41).
struct AdamWOp {
  static constexpr OpFlags F = {true, true, true, true, true, true, true, true, 0};
  float beta1, beta2, eps, clip;
  const float* rescale;
  struct PT { float lr, wd, eta, rs; bool skip; };
  __device__ __forceinline__ PT prep(float lr, float wd, float a0, float) const {
    const float rs = *rescale;
    return {lr, wd, a0, rs, !isfinite(rs) || rs == 0.f};
  }
  template <typename T>
  __device__ __forceinline__ void elem(const PT& p, float& w, float g, float& m, float& v,
                                       float*) const {
    if (p.skip) return;
    float sg = __fmul_rn(p.rs, g);
    if (clip >= 0.0f) sg = clipf(sg, clip);
    const float mean = __fadd_rn(__fmul_rn(beta1, __fsub_rn(m, sg)), sg);
    const float adj = __fmul_rn(sg, sg);
    const float var = __fadd_rn(__fmul_rn(beta2, __fsub_rn(v, adj)), adj);
    m = mean;
    v = var;
    const float upd = __fadd_rn(__fdiv_rn(__fmul_rn(p.lr, mean), __fadd_rn(__fsqrt_rn(var), eps)),
                                __fmul_rn(p.wd, w));
    w = __fsub_rn(w, __fmul_rn(p.eta, upd));
  }
};

// multi_all_finite (src/operator/contrib/all_finite.cc:64-95): counts non-finite values
struct NonFiniteOp {
  static constexpr OpFlags F = {true, false, false, false, false, false, false, false, 1};
  struct PT {};
  __device__ __forceinline__ PT prep(float, float, float, float) const { return {}; }
  template <typename T>
  __device__ __forceinline__ void elem(const PT&, float&, float g, float&, float&, float* red) const {
    if (!isfinite(g)) red[0] = __fadd_rn(red[0], 1.0f);
  }
};

// reset_arrays (src/operator/contrib/reset_arrays.cc): zero every array (the out copies)
struct ZeroOp {
  static constexpr OpFlags F = {false, false, false, true, false, false, false, false, 0};
  struct PT {};
  __device__ __forceinline__ PT prep(float, float, float, float) const { return {}; }
  template <typename T>
  __device__ __forceinline__ void elem(const PT&, float& w, float, float&, float&, float*) const { w = 0.f; }
};

// multi_lars: MultiLARSKernel::Map (src/operator/contrib/multi_lars-inl.h:60-72)
static __global__ void multi_lars_kernel(int n, const float* __restrict__ lrs, const float* __restrict__ w2,
                                         const float* __restrict__ g2, const float* __restrict__ wds,
                                         float eta, float eps, float rescale, float* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float w_norm = __fsqrt_rn(w2[i]);
  const bool valid = w_norm > 0.f && g2[i] > 0.f;
  const float num = __fmul_rn(__fmul_rn(lrs[i], eta), w_norm);
  const float den = __fadd_rn(__fadd_rn(__fmul_rn(__fsqrt_rn(g2[i]), rescale), __fmul_rn(wds[i], w_norm)), eps);
  out[i] = valid ? __fdiv_rn(num, den) : lrs[i];
}

// preloaded_multi_*sgd*: learning rates / weight decays come from device arrays
// (src/operator/contrib/preloaded_multi_sgd-inl.h:160-195): patch the table's (lr, wd) section
static __global__ void fill_lrwd_kernel(uint64_t* __restrict__ section, const float* __restrict__ lrs,
                                        const float* __restrict__ wds, int nt) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nt) reinterpret_cast<float2*>(section)[t] = make_float2(lrs[t], wds[t]);
}

// out[0] = 0 when any count is positive; = 1 first when init_output (all_finite.cc:84-86)
static __global__ void all_finite_final_kernel(const float* __restrict__ sums, int nt, int init_output,
                                               float* __restrict__ out) {
  __shared__ int bad;
  if (threadIdx.x == 0) bad = 0;
  __syncthreads();
  for (int t = threadIdx.x; t < nt; t += blockDim.x)
    if (sums[t] > 0.f) bad = 1;
  __syncthreads();
  if (threadIdx.x == 0) {
    if (bad) out[0] = 0.f;
    else if (init_output) out[0] = 1.f;
  }
}

// sgd_update / mp_sgd_update: SGDKernel / MP_SGDKernel (src/operator/optimizer_op-inl.h:382-399,
// 656-679) — the momentum-free single-tensor form (1 - lr*wd)*w - lr*clip(rescale*g) | (lr*rescale)*g,
// which rounds differently from MultiSGDKernel's mom = 0 form
struct SgdPlainOp {
  static constexpr OpFlags F = {true, true, true, true, false, false, false, false, 0};
  float rescale, clip;
  struct PT { float lr, keep, lrrs; };
  __device__ __forceinline__ PT prep(float lr, float wd, float, float) const {
    return {lr, __fsub_rn(1.f, __fmul_rn(lr, wd)), __fmul_rn(lr, rescale)};
  }
  template <typename T>
  __device__ __forceinline__ void elem(const PT& p, float& w, float g, float&, float&, float*) const {
    const float step = clip >= 0.0f ? __fmul_rn(p.lr, clipf(__fmul_rn(rescale, g), clip)) : __fmul_rn(p.lrrs, g);
    w = __fsub_rn(__fmul_rn(p.keep, w), step);
  }
};

// lamb_update_phase1: LambUpdatePhaseOneKernel (optimizer_op-inl.h:1622-1649), fp32.  aux0 / aux1 =
// beta1^t / beta2^t (std::pow on the host, :1661-1662).  Note the reference's `1.` literal: the
// mean's bias correction divides in DOUBLE, the variance's in float.  The update direction g goes to
// the X section.
struct LambPhase1Op {
  static constexpr OpFlags F = {true, true, false, false, true, true, true, true, 0, false, true};
  float beta1, beta2, eps, rescale, clip;
  int bias_correction;
  struct PT { float wd, b1t, b2t, omb1, omb2; };
  __device__ __forceinline__ PT prep(float, float wd, float a0, float a1) const {
    return {wd, a0, a1, __fsub_rn(1.f, beta1), __fsub_rn(1.f, beta2)};
  }
  template <typename T>
  __device__ __forceinline__ void elem(const PT& p, float& w, float g, float& m, float& v, float* red) const {
    float gr = __fmul_rn(g, rescale);
    if (clip >= 0.f) gr = clipf(gr, clip);
    m = __fadd_rn(__fmul_rn(beta1, m), __fmul_rn(p.omb1, gr));
    v = __fadd_rn(__fmul_rn(beta2, v), __fmul_rn(__fmul_rn(p.omb2, gr), gr));
    float u;
    if (bias_correction) {
      const float mh = static_cast<float>(__ddiv_rn(static_cast<double>(m), __dsub_rn(1.0, static_cast<double>(p.b1t))));
      const float vh = __fdiv_rn(v, __fsub_rn(1.f, p.b2t));
      u = __fadd_rn(__fdiv_rn(mh, __fadd_rn(__fsqrt_rn(vh), eps)), __fmul_rn(p.wd, w));
    } else {
      u = __fadd_rn(__fdiv_rn(m, __fadd_rn(__fsqrt_rn(v), eps)), __fmul_rn(p.wd, w));
    }
    red[kXSlot] = u;
  }
};

// lamb_update_phase2: LambUpdatePhaseTwoKernel (optimizer_op-inl.h:1706-1730): r1, r2 are DEVICE
// scalars (the norms of weight and g); lr' = lr * clamp(r1) / r2 unless either is 0; w -= lr' * g
struct LambPhase2Op {
  static constexpr OpFlags F = {false, true, true, true, false, false, false, false, 0, true, false};
  const float* r1;
  const float* r2;
  float lower, upper;
  struct PT { float lr; };
  __device__ __forceinline__ PT prep(float lr, float, float, float) const {
    float n1 = *r1;
    const float n2 = *r2;
    if (lower >= 0.f) n1 = fmaxf(n1, lower);
    if (upper >= 0.f) n1 = fminf(n1, upper);
    if (!(n1 == 0.0f || n2 == 0.0f)) lr = __fdiv_rn(__fmul_rn(lr, n1), n2);
    return {lr};
  }
  template <typename T>
  __device__ __forceinline__ void elem(const PT& p, float& w, float, float&, float&, float* red) const {
    w = __fsub_rn(w, __fmul_rn(p.lr, red[kXSlot]));
  }
};

// AdaGrad, dense branch of AdaGrad.update (python/mxnet/optimizer/optimizer.py:1681-1687), the same
// chain of fp32 operations the NDArray expressions perform:
//   g = g*rescale; clip; history += g*g; div = g / sqrt(history + eps); w += (div + w*wd) * -lr
struct AdagradOp {
  static constexpr OpFlags F = {true, true, true, true, true, true, false, false, 0};
  float eps, rescale, clip;
  struct PT { float neg_lr, wd; };
  __device__ __forceinline__ PT prep(float lr, float wd, float, float) const { return {-lr, wd}; }
  template <typename T>
  __device__ __forceinline__ void elem(const PT& p, float& w, float g, float& h, float&, float*) const {
    float gr = __fmul_rn(g, rescale);
    if (clip >= 0.f) gr = clipf(gr, clip);
    h = __fadd_rn(h, __fmul_rn(gr, gr));
    const float div = __fdiv_rn(gr, __fsqrt_rn(__fadd_rn(h, eps)));
    w = __fadd_rn(w, __fmul_rn(__fadd_rn(div, __fmul_rn(w, p.wd)), p.neg_lr));
  }
};

// LAMB update direction u = m_hat/(sqrt(v_hat)+eps) + wd*w   (multi_lamb.cc:56-72)
// aux0 = 1 - beta1^t, aux1 = 1 - beta2^t (computed with the host's powf, as the CPU kernel does)
__device__ __forceinline__ float lamb_dir(float m, float v, float w, float wd, float c1, float c2,
                                          float eps, bool bias_correction) {
  if (bias_correction) {
    const float mh = __fdiv_rn(m, c1);
    const float vh = __fdiv_rn(v, c2);
    return __fadd_rn(__fdiv_rn(mh, __fadd_rn(__fsqrt_rn(vh), eps)), __fmul_rn(wd, w));
  }
  return __fadd_rn(__fdiv_rn(m, __fadd_rn(__fsqrt_rn(v), eps)), __fmul_rn(wd, w));
}

// pass 1: moments, and per-chunk partial sums of w^2 (in the weight dtype, multi_lamb-inl.h:286-289)
// and u^2.  STORE_U: u is also written to the X scratch section (the reference's temp_g) and pass
// 2 (Lamb2UOp) streams w, u -> w; otherwise pass 2 (Lamb2Op) recomputes u from the new m, v and the
// unchanged w — no scratch memory, more arithmetic and one more array read.
template <bool STORE_U>
struct Lamb1Op {
  static constexpr OpFlags F = {true, true, false, false, true, true, true, true, 2, false, STORE_U};
  float beta1, beta2, eps, rescale, clip;
  int bias_correction;
  struct PT { float wd, c1, c2, omb1, omb2; };
  __device__ __forceinline__ PT prep(float, float wd, float a0, float a1) const {
    return {wd, a0, a1, __fsub_rn(1.f, beta1), __fsub_rn(1.f, beta2)};
  }
  template <typename T>
  __device__ __forceinline__ void elem(const PT& p, float& w, float g, float& m, float& v,
                                       float* red) const {
    float sg = __fmul_rn(g, rescale);
    if (clip >= 0.0f) sg = clipf(sg, clip);
    m = __fadd_rn(__fmul_rn(beta1, m), __fmul_rn(p.omb1, sg));
    v = __fadd_rn(__fmul_rn(beta2, v), __fmul_rn(__fmul_rn(p.omb2, sg), sg));
    const float u = lamb_dir(m, v, w, p.wd, p.c1, p.c2, eps, bias_correction != 0);
    const float wt = round_to<T>(w);
    red[0] = __fadd_rn(red[0], round_to<T>(__fmul_rn(wt, wt)));
    red[1] = __fadd_rn(red[1], __fmul_rn(u, u));
    if (STORE_U) red[kXSlot] = u;   // kept for pass 2 (the reference's temp_g, multi_lamb.cc:77)
  }
};

// pass 2: w -= lr * ratio * u   (multi_lamb.cc:79-118).  aux0/aux1 as in pass 1; the per-tensor
// trust ratio is read from `ratio[t]` (device array produced by lamb_ratio_kernel).
struct Lamb2Op {
  static constexpr OpFlags F = {false, true, true, true, true, false, true, false, 0};
  float eps;
  int bias_correction;
  const float* ratio;
  struct PT { float wd, c1, c2, lr_adj; };
  __device__ __forceinline__ PT prep_t(float lr, float wd, float a0, float a1, int t) const {
    return {wd, a0, a1, __fmul_rn(lr, ratio[t])};
  }
  template <typename T>
  __device__ __forceinline__ void elem(const PT& p, float& w, float, float& m, float& v,
                                       float*) const {
    const float u = lamb_dir(m, v, w, p.wd, p.c1, p.c2, eps, bias_correction != 0);
    w = __fsub_rn(w, __fmul_rn(p.lr_adj, u));
  }
};

// pass 2 when pass 1 stored u: w -= lr * ratio * u, a pure stream (reads w, u; writes w)
struct Lamb2UOp {
  static constexpr OpFlags F = {false, true, true, true, false, false, false, false, 0, true, false};
  const float* ratio;
  struct PT { float lr_adj; };
  __device__ __forceinline__ PT prep_t(float lr, float, float, float, int t) const {
    return {__fmul_rn(lr, ratio[t])};
  }
  template <typename T>
  __device__ __forceinline__ void elem(const PT& p, float& w, float, float&, float&, float* red) const {
    w = __fsub_rn(w, __fmul_rn(p.lr_adj, red[kXSlot]));
  }
};

// multi_sum_sq: per-chunk partial sums of g^2 (product rounded to the element type first, as
// half_t * half_t does on the CPU path)
struct SumSqOp {
  static constexpr OpFlags F = {true, false, false, false, false, false, false, false, 1};
  struct PT {};
  __device__ __forceinline__ PT prep(float, float, float, float) const { return {}; }
  template <typename T>
  __device__ __forceinline__ void elem(const PT&, float&, float g, float&, float&,
                                       float* red) const {
    red[0] = __fadd_rn(red[0], round_to<T>(__fmul_rn(g, g)));
  }
};

// copy (pack into a flat bucket / assign merged value / unpack to the device copies): w = g
template <bool TO_W, bool TO_OUT>
struct CopyOp {
  static constexpr OpFlags F = {true, false, TO_W, TO_OUT, false, false, false, false, 0};
  struct PT {};
  __device__ __forceinline__ PT prep(float, float, float, float) const { return {}; }
  template <typename T>
  __device__ __forceinline__ void elem(const PT&, float& w, float g, float&, float&, float*) const {
    w = g;
  }
};

// out copies <- W section (fp32 master -> weight dtype), nothing else touched
struct MasterToOutOp {
  static constexpr OpFlags F = {false, true, false, true, false, false, false, false, 0};
  struct PT {};
  __device__ __forceinline__ PT prep(float, float, float, float) const { return {}; }
  template <typename T>
  __device__ __forceinline__ void elem(const PT&, float&, float, float&, float&, float*) const {}
};

template <class Op, class = void>
struct HasPrepT { static constexpr bool value = false; };
template <class Op>
struct HasPrepT<Op, decltype((void)&Op::prep_t)> { static constexpr bool value = true; };

// fp16 buckets over the NVSwitch (one fp32-accumulated sum, rounded once) stay inside the fp16 tolerance
// only where the step is linear in the gradient; Adam / AdaGrad normalise by it and never take that
// exchange (Store::run_class_p2p), so their fp16 instantiations carry no multicast code
template <class Op> struct Mc16 { static constexpr bool value = true; };
template <> struct Mc16<AdamOp> { static constexpr bool value = false; };
template <> struct Mc16<AdagradOp> { static constexpr bool value = false; };

template <class Op>
__device__ __forceinline__ typename Op::PT op_prep(const Op& op, float lr, float wd, float a0,
                                                   float a1, int t) {
  if constexpr (HasPrepT<Op>::value) return op.prep_t(lr, wd, a0, a1, t);
  else return op.prep(lr, wd, a0, a1);
}

// ------------------------------------------------------------------------------------------------
// The generic chunk kernel.
//   T   element type of gradients and of the `out` copies (float | __half)
//   WT  element type of the W section (float master/weight, or T for pack/assign buckets)
//   NG  1 = single gradient source, 0 = a.ngrad sources summed in CommCPU order
// gscale (optional device scalar): global-norm clip factor, g = g*scale rounded in T like
// `arr *= scale` in gluon/utils.py:155-156.
// partials: [nchunks][nred] per-chunk reduction outputs (deterministic two-stage reduction, like
// multi_sum_sq.cu's no-atomics design, NEWS.md:795).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}


__device__ __forceinline__ void spin_until(const uint32_t* p, uint32_t epoch, uint32_t* err) {
  const long long t0 = clock64();
  while (static_cast<int32_t>(ld_acquire_sys(p) - epoch) < 0) {
    if (clock64() - t0 > 40000000000LL) {   // ~20 s: give up rather than hang the GPU
      *err = 1;
      break;
    }
  }
}

template <class Op, typename T, typename WT, int NG>
__global__ void __launch_bounds__(kThreads, 2)
mta_kernel(const MtaArgs a, const Op op, const float* __restrict__ gscale,
           float* __restrict__ partials) {
  constexpr OpFlags F = Op::F;
  constexpr bool kMc = sizeof(T) == 4 || Mc16<Op>::value;   // may this instantiation run over the NVSwitch?
  __shared__ float sred[2][kThreads / 32];
  const int tid = threadIdx.x;
  const float gs = (F.read_g && gscale != nullptr) ? *gscale : 1.0f;
  const bool use_gs = F.read_g && gscale != nullptr;
  float red[3] = {0.f, 0.f, 0.f};

  if (a.sync != nullptr && (a.sync_mode & 1)) {
    // every rank's gradient bucket must be complete before this rank reads its shard of it
    const P2PFlags& f = *a.sync;
    if (tid < f.world) {
      if (blockIdx.x == 0) st_release_sys(f.flags[tid] + f.rank, a.epoch);
      spin_until(f.flags[f.rank] + tid, a.epoch, f.flags[f.rank] + 16);
    }
    __syncthreads();
  }

  for (int c = blockIdx.x; c < a.nchunks; c += gridDim.x) {
    const Chunk ch = a.chunks[c];
    const int t = ch.tensor;
    const int64_t base = ch.off;
    WT* const wp = reinterpret_cast<WT*>(sec(a, kSecW)[t]);
    float* const s0p = reinterpret_cast<float*>(sec(a, kSecS0)[t]);
    float* const s1p = reinterpret_cast<float*>(sec(a, kSecS1)[t]);
    float* const xp = reinterpret_cast<float*>(sec(a, kSecX)[t]);
    const float2 lrwd = *reinterpret_cast<const float2*>(&sec(a, kSecLrWd)[t]);
    const float2 aux = *reinterpret_cast<const float2*>(&sec(a, kSecAux)[t]);
    const typename Op::PT pt = op_prep(op, lrwd.x, lrwd.y, aux.x, aux.y, t);

    // 16-byte alignment (8 for a 2-byte type) of every array at this chunk's first element
    uint64_t bits = 0;
    const uint64_t o4 = static_cast<uint64_t>(base) * 4, oT = static_cast<uint64_t>(base) * sizeof(T),
                   oW = static_cast<uint64_t>(base) * sizeof(WT);
    constexpr uint64_t mT = sizeof(T) == 2 ? 2 : 1, mW = sizeof(WT) == 2 ? 2 : 1;
    if (F.read_w || F.write_w) bits |= (reinterpret_cast<uint64_t>(wp) + oW) * mW;
    if (F.read_s0 || F.write_s0) bits |= reinterpret_cast<uint64_t>(s0p) + o4;
    if (F.read_s1 || F.write_s1) bits |= reinterpret_cast<uint64_t>(s1p) + o4;
    if (F.read_x || F.write_x) bits |= reinterpret_cast<uint64_t>(xp) + o4;
    constexpr int NGC = NG > 0 ? NG : 1;
    const T* gp[NGC];
    if (F.read_g) {
#pragma unroll
      for (int k = 0; k < NGC; ++k) gp[k] = reinterpret_cast<const T*>(sec_grad(a, k)[t]);
      for (int k = 0; k < a.ngrad; ++k) bits |= (sec_grad(a, k)[t] + oT) * mT;
    }
    if (F.write_out)
      for (int k = 0; k < a.nout; ++k) bits |= (sec_out(a, k)[t] + oT) * mT;
    const bool vec = (bits & 15) == 0;

    if (vec) {
      const int nvec = ch.len >> 2;
      constexpr int U = Unroll<NG>::value;
      for (int v0 = 0; v0 < nvec; v0 += U * kThreads) {
      float w[U][4], g[U][4], s0[U][4], s1[U][4], x[U][4];
      Raw<T> gr[U][NGC];
      Raw<WT> wr[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int v = v0 + u * kThreads + tid;
        if (v < nvec) {
          const int64_t e = base + 4 * v;
          if (F.read_x) ld4(xp + e, x[u]);
          if (F.read_g) grad_load<T, NG, kMc>(a, gp, e, gr[u]);
          if (F.read_w) ldraw(wp + e, wr[u]);
          if (F.read_s0) ld4(s0p + e, s0[u]);
          if (F.read_s1) ld4(s1p + e, s1[u]);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int v = v0 + u * kThreads + tid;
        if (v < nvec) {
          const int64_t e = base + 4 * v;
          if (F.read_g) grad_finish<T, NG>(a, t, e, gr[u], g[u]);
          if (F.read_w) cvt(wr[u], w[u]);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (use_gs) g[u][j] = round_to<T>(__fmul_rn(g[u][j], gs));
            if (F.read_x) red[kXSlot] = x[u][j];
            op.template elem<T>(pt, w[u][j], g[u][j], s0[u][j], s1[u][j], red);
            if (F.write_x) x[u][j] = red[kXSlot];
          }
          if (F.write_x) st4(xp + e, x[u]);
          if (F.write_s0) st4(s0p + e, s0[u]);
          if (F.write_s1) st4(s1p + e, s1[u]);
          if (F.write_w) st4(wp + e, w[u]);
          if (F.write_out) {
            bool mc = false;
            if constexpr (kMc) {
              mc = (a.sync_mode & kSyncMcOut) != 0;
              if (mc) mc_st4(reinterpret_cast<T*>(sec_out(a, 0)[t]) + e, w[u]);
            }
            if (!mc)
              for (int k = 0; k < a.nout; ++k) st4(reinterpret_cast<T*>(sec_out(a, k)[t]) + e, w[u]);
          }
        }
      }
      }
    }
    // scalar path: the <4-element tail of a vector chunk, or the whole chunk when unaligned
    for (int i = (vec ? (ch.len & ~3) : 0) + tid; i < ch.len; i += kThreads) {
      const int64_t e = base + i;
      float w = 0.f, g = 0.f, s0 = 0.f, s1 = 0.f;
      if (F.read_g) g = grad1<T, NG>(a, t, e);
      if (use_gs) g = round_to<T>(__fmul_rn(g, gs));
      if (F.read_w) w = ld1(wp + e);
      if (F.read_s0) s0 = s0p[e];
      if (F.read_s1) s1 = s1p[e];
      if (F.read_x) red[kXSlot] = xp[e];
      op.template elem<T>(pt, w, g, s0, s1, red);
      if (F.write_x) xp[e] = red[kXSlot];
      if (F.write_s0) s0p[e] = s0;
      if (F.write_s1) s1p[e] = s1;
      if (F.write_w) st1(wp + e, w);
      if (F.write_out) {
        bool mc = false;
        if constexpr (sizeof(T) == 4) {
          mc = (a.sync_mode & kSyncMcOut) != 0;
          if (mc) mc_st1(reinterpret_cast<T*>(sec_out(a, 0)[t]) + e, w);
        } else if (a.sync_mode & kSyncMcOut) {
          __trap();
        }
        if (!mc)
          for (int k = 0; k < a.nout; ++k) st1(reinterpret_cast<T*>(sec_out(a, k)[t]) + e, w);
      }
    }

    if (F.nred > 0) {
      // A CTA keeps accumulating in registers while its next chunk belongs to the same tensor
      // (writing 0 for this chunk) and reduces across the block only at the end of such a run:
      // big tensors cost one block reduction per CTA instead of one per 16 KB chunk.  The
      // assignment of chunks to CTAs is fixed, so results stay run-to-run identical.
      const int cn = c + gridDim.x;
      const bool run_continues = cn < a.nchunks && a.chunks[cn].tensor == t;
      if (run_continues) {
        if (tid < F.nred) partials[static_cast<size_t>(c) * F.nred + tid] = 0.f;
      } else {
#pragma unroll
        for (int r = 0; r < F.nred; ++r) {
          float x = red[r];
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
          if ((tid & 31) == 0) sred[r][tid >> 5] = x;
          red[r] = 0.f;
        }
        __syncthreads();
        if (tid < F.nred) {
          float x = 0.f;
#pragma unroll
          for (int k = 0; k < kThreads / 32; ++k) x += sred[tid][k];
          partials[static_cast<size_t>(c) * F.nred + tid] = x;
        }
        __syncthreads();
      }
    }
  }

  if (a.sync != nullptr && (a.sync_mode & 2)) {
    // "my shard of the new weights has landed in every peer's buffer": the last CTA to finish
    // signals every rank and waits for theirs, so the grid completes only when the whole
    // all-gather has — the caller's next kernel may then read the weights / refill the bucket.
    const P2PFlags& f = *a.sync;
    __shared__ int s_last;
    __threadfence_system();
    __syncthreads();
    if (tid == 0) {
      uint32_t* counter = f.flags[f.rank] + 17;
      const uint32_t prev = atomicAdd(counter, 1u);
      s_last = prev == gridDim.x - 1;
      if (s_last) *counter = 0;
    }
    __syncthreads();
    if (s_last && tid < f.world) {
      __threadfence_system();
      st_release_sys(f.flags[tid] + 8 + f.rank, a.epoch);
      spin_until(f.flags[f.rank] + 8 + tid, a.epoch, f.flags[f.rank] + 16);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Cross-GPU barriers for the peer-memory (NVLink) path.  Every rank owns a flag block
// [ready[8] | done[8]] that its peers write with system-scope stores; `flags[k]` is rank k's block
// as mapped into this process (cudaIpcOpenMemHandle), flags[rank] the local one.
//   ready: "my gradient bucket for step `epoch` is complete — you may read your shard of it"
//   done : "I have written my shard of the new weights into your weight buffer"
// Each kernel is one warp: lane k signals rank k, then waits for rank k's signal.  The ranks run
// on different GPUs, so waiting here cannot starve the signaller.  A bounded spin (~20 s) turns a
// lost peer into an error flag instead of a hung GPU.
// ------------------------------------------------------------------------------------------------
// Reduce-scatter, push half: every rank stores slice r of its own gradient bucket into rank r's
// staging area (region `rank` of it).  Pure NVLink writes — measured 677 GB/s per direction with
// both directions busy, against 624 GB/s for peer reads and ~590 GB/s for mixed read+write traffic.
// dst[r] = rank r's staging region for this rank (nullptr for r == rank); 16-byte vectors.
struct PushArgs {
  uint4* dst[kMaxCopies];
  const uint4* src;        // this rank's bucket
  int64_t slice_vecs;      // 16-byte vectors per slice
  int world, rank;
};

static __global__ void __launch_bounds__(256) p2p_push_kernel(const PushArgs a) {
  // tile = 256 threads x 4 vectors (16 KB); tiles are dealt round-robin over the W-1 peers so that
  // all links carry traffic all the time
  constexpr int kVecPerTile = 1024;
  const int npeer = a.world - 1;
  const int64_t tiles_per_slice = (a.slice_vecs + kVecPerTile - 1) / kVecPerTile;
  const int64_t ntiles = tiles_per_slice * npeer;
  for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const int pi = static_cast<int>(t % npeer);
    const int r = pi < a.rank ? pi : pi + 1;
    const int64_t v0 = (t / npeer) * kVecPerTile;
    const uint4* src = a.src + static_cast<int64_t>(r) * a.slice_vecs;
    uint4 x[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t v = v0 + u * 256 + threadIdx.x;
      if (v < a.slice_vecs) x[u] = src[v];
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t v = v0 + u * 256 + threadIdx.x;
      if (v < a.slice_vecs) a.dst[r][v] = x[u];
    }
  }
}

// which = 0: ready slots, 1: done slots.  flags block layout: [0..8) ready, [8..16) done, [16] error
static __global__ void p2p_barrier_kernel(const P2PFlags f, int which) {
  const int k = threadIdx.x;
  if (k >= f.world) return;
  __threadfence_system();
  st_release_sys(f.flags[k] + which * 8 + f.rank, f.epoch);
  const uint32_t* mine = f.flags[f.rank] + which * 8 + k;
  spin_until(mine, f.epoch, f.flags[f.rank] + 16);
  __threadfence_system();
}

// Tiny all-reduce over peer memory (replaces an ncclAllReduce of a few hundred floats between LAMB's
// two passes / before the global-norm scale): every rank stores its n partial sums into slot `rank`
// of every rank's exchange area, signals, waits for all, and adds the W slots IN RANK ORDER — so the
// result is deterministic and identical on every rank.  The exchange area lives in the flag
// allocation (byte 4096 onwards): 2 buffers (alternating with the epoch, so a fast rank's next round
// never overwrites what a slow rank is still reading) x 8 ranks x kXsumCap floats.
constexpr int kXsumCap = 16384;
constexpr int kXsumFlagBase = 32;      // words [32, 40) of the flag block
__device__ __forceinline__ float* xsum_area(uint32_t* flags) {
  return reinterpret_cast<float*>(reinterpret_cast<char*>(flags) + 4096);
}
static __global__ void __launch_bounds__(256) xsum_kernel(float* __restrict__ sums, int n, const P2PFlags f) {
  const int tid = threadIdx.x;
  const int buf = f.epoch & 1;
  for (int k = 0; k < f.world; ++k) {
    float* dst = xsum_area(f.flags[k]) + static_cast<size_t>(buf * kMaxCopies + f.rank) * kXsumCap;
    for (int i = tid; i < n; i += 256) dst[i] = sums[i];
  }
  __threadfence_system();
  __syncthreads();
  if (tid < f.world) {
    st_release_sys(f.flags[tid] + kXsumFlagBase + f.rank, f.epoch);
    spin_until(f.flags[f.rank] + kXsumFlagBase + tid, f.epoch, f.flags[f.rank] + 16);
  }
  __syncthreads();
  const float* mine = xsum_area(f.flags[f.rank]) + static_cast<size_t>(buf) * kMaxCopies * kXsumCap;
  for (int i = tid; i < n; i += 256) {
    float acc = __ldcv(mine + i);
    for (int k = 1; k < f.world; ++k) acc = __fadd_rn(acc, __ldcv(mine + static_cast<size_t>(k) * kXsumCap + i));
    sums[i] = acc;
  }
}

// second stage of the deterministic reduction: one warp per tensor sums that tensor's chunk
// partials (chunks of a tensor are consecutive: [first[t], first[t+1])).  sums is [nred][nt];
// accumulate != 0 adds to the existing value (used when a tensor list is processed in pieces).
static __global__ void reduce_partials_kernel(const float* __restrict__ partials,
                                       const int* __restrict__ first, int nt, int nred,
                                       float* __restrict__ sums) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= nt) return;
  const int b = first[warp], e = first[warp + 1];
  for (int r = 0; r < nred; ++r) {
    float x = 0.f;
    for (int c = b + lane; c < e; c += 32) x += partials[static_cast<size_t>(c) * nred + r];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    if (lane == 0) sums[static_cast<size_t>(r) * nt + warp] = x;
  }
}

// LAMB trust ratio per tensor (multi_lamb.cc:92-104): sums = [sum w^2 | sum u^2], each [nt]
static __global__ void lamb_ratio_kernel(const float* __restrict__ sums, int nt, float lower, float upper,
                                  float* __restrict__ ratio) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nt) return;
  float r1 = __fsqrt_rn(sums[t]);
  const float r2 = __fsqrt_rn(sums[nt + t]);
  if (lower >= 0) r1 = fmaxf(r1, lower);
  if (upper >= 0) r1 = fminf(r1, upper);
  ratio[t] = (r1 == 0.0f || r2 == 0.0f) ? 1.0f : __fdiv_rn(r1, r2);
}

// clip_global_norm scalar (gluon/utils.py:149-154): out[0] = scale = min(max_norm/(norm+1e-8), 1),
// out[1] = norm.  One block; per-tensor sums are added in a fixed tree order.
static __global__ void global_norm_kernel(const float* __restrict__ sums, int nt, float max_norm,
                                   float* __restrict__ out) {
  __shared__ float s[kThreads];
  float x = 0.f;
  for (int t = threadIdx.x; t < nt; t += kThreads) x += sums[t];
  s[threadIdx.x] = x;
  __syncthreads();
  for (int o = kThreads / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const float norm = __fsqrt_rn(s[0]);
    const float scale = __fdiv_rn(max_norm, __fadd_rn(norm, 1e-8f));
    out[0] = fminf(scale, 1.0f);
    out[1] = norm;
  }
}

}  // namespace b200kv
