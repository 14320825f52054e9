// Gradient-staged variant of the multi-tensor kernel for the peer-memory (NVLink) path.
//
// When a rank sums its shard of W gradient buckets that live on W different GPUs, a load from a
// peer takes several microseconds; in the plain kernel every thread sits on its local weight /
// momentum registers until the slowest peer load returns, so neither HBM nor NVLink is kept busy.
// Here one elected thread streams the gradient copies (local and peer) into a shared-memory ring
// with TMA bulk copies (cp.async.bulk + mbarrier complete_tx), several stages ahead of the math;
// the 256 consumer threads issue their local loads, wait on the stage's mbarrier, sum the copies
// from shared memory in the reference's CommCPU order, run the optimizer functor and store.  The
// new weights go to every peer (the all-gather half of the fused step) the same way in reverse:
// they are collected in a small shared-memory ring and pushed out with TMA bulk stores
// (cp.async.bulk.global.shared::cta), because ordinary st.global to a peer holds a load/store-unit
// slot for the whole NVLink round trip and stalls the local traffic behind it (measured: 157 GB/s
// with 8-byte and 310 GB/s with 16-byte peer stores against ~740 GB/s for bulk copies).
#include "optim.cuh"

#include "launch.h"

namespace b200kv {

namespace {

constexpr int kSub = 1024;   // elements per copy per stage (256 threads x one 4-element vector)

template <int NG> struct Stages { static constexpr int value = NG <= 2 ? 8 : (NG <= 4 ? 6 : 4); };

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!ok);
}
// 1-D TMA bulk copy global (local or peer) -> shared, completion counted in bytes on `bar`
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
                   "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// 1-D TMA bulk copy shared -> global (local or peer), tracked by bulk async-groups
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

constexpr int kOutBufs = 3;

struct Cursor {
  int c, sub, nsub;
  Chunk ch;
};

template <class Op, typename T, typename WT, int NG>
__global__ void __launch_bounds__(kThreads)
mta_staged_kernel(const MtaArgs a, const Op op, const float* __restrict__ gscale,
                  float* __restrict__ partials) {
  constexpr OpFlags F = Op::F;
  constexpr int NS = Stages<NG>::value;
  constexpr int E = 16 / sizeof(T);            // elements per 16 bytes: TMA sizes are 16-byte multiples
  extern __shared__ __align__(128) unsigned char smem_raw[];
  T* const ring = reinterpret_cast<T*>(smem_raw);                                  // [NS][NG][kSub]
  T* const oring = ring + NS * NG * kSub;                                          // [kOutBufs][kSub]
  uint64_t* const full = reinterpret_cast<uint64_t*>(smem_raw + sizeof(T) * (NS * NG + kOutBufs) * kSub);
  __shared__ float sred[2][kThreads / 32];
  const int tid = threadIdx.x;
  const float gs = gscale != nullptr ? *gscale : 1.0f;
  const bool use_gs = gscale != nullptr;

  if (tid == 0) {
    for (int s = 0; s < NS; ++s) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  auto load_chunk = [&](Cursor& q) {
    if (q.c < a.nchunks) {
      q.ch = a.chunks[q.c];
      q.nsub = (q.ch.len + kSub - 1) / kSub;
    }
  };
  auto advance = [&](Cursor& q) {
    if (++q.sub >= q.nsub) {
      q.c += gridDim.x;
      q.sub = 0;
      load_chunk(q);
    }
  };
  auto issue = [&](const Cursor& q, int stage) {   // elected thread only
    const int t = q.ch.tensor;
    const int64_t e0 = q.ch.off + static_cast<int64_t>(q.sub) * kSub;
    const int n = min(kSub, q.ch.len - q.sub * kSub);
    const uint32_t bytes = static_cast<uint32_t>(n & ~(E - 1)) * sizeof(T);
    mbar_expect_tx(&full[stage], bytes * NG);
    if (bytes)
#pragma unroll
      for (int k = 0; k < NG; ++k)
        bulk_g2s(ring + (stage * NG + k) * kSub, reinterpret_cast<const T*>(sec_grad(a, k)[t]) + e0, bytes,
                 &full[stage]);
  };

  Cursor cons{static_cast<int>(blockIdx.x), 0, 0, {}};
  load_chunk(cons);
  Cursor prod = cons;
  if (tid == 0)
    for (int s = 0; s < NS && prod.c < a.nchunks; ++s) {
      issue(prod, s);
      advance(prod);
    }

  float red[2] = {0.f, 0.f};
  WT* wp = nullptr;
  float *s0p = nullptr, *s1p = nullptr;
  typename Op::PT pt{};
  for (int it = 0; cons.c < a.nchunks; ++it) {
    const int stage = it % NS;
    const uint32_t parity = (it / NS) & 1;
    const int t = cons.ch.tensor;
    if (cons.sub == 0) {
      wp = reinterpret_cast<WT*>(sec(a, kSecW)[t]);
      s0p = reinterpret_cast<float*>(sec(a, kSecS0)[t]);
      s1p = reinterpret_cast<float*>(sec(a, kSecS1)[t]);
      const float2 lrwd = *reinterpret_cast<const float2*>(&sec(a, kSecLrWd)[t]);
      const float2 aux = *reinterpret_cast<const float2*>(&sec(a, kSecAux)[t]);
      pt = op_prep(op, lrwd.x, lrwd.y, aux.x, aux.y, t);
    }
    const int64_t e0 = cons.ch.off + static_cast<int64_t>(cons.sub) * kSub;
    const int n = min(kSub, cons.ch.len - cons.sub * kSub);
    const int ntma = n & ~(E - 1);
    const bool act = 4 * tid < ntma;
    const int64_t e = e0 + 4 * tid;
    float w[4], g[4], s0[4], s1[4];
    if (act) {   // local operands first: their latency hides behind the wait below
      if (F.read_w) ld4(wp + e, w);
      if (F.read_s0) ld4(s0p + e, s0);
      if (F.read_s1) ld4(s1p + e, s1);
    }
    mbar_wait(&full[stage], parity);
    if (act) {
      float x[NG][4];
#pragma unroll
      for (int k = 0; k < NG; ++k) ld4(ring + (stage * NG + k) * kSub + 4 * tid, x[k]);
#pragma unroll
      for (int j = 0; j < 4; ++j) {   // CommCPU order: acc = x0; acc += x1 (+ x2 (+ x3 (+ x4))); ...
        float acc = x[0][j];
#pragma unroll
        for (int i = 1; i < NG; i += 4) {
          float sum = x[i][j];
#pragma unroll
          for (int k = 1; k < 4; ++k)
            if (i + k < NG) sum = add_as<T>(sum, x[i + k][j]);
          acc = add_as<T>(acc, sum);
        }
        g[j] = use_gs ? round_to<T>(__fmul_rn(acc, gs)) : acc;
        op.template elem<T>(pt, w[j], g[j], s0[j], s1[j], red);
      }
      if (F.write_s0) st4(s0p + e, s0);
      if (F.write_s1) st4(s1p + e, s1);
      if (F.write_w) st4(wp + e, w);
      if (F.write_out && a.nout > 0) st4(oring + (it % kOutBufs) * kSub + 4 * tid, w);
    }
    if (F.write_out && a.nout > 0) {
      fence_proxy_async();                 // my shared-memory writes -> visible to the TMA engine
      if (tid == 0) bulk_wait_read<kOutBufs - 2>();   // the buffer the NEXT iteration fills is free
    }
    // the < 16-byte tail of a tensor's last sub-chunk goes straight to global memory
    for (int i = ntma + tid; i < n; i += kThreads) {
      const int64_t ee = e0 + i;
      float ws = 0.f, gsc = grad1<T, 0>(a, t, ee), a0 = 0.f, a1 = 0.f;
      if (use_gs) gsc = round_to<T>(__fmul_rn(gsc, gs));
      if (F.read_w) ws = ld1(wp + ee);
      if (F.read_s0) a0 = s0p[ee];
      if (F.read_s1) a1 = s1p[ee];
      op.template elem<T>(pt, ws, gsc, a0, a1, red);
      if (F.write_s0) s0p[ee] = a0;
      if (F.write_s1) s1p[ee] = a1;
      if (F.write_w) st1(wp + ee, ws);
      if (F.write_out)
        for (int k = 0; k < a.nout; ++k) st1(reinterpret_cast<T*>(sec_out(a, k)[t]) + ee, ws);
    }
    __syncthreads();   // every thread has finished reading this stage / filling the out buffer
    if (tid == 0) {
      if (F.write_out && a.nout > 0) {
        const uint32_t bytes = static_cast<uint32_t>(ntma) * sizeof(T);
        if (bytes)
          for (int k = 0; k < a.nout; ++k)
            bulk_s2g(reinterpret_cast<T*>(sec_out(a, k)[t]) + e0, oring + (it % kOutBufs) * kSub, bytes);
        bulk_commit();
      }
      if (prod.c < a.nchunks) {
        issue(prod, stage);
        advance(prod);
      }
    }
    if (F.nred > 0 && cons.sub == cons.nsub - 1) {
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
        partials[static_cast<size_t>(cons.c) * F.nred + tid] = x;
      }
      __syncthreads();
    }
    advance(cons);
  }
  if (F.write_out && tid == 0) bulk_wait_all();   // every peer store has completed before the grid ends
}

template <class Op, typename T, typename WT, int NG>
int launch_staged_ng(const MtaCall& c, const Op& op, const float* gscale, float* partials) {
  constexpr int NS = Stages<NG>::value;
  const size_t smem = sizeof(T) * (NS * NG + kOutBufs) * kSub + sizeof(uint64_t) * NS;
  auto kernel = mta_staged_kernel<Op, T, WT, NG>;
  static int per_sm = 0;
  if (per_sm == 0) {
    CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kThreads, smem) != cudaSuccess || n < 1) n = 1;
    per_sm = n;
  }
  MtaArgs a;
  a.blob = c.blob;
  a.chunks = c.plan->chunks.as<Chunk>();
  a.nchunks = c.plan->nchunks;
  a.nt = c.nt;
  a.ngrad = c.ngrad;
  a.nout = c.nout;
  a.sync = nullptr;
  a.epoch = 0;
  a.sync_mode = 0;
  const int g = sm_count(c.device) * per_sm;
  const int grid = a.nchunks < g ? a.nchunks : g;
  kernel<<<grid, kThreads, smem, c.stream>>>(a, op, gscale, partials);
  CUDA_OK(cudaGetLastError());
  return 1;
}

template <class Op, typename T, typename WT>
int launch_staged_t(const MtaCall& c, const Op& op, const float* gscale, float* partials) {
  if (c.plan->nchunks == 0) return 0;
  if (c.ngrad == 2) return launch_staged_ng<Op, T, WT, 2>(c, op, gscale, partials);
  if (c.ngrad == 4) return launch_staged_ng<Op, T, WT, 4>(c, op, gscale, partials);
  if (c.ngrad == 8) return launch_staged_ng<Op, T, WT, 8>(c, op, gscale, partials);
  return -1;   // not a staged shape: caller falls back to the plain kernel
}

}  // namespace

// Staged forms; return -1 when the call is not eligible (copy count not 2/4/8).
int launch_sgd_staged(const MtaCall& c, bool has_mom, const OptScalars& s, const float* gscale) {
  if (has_mom) {
    SgdOp<true> op{s.momentum, s.rescale, s.clip};
    return c.dtype == kF32 ? launch_staged_t<SgdOp<true>, float, float>(c, op, gscale, nullptr)
                           : launch_staged_t<SgdOp<true>, __half, float>(c, op, gscale, nullptr);
  }
  SgdOp<false> op{0.f, s.rescale, s.clip};
  return c.dtype == kF32 ? launch_staged_t<SgdOp<false>, float, float>(c, op, gscale, nullptr)
                         : launch_staged_t<SgdOp<false>, __half, float>(c, op, gscale, nullptr);
}

int launch_adam_staged(const MtaCall& c, const OptScalars& s, const float* gscale) {
  AdamOp op{s.beta1, s.beta2, s.eps, s.rescale, s.clip};
  return c.dtype == kF32 ? launch_staged_t<AdamOp, float, float>(c, op, gscale, nullptr)
                         : launch_staged_t<AdamOp, __half, float>(c, op, gscale, nullptr);
}

// W section (same dtype) <- sum of the gradient copies: the reduce-scatter half on its own
int launch_copy_staged(const MtaCall& c) {
  CopyOp<true, false> op;
  return c.dtype == kF32 ? launch_staged_t<CopyOp<true, false>, float, float>(c, op, nullptr, nullptr)
                         : launch_staged_t<CopyOp<true, false>, __half, __half>(c, op, nullptr, nullptr);
}

}  // namespace b200kv
