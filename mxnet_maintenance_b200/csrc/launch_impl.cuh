// Shared launch plumbing for the multi-tensor kernels: persistent-grid sizing, MtaArgs, and the
// dispatch over the number of gradient copies.  Included by the optim_*.cu translation units, which
// each instantiate one family of kernels so that the build parallelises.
#pragma once
#include <algorithm>

#include "optim.cuh"

#include "launch.h"

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
  int g = sm_count(c.device) * (c.max_per_sm > 0 ? std::min(per_sm, c.max_per_sm) : per_sm);
  if (c.grid_div > 1) g = std::max(1, g / c.grid_div);
  return c.plan->nchunks < g ? (c.plan->nchunks > 0 ? c.plan->nchunks : 1) : g;
}

// every instantiation that can be launched registers itself when the library is loaded
template <class Op, typename T, typename WT, int NG>
struct KernelReg {
  static const bool done;
};
template <class Op, typename T, typename WT, int NG>
const bool KernelReg<Op, T, WT, NG>::done =
    register_kernel(reinterpret_cast<const void*>(&mta_kernel<Op, T, WT, NG>));

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
  // a rank whose shard of this key set is empty still takes part in the in-kernel barriers
  if (c.plan->nchunks == 0 && (c.sync == nullptr || c.sync_mode == 0)) return 0;
  KV_CHECK(c.ngrad >= 1 && c.ngrad <= kMaxCopies, "unsupported number of gradient copies " << c.ngrad);
  KV_CHECK(c.nout >= 0 && c.nout <= kMaxCopies, "unsupported number of output copies " << c.nout);
  const MtaArgs a = args_of(c);
#define GO(NG)                                                                                     \
  (void)KernelReg<Op, T, WT, NG>::done;                                                            \
  mta_kernel<Op, T, WT, NG><<<grid_for(c, mta_kernel<Op, T, WT, NG>), kThreads, 0, c.stream>>>(a, op, gscale, \
                                                                                              partials)
  if (!Op::F.read_g || c.ngrad == 1) { GO(1); }
  else if (c.ngrad == 2) { GO(2); }
  else if (c.ngrad == 4) { GO(4); }
  else if (c.ngrad == 8) { GO(8); }
  else { GO(0); }
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

}  // namespace b200kv
