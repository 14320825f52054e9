// Host launchers of the multi-tensor kernels (defined in optim.cu).
#pragma once
#include "common.h"

namespace b200kv {

struct MtaCall {
  const uint64_t* blob = nullptr;   // device table (see mta.h)
  const ChunkPlan* plan = nullptr;
  int nt = 0, ngrad = 1, nout = 0;
  int dtype = kF32;                 // element type of grads / outs: kF32 | kF16
  cudaStream_t stream = nullptr;
  int device = 0;
  // gradient copies are 16-byte aligned bucket slots on several GPUs: use the TMA-staged kernel
  bool staged = false;
  // in-kernel cross-GPU barriers (MtaArgs::sync*)
  const P2PFlags* sync = nullptr;
  uint32_t epoch = 0;
  int sync_mode = 0;
  // several ranks of one process share this GPU ("virtual ranks", used to exercise the multi-rank
  // path on one device): every kernel that may wait on a peer takes 1/grid_div of the persistent
  // grid so that all ranks' kernels are resident together
  int grid_div = 1;
  // cap on resident CTAs per SM for this launch (0 = what the occupancy API allows): the NVSwitch
  // multicast exchange runs faster with fewer requests in flight (tools/nvls_probe.cu)
  int max_per_sm = 0;
};

struct OptScalars {
  float momentum = 0.f, beta1 = 0.9f, beta2 = 0.999f, eps = 1e-8f;
  float rescale = 1.f, clip = -1.f, lower = -1.f, upper = -1.f;
  int bias_correction = 1;
};

// Each launcher returns the number of kernels it enqueued.
// Update kernels: W section is the fp32 weight (dtype kF32) or the fp32 master (dtype kF16).
int launch_sgd(const MtaCall& c, bool has_mom, const OptScalars& s, const float* gscale);
int launch_adam(const MtaCall& c, const OptScalars& s, const float* gscale);
// single-tensor forms: sgd_update / mp_sgd_update, lamb_update_phase1 / phase2 (fp32)
int launch_sgd_plain(const MtaCall& c, const OptScalars& s);
int launch_lamb_phase1(const MtaCall& c, const OptScalars& s);
int launch_lamb_phase2(const MtaCall& c, const float* r1, const float* r2, float lower, float upper);
int launch_adagrad(const MtaCall& c, const OptScalars& s, const float* gscale);
int launch_adamw(const MtaCall& c, const OptScalars& s, const float* rescale_dev);
// partials [nchunks]: number of non-finite values per chunk
int launch_nonfinite(const MtaCall& c, float* partials);
int launch_all_finite_final(const float* sums, int nt, int init_output, float* out, cudaStream_t stream);
// out copies <- 0
int launch_zero(const MtaCall& c);
int launch_multi_lars(int n, const float* lrs, const float* w2, const float* g2, const float* wds, float eta,
                      float eps, float rescale, float* out, cudaStream_t stream);
// overwrite the (lr, wd) section of an uploaded table from device arrays
int launch_fill_lrwd(const MtaCall& c, const float* lrs_dev, const float* wds_dev);
// partials: [nchunks][2] floats
// store_u: pass 1 also writes the update direction u to the X section, and pass 2 (launch_lamb2u)
// reads it back instead of recomputing it from m, v, w
int launch_lamb1(const MtaCall& c, const OptScalars& s, const float* gscale, float* partials, bool store_u = false);
int launch_lamb2(const MtaCall& c, const OptScalars& s, const float* ratio);
int launch_lamb2u(const MtaCall& c, const float* ratio);
// partials: [nchunks] floats; squares of the G section (ngrad copies summed first)
int launch_sumsq(const MtaCall& c, float* partials);
// W section <- sum of the G copies; optionally also to the out copies.  The W section holds
// `dtype` elements, or fp32 when w_is_f32 (fp16 value -> fp32 master).
int launch_copy(const MtaCall& c, bool to_w, bool to_out, bool w_is_f32 = false);
// out copies <- G section * (*gscale), rounded in dtype (clip_global_norm's `arr *= scale`)
int launch_scale(const MtaCall& c, const float* gscale);
// out copies (dtype) <- W section (fp32 master)
int launch_master_to_out(const MtaCall& c);

// push slice r of `src` (this rank's bucket, `slice_bytes` per slice) into dst[r] for every r != rank
int launch_p2p_push(const void* src, void* const* dst, int64_t slice_bytes, int world, int rank, int device,
                    cudaStream_t stream);
// stand-alone peer-memory barrier kernel (see optim.cuh): which = 0 "gradients ready",
// 1 "weights written"; used around the TMA-staged kernels, which carry no in-kernel barrier
int launch_p2p_barrier(const P2PFlags& f, int which, cudaStream_t stream);

// TMA-staged forms (staged.cu); -1 = not eligible, use the plain launcher
int launch_sgd_staged(const MtaCall& c, bool has_mom, const OptScalars& s, const float* gscale);
int launch_adam_staged(const MtaCall& c, const OptScalars& s, const float* gscale);
int launch_copy_staged(const MtaCall& c);

// deterministic all-reduce (rank order) of n <= 16384 floats through the peers' flag allocations;
// f.epoch is the exchange's own round counter
int launch_xsum(float* sums, int n, const P2PFlags& f, cudaStream_t stream);
int launch_reduce_partials(const float* partials, const ChunkPlan& plan, int nred, float* sums,
                           cudaStream_t stream);
int launch_lamb_ratio(const float* sums, int nt, float lower, float upper, float* ratio,
                      cudaStream_t stream);
int launch_global_norm(const float* sums, int nt, float max_norm, float* out2, cudaStream_t stream);

}  // namespace b200kv
