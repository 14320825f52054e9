// Dense part of the store.  Reference semantics: src/kvstore/kvstore_local.h:226-343 (init / push /
// pull / pushpull / broadcast, grouping, key checks); the reduce + broadcast it replaces:
// CommDevice::Reduce/Broadcast (src/kvstore/comm.h:503-616) and KVStoreNCCL::Reduce/Broadcast
// (src/kvstore/kvstore_nccl.h:213-421).
#include "store.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "nccl_dl.h"

namespace b200kv {

int launch_cast(const void* src, int src_dtype, void* dst, int dst_dtype, int64_t n,
                cudaStream_t stream);  // optim.cu

namespace {

int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }
constexpr int kXsumCapHost = 16384;   // optim.cuh kXsumCap

struct Blob {
  uint64_t* p;
  int nt, ngrad, nout;
  Blob(uint64_t* base, int nt_, int ng, int no) : p(base), nt(nt_), ngrad(ng), nout(no) {
    std::memset(p, 0, words() * 8);
  }
  size_t words() const { return mta_blob_words(nt, ngrad, nout); }
  void G(int k, int t, const void* ptr) { p[static_cast<size_t>(k) * nt + t] = reinterpret_cast<uint64_t>(ptr); }
  void O(int k, int t, const void* ptr) { p[static_cast<size_t>(ngrad + k) * nt + t] = reinterpret_cast<uint64_t>(ptr); }
  void S(int sec, int t, const void* ptr) {
    p[static_cast<size_t>(ngrad + nout + sec) * nt + t] = reinterpret_cast<uint64_t>(ptr);
  }
  void F2(int sec, int t, float a, float b) {
    float v[2] = {a, b};
    std::memcpy(&p[static_cast<size_t>(ngrad + nout + sec) * nt + t], v, 8);
  }
};

inline char* at(const DevBuf& b, int64_t elem, size_t esize) {
  // may point outside the allocation for shard-relative arrays; only in-range chunks dereference it
  return static_cast<char*>(b.p) + elem * static_cast<int64_t>(esize);
}

}  // namespace

Store::Store(int device, int rank, int world, bool in_process, int grid_share)
    : device_(device), rank_(rank), world_(world), in_process_(in_process), grid_share_(grid_share) {
  KV_CHECK(world >= 1 && rank >= 0 && rank < world, "invalid rank " << rank << " / world " << world);
  int n = 0;
  CUDA_OK(cudaGetDeviceCount(&n));
  KV_CHECK(device >= 0 && device < n, "invalid CUDA device " << device << " (" << n << " visible)");
  std::memset(&opt_, 0, sizeof(opt_));
  opt_.kind = B200KV_OPT_NONE;
  opt_.clip_global_norm = -1.f;
  use_device();
  gnorm_.reserve(2 * sizeof(float));
}

Store::~Store() {
  int prev = 0;
  cudaGetDevice(&prev);
  struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};   // the caller's current device stays
  cudaSetDevice(device_);
  cudaStreamSynchronize(stream_);
  for (auto& sp : segs_) {   // unmap the peers' buffers before our own are freed
    Segment::P2P& P = sp->p2p;
    if (!P.ok || P.symmetric || in_process_) continue;   // symmetric / sibling memory is not ours to unmap
    for (int k = 0; k < world_; ++k) {
      if (k == rank_) continue;
      if (P.g[k]) cudaIpcCloseMemHandle(P.g[k]);
      if (P.w[k]) cudaIpcCloseMemHandle(P.w[k]);
      if (P.flags[k]) cudaIpcCloseMemHandle(P.flags[k]);
      if (P.stage[k]) cudaIpcCloseMemHandle(P.stage[k]);
    }
  }
  for (auto& e : prof_events_) {
    cudaEventDestroy(e.first);
    cudaEventDestroy(e.second);
  }
  if (comm_) NcclApi::get().CommDestroy(reinterpret_cast<ncclComm_t>(comm_));
}

void Store::use_device() { CUDA_OK(cudaSetDevice(device_)); }

void Store::wait() {
  use_device();
  CUDA_OK(cudaStreamSynchronize(stream_));
  check_rsp_errors();
  // a peer that never arrived at a cross-GPU barrier leaves an error flag instead of a hung GPU
  for (auto& sp : segs_) {
    if (!sp->p2p.ok) continue;
    uint32_t err = 0;
    CUDA_OK(cudaMemcpy(&err, sp->p2p.local_flags.as<uint32_t>() + 16, sizeof(err), cudaMemcpyDeviceToHost));
    KV_CHECK(err == 0, "a peer GPU did not reach the gradient-exchange barrier within the time limit; "
                       "the weights of the last step are not valid");
  }
}

void Store::comm_init(const char id[128]) {
  use_device();
  KV_CHECK(comm_ == nullptr, "communicator already initialised");
  if (world_ == 1) return;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
  ncclUniqueId uid;
  std::memcpy(&uid, id, sizeof(uid));
  ncclComm_t c;
  NCCL_OK(NcclApi::get().CommInitRank(&c, world_, uid, rank_));
  comm_ = reinterpret_cast<ncclComm*>(c);
}

void Store::barrier() {
  use_device();
  if (world_ > 1 && !in_process_) {
    KV_CHECK(comm_, "b200kv_comm_init has not been called");
    DevBuf tmp;
    tmp.reserve(16);
    CUDA_OK(cudaMemsetAsync(tmp.p, 0, 16, stream_));
    NCCL_OK(NcclApi::get().AllReduce(tmp.p, tmp.p, 1, ncclFloat32, ncclSum,
                                     reinterpret_cast<ncclComm_t>(comm_), stream_));
    CUDA_OK(cudaStreamSynchronize(stream_));
  }
  CUDA_OK(cudaStreamSynchronize(stream_));
}

// ---- keys ----------------------------------------------------------------------------------------
void Store::check_key_type(bool is_str) {
  // kvstore_local.h:322-325: a store uses either int or string keys, never both
  const int t = is_str ? 1 : 0;
  if (key_type_ < 0) key_type_ = t;
  KV_CHECK(key_type_ == t, "inconsistent key types: mixing int and string keys is not allowed");
}

void Store::lookup_keys(int num, const char* const* names, bool create, int* out) {
  check_key_type(true);
  for (int i = 0; i < num; ++i) {
    const std::string s(names[i]);
    auto it = str_keys_.find(s);
    if (it == str_keys_.end()) {
      KV_CHECK(create, "key " << s << " has not been initialized");
      const int k = static_cast<int>(str_keys_.size());
      it = str_keys_.emplace(s, k).first;
    }
    out[i] = it->second;
  }
}

KeyInfo& Store::info(int key) {
  auto it = keys_.find(key);
  KV_CHECK(it != keys_.end(), "key " << key << " has not been initialized");
  return it->second;
}

void Store::init(int num, const int* keys, DLManagedTensor* const* vals) {
  use_device();
  if (key_type_ < 0) key_type_ = 0;
  for (int i = 0; i < num; ++i) {
    KV_CHECK(keys_.find(keys[i]) == keys_.end() && rsp_.find(keys[i]) == rsp_.end(),
             "duplicate init of key " << keys[i]);   // kvstore_local.h:229-232
    const TensorView v = view_of(vals[i], device_, "init value");
    KV_CHECK(v.dtype == kF32 || v.dtype == kF16, "init: only float32/float16 values are supported");
    KeyInfo k;
    k.key = keys[i];
    k.dtype = v.dtype;
    k.size = v.numel;
    k.staged.reset(new DevBuf());
    const size_t bytes = static_cast<size_t>(v.numel) * dtype_size(v.dtype);
    k.staged->reserve(bytes ? bytes : 16);
    if (bytes)
      CUDA_OK(cudaMemcpyAsync(k.staged->p, v.data, bytes, cudaMemcpyDeviceToDevice, stream_));
    if (world_ > 1 && bytes && !in_process_) {  // every rank starts from rank 0's value
      KV_CHECK(comm_, "b200kv_comm_init has not been called");
      NCCL_OK(NcclApi::get().Broadcast(k.staged->p, k.staged->p, v.numel, nccl_type(v.dtype), 0,
                                       reinterpret_cast<ncclComm_t>(comm_), stream_));
    }
    keys_.emplace(keys[i], std::move(k));
    unsealed_.push_back(keys[i]);
  }
}

// Move every staged key into a new flat segment per dtype.
void Store::seal() {
  if (unsealed_.empty()) return;
  for (int dt : {static_cast<int>(kF32), static_cast<int>(kF16)}) {
    std::vector<int> ks;
    for (int k : unsealed_)
      if (keys_[k].dtype == dt) ks.push_back(k);
    if (ks.empty()) continue;
    std::unique_ptr<Segment> s(new Segment());
    s->dtype = dt;
    int64_t off = 0;
    for (int k : ks) {
      KeyInfo& ki = keys_[k];
      ki.off = off;
      ki.seg = static_cast<int>(segs_.size());
      off = align_up(off + ki.size, kAlignElems);
    }
    s->total = align_up(std::max<int64_t>(off, kAlignElems), static_cast<int64_t>(world_) * kAlignElems);
    s->slice = s->total / world_;
    s->keys = ks;
    const size_t es = dtype_size(dt);
    {
      void* peers[kMaxCopies] = {nullptr};
      void* mc = nullptr;
      if (alloc_exchange(s->wbuf, static_cast<size_t>(s->total) * es, peers, &mc)) {
        s->p2p.symmetric = true;
        for (int k = 0; k < world_; ++k) s->p2p.w[k] = peers[k];
        s->p2p.mc_w = mc;
      }
    }
    CUDA_OK(cudaMemsetAsync(s->wbuf.p, 0, static_cast<size_t>(s->total) * es, stream_));
    for (int k : ks) {
      KeyInfo& ki = keys_[k];
      if (ki.size)
        CUDA_OK(cudaMemcpyAsync(at(s->wbuf, ki.off, es), ki.staged->p, static_cast<size_t>(ki.size) * es,
                                cudaMemcpyDeviceToDevice, stream_));
    }
    // staged buffers are freed after the copies have run
    CUDA_OK(cudaStreamSynchronize(stream_));
    for (int k : ks) keys_[k].staged.reset();
    segs_.push_back(std::move(s));
  }
  unsealed_.clear();
}

// ---- optimizer -----------------------------------------------------------------------------------
void Store::set_optimizer(const b200kv_optimizer& o) {
  use_device();
  KV_CHECK(o.kind >= B200KV_OPT_NONE && o.kind <= B200KV_OPT_ADAGRAD, "unknown optimizer kind " << o.kind);
  const bool had_mom = opt_.kind == B200KV_OPT_SGD && opt_.momentum != 0.f;
  const bool has_mom = o.kind == B200KV_OPT_SGD && o.momentum != 0.f;
  const bool reset = o.kind != opt_.kind || o.multi_precision != opt_.multi_precision || had_mom != has_mom;
  if (reset) {
    for (auto& sp : segs_) {
      Segment& s = *sp;
      if (s.w_stale) {  // the master is authoritative: materialise the stored fp16 value first
        launches_ += launch_cast(s.master.p, kF32, s.wbuf.p, kF16, s.total, stream_);
        s.w_stale = false;
      }
      CUDA_OK(cudaStreamSynchronize(stream_));
      s.master.release(); s.s0.release(); s.s1.release(); s.xbuf.release();
      s.has_master = s.has_s0 = s.has_s1 = false;
      s.state_len = 0;
    }
  }
  opt_ = o;
  sc_.momentum = o.momentum; sc_.beta1 = o.beta1; sc_.beta2 = o.beta2; sc_.eps = o.epsilon;
  sc_.rescale = o.rescale_grad; sc_.clip = o.clip_gradient; sc_.lower = o.lower_bound;
  sc_.upper = o.upper_bound; sc_.bias_correction = o.bias_correction;
}

void Store::set_hyper(int num, const int* keys, const float* lrs, const float* wds, const int* steps) {
  for (int i = 0; i < num; ++i) {
    auto it = keys_.find(keys[i]);
    if (it != keys_.end()) {
      if (lrs) it->second.lr = lrs[i];
      if (wds) it->second.wd = wds[i];
      if (steps) it->second.step = steps[i];
      continue;
    }
    KV_CHECK(rsp_.find(keys[i]) != rsp_.end(), "key " << keys[i] << " has not been initialized");
    set_hyper_rsp(keys[i], lrs ? lrs[i] : NAN, wds ? wds[i] : NAN, steps ? steps[i] : -1);
  }
}

// States are created at the first update from the stored weight, as Updater.__call__ does
// (python/mxnet/optimizer/optimizer.py:2096-2098); they cover only this rank's shard when world > 1.
void Store::ensure_states(Segment& s) {
  const bool want_s0 = (opt_.kind == B200KV_OPT_SGD && opt_.momentum != 0.f) ||
                       opt_.kind == B200KV_OPT_ADAM || opt_.kind == B200KV_OPT_LAMB ||
                       opt_.kind == B200KV_OPT_ADAGRAD;
  const bool want_s1 = opt_.kind == B200KV_OPT_ADAM || opt_.kind == B200KV_OPT_LAMB;
  const bool want_master = use_mp(s);
  if (s.state_len == 0) {
    s.state_begin = world_ == 1 ? 0 : static_cast<int64_t>(rank_) * s.slice;
    s.state_len = world_ == 1 ? s.total : s.slice;
  }
  const size_t bytes = static_cast<size_t>(s.state_len) * 4;
  if (want_s0 && !s.has_s0) {
    s.s0.reserve(bytes);
    CUDA_OK(cudaMemsetAsync(s.s0.p, 0, bytes, stream_));
    s.has_s0 = true;
  }
  if (want_s1 && !s.has_s1) {
    s.s1.reserve(bytes);
    CUDA_OK(cudaMemsetAsync(s.s1.p, 0, bytes, stream_));
    s.has_s1 = true;
  }
  if (opt_.kind == B200KV_OPT_LAMB && s.xbuf.p == nullptr) {
    static const bool keep_u = [] { const char* e = getenv("B200KV_LAMB_RECOMPUTE"); return !(e && e[0] == '1'); }();
    if (keep_u) s.xbuf.reserve(bytes);
  }
  if (want_master && !s.has_master) {
    s.master.reserve(bytes);
    launches_ += launch_cast(at(s.wbuf, s.state_begin, 2), kF16, s.master.p, kF32, s.state_len, stream_);
    s.has_master = true;
  }
}

// ---- plans ---------------------------------------------------------------------------------------
DensePlan& Store::plan_for(int seg, const std::vector<int>& keys, int ngrad, int nout) {
  std::string id = std::to_string(seg) + ":" + std::to_string(ngrad) + ":" + std::to_string(nout) + ":";
  id.reserve(id.size() + keys.size() * 6);
  for (int k : keys) { id += std::to_string(k); id += ','; }
  auto it = plans_.find(id);
  if (it != plans_.end()) return *it->second;

  std::unique_ptr<DensePlan> p(new DensePlan());
  Segment& s = *segs_[seg];
  p->keys = keys;
  p->seg = seg;
  p->ngrad = ngrad;
  p->nout = nout;
  p->full_segment = keys.size() == s.keys.size();
  const int nt = static_cast<int>(keys.size());
  std::vector<int64_t> b(nt, 0), e(nt), sb(nt), se(nt);
  const int64_t lo = static_cast<int64_t>(rank_) * s.slice, hi = lo + s.slice;
  for (int t = 0; t < nt; ++t) {
    const KeyInfo& ki = keys_[keys[t]];
    e[t] = ki.size;
    // shard range relative to the tensor's first element
    sb[t] = std::min(std::max(lo - ki.off, int64_t(0)), ki.size);
    se[t] = std::max(std::min(hi - ki.off, ki.size), sb[t]);
    // whole 4-element vectors: a tensor's last vector may reach into the zero padding of its
    // 64-element slot (multicast accesses have no scalar fp16 form; zeros stay zeros under every
    // optimizer here)
    if (se[t] == ki.size && se[t] > sb[t]) se[t] = (se[t] + 3) / 4 * 4;
    if (world_ > 1)
      for (int r = 0; r < world_; ++r) {
        const int64_t plo = std::max(ki.off, r * s.slice), phi = std::min(ki.off + ki.size, (r + 1) * s.slice);
        if (plo < phi) p->pieces.push_back({plo, phi - plo, r});
      }
  }
  p->full.build(b, e);
  if (world_ > 1) {
    // a small shard gets finer chunks so that the persistent CTAs stay balanced (a ResNet-50 shard
    // at 8 GPUs is only ~780 chunks of 4096 elements for ~450 resident CTAs)
    int64_t shard_elems = 0;
    for (int t = 0; t < nt; ++t) shard_elems += se[t] - sb[t];
    const int chunk = shard_elems / kChunk < 16 * static_cast<int64_t>(sm_count(device_)) ? 1024 : kChunk;
    p->shard.build(sb, se, chunk);
  }
  const int maxc = std::max(p->full.nchunks, p->shard.nchunks);
  p->partials.reserve(sizeof(float) * 2 * static_cast<size_t>(std::max(maxc, 1)));
  p->sums.reserve(sizeof(float) * (3 * static_cast<size_t>(nt) + 4));
  DensePlan& ref = *p;
  plans_.emplace(id, std::move(p));
  return ref;
}

// ---- the update sequence on a chunk list -----------------------------------------------------------
const float* Store::global_norm_scale(DensePlan& p, const ChunkPlan& cp, const MtaCall& gcall) {
  // clip_global_norm (python/mxnet/gluon/utils.py:122-161) on the merged gradients of this push
  MtaCall c = gcall;
  c.plan = &cp;
  c.sync = nullptr;      // reads the locally reduced bucket only: no cross-GPU barrier here
  c.sync_mode = 0;
  c.staged = false;
  float* partials = p.partials.as<float>();
  float* sums = p.sums.as<float>();
  launches_ += launch_sumsq(c, partials);
  launches_ += launch_reduce_partials(partials, cp, 1, sums, stream_);
  if (world_ > 1) allreduce_sums(*segs_[p.seg], sums, c.nt);
  launches_ += launch_global_norm(sums, c.nt, opt_.clip_global_norm, gnorm_.as<float>(), stream_);
  gnorm_valid_ = true;
  return gnorm_.as<float>();
}

void Store::update_chunks(DensePlan& p, const ChunkPlan& cp, const MtaCall& call, const float* gscale) {
  MtaCall c = call;
  c.plan = &cp;
  switch (opt_.kind) {
    case B200KV_OPT_SGD:
      launches_ += launch_sgd(c, opt_.momentum != 0.f, sc_, gscale);
      break;
    case B200KV_OPT_ADAM:
      launches_ += launch_adam(c, sc_, gscale);
      break;
    case B200KV_OPT_ADAGRAD:
      launches_ += launch_adagrad(c, sc_, gscale);
      break;
    case B200KV_OPT_LAMB: {
      float* partials = p.partials.as<float>();
      float* sums = p.sums.as<float>();
      float* ratio = sums + 2 * static_cast<size_t>(c.nt);
      MtaCall c1 = c, c2 = c;
      c1.sync_mode = c.sync_mode & (kSyncWaitReady | kSyncMcGrad);   // pass 1: first reader of the gradients
      c2.sync_mode = c.sync_mode & (kSyncSignalDone | kSyncMcOut);   // pass 2: last writer of the weights
      const bool keep_u = segs_[p.seg]->xbuf.p != nullptr;
      launches_ += launch_lamb1(c1, sc_, gscale, partials, keep_u);
      launches_ += launch_reduce_partials(partials, cp, 2, sums, stream_);
      if (world_ > 1)  // per-tensor norms span shards: one tiny all-reduce between the two passes
        allreduce_sums(*segs_[p.seg], sums, 2 * c.nt);
      launches_ += launch_lamb_ratio(sums, c.nt, sc_.lower, sc_.upper, ratio, stream_);
      launches_ += keep_u ? launch_lamb2u(c2, ratio) : launch_lamb2(c2, sc_, ratio);
      break;
    }
    default:
      KV_CHECK(false, "no optimizer set");
  }
}

void Store::allreduce_sums(Segment& s, float* sums, int n) {
  static const bool use_nccl = [] { const char* e = getenv("B200KV_XSUM"); return e && e[0] == '0'; }();
  if (s.p2p.ok && (in_process_ || !use_nccl) && n <= kXsumCapHost) {
    P2PFlags f;
    for (int k = 0; k < kMaxCopies; ++k) f.flags[k] = s.p2p.flags[k];
    f.world = world_; f.rank = rank_; f.epoch = ++s.p2p.xepoch;
    launches_ += launch_xsum(sums, n, f, stream_);
    return;
  }
  KV_CHECK(comm_ != nullptr, "no communicator for the partial-sum all-reduce");
  NCCL_OK(NcclApi::get().AllReduce(sums, sums, static_cast<size_t>(n), ncclFloat32, ncclSum,
                                   reinterpret_cast<ncclComm_t>(comm_), stream_));
}

float Store::last_global_norm() {
  use_device();
  KV_CHECK(gnorm_valid_, "no push with clip_global_norm has run yet");
  float h[2];
  CUDA_OK(cudaMemcpyAsync(h, gnorm_.p, sizeof(h), cudaMemcpyDeviceToHost, stream_));
  CUDA_OK(cudaStreamSynchronize(stream_));
  return h[1];
}

// ---- one class of keys (same segment, same number of value / out copies) ---------------------------
void Store::run_class(Mode mode, DensePlan& p, const std::vector<const Entry*>& ents) {
  Segment& s = *segs_[p.seg];
  const int nt = static_cast<int>(p.keys.size());
  const size_t es = dtype_size(s.dtype);
  const bool push = (mode & kPush) != 0, pull = (mode & kPull) != 0;
  const bool mp = has_opt() && use_mp(s);
  if (push && has_opt()) {
    KV_CHECK(!(s.dtype == kF16 && !opt_.multi_precision),
             "float16 weights need multi_precision=True (pure fp16 optimizer state is not supported)");
    ensure_states(s);
  }
  NcclApi* nccl = world_ > 1 && !in_process_ ? &NcclApi::get() : nullptr;
  ncclComm_t comm = reinterpret_cast<ncclComm_t>(comm_);
  if (world_ > 1 && !in_process_) KV_CHECK(comm_, "b200kv_comm_init has not been called");

  // carve all tables of this call out of one ring slot
  const int ng = push ? p.ngrad : 1, no = pull ? p.nout : 0;
  // sharded update (world > 1): one gradient source (the reduced bucket); the fp16 weight shard
  // for the all-gather is its single "out" copy in multi-precision mode
  const int upd_ng = world_ > 1 ? 1 : ng, upd_no = world_ > 1 ? (mp ? 1 : 0) : no;
  const size_t w_pack = mta_blob_words(nt, ng, 0), w_upd = mta_blob_words(nt, upd_ng, upd_no),
               w_unpack = mta_blob_words(nt, 1, no);
  uint64_t* host = ring_.begin((w_pack + w_upd + w_unpack) * 8);
  const uint64_t* dev = nullptr;  // set after commit; offsets are fixed now
  const size_t o_pack = 0, o_upd = w_pack, o_unpack = w_pack + w_upd;

  MtaCall base;
  base.nt = nt;
  base.dtype = s.dtype;
  base.stream = stream_;
  base.device = device_;
  base.grid_div = grid_share_;

  auto fill_update_scalars = [&](Blob& b) {
    for (int t = 0; t < nt; ++t) {
      const KeyInfo& ki = keys_[p.keys[t]];
      b.S(kSecW, t, mp ? at(s.master, ki.off - s.state_begin, 4) : at(s.wbuf, ki.off, 4));
      if (s.has_s0) b.S(kSecS0, t, at(s.s0, ki.off - s.state_begin, 4));
      if (s.has_s1) b.S(kSecS1, t, at(s.s1, ki.off - s.state_begin, 4));
      if (s.xbuf.p) b.S(kSecX, t, at(s.xbuf, ki.off - s.state_begin, 4));
      b.F2(kSecLrWd, t, ki.lr, ki.wd);
      if (opt_.kind == B200KV_OPT_LAMB) {
        // 1 - beta^t with the host's powf in fp32, as MultiLAMBKernelStep1 does per element
        // (multi_lamb.cc:59-67)
        const float tf = static_cast<float>(ki.step);
        b.F2(kSecAux, t, 1.0f - powf(opt_.beta1, tf), 1.0f - powf(opt_.beta2, tf));
      }
    }
  };

  // Arrays that ARE the store's own buffers (b200kv_view) need no copy: a gradient that already
  // lives in its bucket slot is not packed, an out that is the stored weight is not written twice.
  bool g_alias = push && ng == 1, o_alias = pull && no == 1;
  for (int t = 0; t < nt && (g_alias || o_alias); ++t) {
    const int64_t off = keys_[p.keys[t]].off;
    if (g_alias && (s.gbuf.p == nullptr || ents[t]->vals[0].data != at(s.gbuf, off, es))) g_alias = false;
    if (o_alias && ents[t]->outs[0].data != at(s.wbuf, off, es)) o_alias = false;
  }
  if (world_ == 1) {
    if (push && has_opt() && o_alias) {
      // the caller's weights are the stored weights: update in place, no separate out copy
      Blob b(host + o_upd, nt, ng, mp ? 1 : 0);
      for (int t = 0; t < nt; ++t) {
        for (int k = 0; k < ng; ++k) b.G(k, t, ents[t]->vals[k].data);
        if (mp) b.O(0, t, ents[t]->outs[0].data);
      }
      fill_update_scalars(b);
      dev = ring_.commit(stream_);
      MtaCall c = base;
      c.blob = dev + o_upd; c.ngrad = ng; c.nout = mp ? 1 : 0; c.plan = &p.full;
      prof_begin();
      const float* gscale = opt_.clip_global_norm >= 0.f ? global_norm_scale(p, p.full, c) : nullptr;
      update_chunks(p, p.full, c, gscale);
      prof_end();
      // wbuf was written as the (fp16) out, or is the fp32 weight itself — for THESE keys only: an
      // earlier non-aliased multi-precision push may have left other keys of the segment stale
      if (p.full_segment) s.w_stale = false;
      ring_.end(stream_);
      return;
    }
    if (push) {
      Blob b(host + o_upd, nt, ng, no);
      for (int t = 0; t < nt; ++t) {
        for (int k = 0; k < ng; ++k) b.G(k, t, ents[t]->vals[k].data);
        for (int k = 0; k < no; ++k) b.O(k, t, ents[t]->outs[k].data);
      }
      if (has_opt()) fill_update_scalars(b);
      else for (int t = 0; t < nt; ++t) b.S(kSecW, t, at(s.wbuf, keys_[p.keys[t]].off, es));
      dev = ring_.commit(stream_);
      MtaCall c = base;
      c.blob = dev + o_upd; c.ngrad = ng; c.nout = no; c.plan = &p.full;
      if (has_opt()) {
        prof_begin();
        const float* gscale = opt_.clip_global_norm >= 0.f ? global_norm_scale(p, p.full, c) : nullptr;
        update_chunks(p, p.full, c, gscale);
        prof_end();
        if (mp) s.w_stale = true;
      } else {
        // no updater: the merged value replaces the stored one (kvstore_local.h:268-274)
        launches_ += launch_copy(c, true, no > 0);
      }
    } else {  // pull only
      Blob b(host + o_unpack, nt, 1, no);
      for (int t = 0; t < nt; ++t) {
        const KeyInfo& ki = keys_[p.keys[t]];
        b.G(0, t, at(s.wbuf, ki.off, es));
        b.S(kSecW, t, s.w_stale ? at(s.master, ki.off - s.state_begin, 4) : nullptr);
        for (int k = 0; k < no; ++k) b.O(k, t, ents[t]->outs[k].data);
      }
      dev = ring_.commit(stream_);
      MtaCall c = base;
      c.blob = dev + o_unpack; c.ngrad = 1; c.nout = no; c.plan = &p.full;
      if (s.w_stale) launches_ += launch_master_to_out(c);
      else launches_ += launch_copy(c, false, true);
    }
    ring_.end(stream_);
    return;
  }

  // ---------------- world > 1: pack -> reduce-scatter -> sharded update -> all-gather -> unpack
  if (push) last_exchange_ = "nccl";
  if (s.gbuf.p == nullptr && push) alloc_gbuf(s);
  if (push && in_process_) {
    // one process, several GPUs: every push — whole model or one key, with or without an optimizer —
    // runs as the fused peer-memory kernel (the Group linked the members' buffers beforehand)
    KV_CHECK(s.p2p.ok, "internal: segment not linked to its peers");
    ring_.end(stream_);
    run_class_p2p(p, s, ents, pull, g_alias, o_alias);
    return;
  }
  if (push && has_opt() && p.full_segment) {
    if (!s.p2p.tried) setup_p2p(s);
    if (s.p2p.ok) {
      ring_.end(stream_);   // hand the (unused) slot back; the peer path carves its own
      run_class_p2p(p, s, ents, pull, g_alias, o_alias);
      return;
    }
  }
  if (push) {
    Blob b(host + o_pack, nt, ng, 0);
    for (int t = 0; t < nt; ++t) {
      for (int k = 0; k < ng; ++k) b.G(k, t, ents[t]->vals[k].data);
      b.S(kSecW, t, at(s.gbuf, keys_[p.keys[t]].off, es));
    }
    if (has_opt()) {
      Blob u(host + o_upd, nt, upd_ng, upd_no);
      for (int t = 0; t < nt; ++t) {
        const KeyInfo& ki = keys_[p.keys[t]];
        u.G(0, t, at(s.gbuf, ki.off, es));
        if (mp) u.O(0, t, at(s.wbuf, ki.off, es));
      }
      fill_update_scalars(u);
    }
  }
  if (pull) {
    Blob b(host + o_unpack, nt, 1, no);
    for (int t = 0; t < nt; ++t) {
      b.G(0, t, at(s.wbuf, keys_[p.keys[t]].off, es));
      for (int k = 0; k < no; ++k) b.O(k, t, ents[t]->outs[k].data);
    }
  }
  dev = ring_.commit(stream_);
  const ncclDataType_t nt_type = nccl_type(s.dtype);
  if (push) {
    MtaCall c = base;
    c.blob = dev + o_pack; c.ngrad = ng; c.nout = 0; c.plan = &p.full;
    if (!g_alias) {
      prof_begin(1);
      launches_ += launch_copy(c, true, false);   // user gradients (local copies summed) -> fusion bucket
      prof_end();
    }
    if (has_opt()) {
      prof_begin(2);
      if (p.full_segment) {
        NCCL_OK(nccl->ReduceScatter(s.gbuf.p, at(s.gbuf, rank_ * s.slice, es), s.slice, nt_type, ncclSum,
                                    comm, stream_));
      } else {
        NCCL_OK(nccl->GroupStart());
        for (const auto& pc : p.pieces)
          NCCL_OK(nccl->Reduce(at(s.gbuf, pc.off, es), at(s.gbuf, pc.off, es), pc.len, nt_type, ncclSum,
                               pc.root, comm, stream_));
        NCCL_OK(nccl->GroupEnd());
      }
      prof_end();
      MtaCall u = base;
      u.blob = dev + o_upd; u.ngrad = upd_ng; u.nout = upd_no; u.plan = &p.shard;
      prof_begin();
      const float* gscale = opt_.clip_global_norm >= 0.f ? global_norm_scale(p, p.shard, u) : nullptr;
      update_chunks(p, p.shard, u, gscale);
      prof_end();
      prof_begin(3);
      if (p.full_segment) {
        NCCL_OK(nccl->AllGather(at(s.wbuf, rank_ * s.slice, es), s.wbuf.p, s.slice, nt_type, comm, stream_));
      } else {
        NCCL_OK(nccl->GroupStart());
        for (const auto& pc : p.pieces)
          NCCL_OK(nccl->Broadcast(at(s.wbuf, pc.off, es), at(s.wbuf, pc.off, es), pc.len, nt_type, pc.root,
                                  comm, stream_));
        NCCL_OK(nccl->GroupEnd());
      }
      prof_end();
    } else {
      // no updater: all-reduce the merged gradients straight into the stored values
      if (p.full_segment) {
        NCCL_OK(nccl->AllReduce(s.gbuf.p, s.wbuf.p, s.total, nt_type, ncclSum, comm, stream_));
      } else {
        NCCL_OK(nccl->GroupStart());
        for (int t = 0; t < nt; ++t) {
          const KeyInfo& ki = keys_[p.keys[t]];
          NCCL_OK(nccl->AllReduce(at(s.gbuf, ki.off, es), at(s.wbuf, ki.off, es), ki.size, nt_type, ncclSum,
                                  comm, stream_));
        }
        NCCL_OK(nccl->GroupEnd());
      }
    }
  }
  if (pull && !o_alias) {
    MtaCall c = base;
    c.blob = dev + o_unpack; c.ngrad = 1; c.nout = no; c.plan = &p.full;
    prof_begin(4);
    launches_ += launch_copy(c, false, true);
    prof_end();
  }
  ring_.end(stream_);
}

// ---- peer-memory path --------------------------------------------------------------------------------
// Map every rank's gradient bucket, weight buffer and flag block into this process (CUDA IPC over
// NVLink peer access).  All ranks must agree on the outcome, so the result is min-reduced.
bool Store::alloc_exchange(DevBuf& buf, size_t bytes, void** peers, void** mc) {
  if (symm_alloc_ != nullptr && world_ > 1 && world_ <= kMaxCopies) {
    void* local = nullptr;
    void* pp[kMaxCopies] = {nullptr};
    void* m = nullptr;
    if (symm_alloc_(symm_ctx_, bytes, &local, pp, &m) == 0 && local != nullptr) {
      buf.adopt(local, bytes);
      for (int k = 0; k < world_; ++k) peers[k] = pp[k];
      *mc = m;
      return true;
    }
  }
  buf.reserve(bytes);
  return false;
}

void Store::alloc_gbuf(Segment& s) {
  const size_t bytes = static_cast<size_t>(s.total) * dtype_size(s.dtype);
  void* peers[kMaxCopies] = {nullptr};
  void* mc = nullptr;
  if (alloc_exchange(s.gbuf, bytes, peers, &mc) && s.p2p.symmetric) {
    for (int k = 0; k < world_; ++k) s.p2p.g[k] = peers[k];
    s.p2p.mc_g = mc;
  } else {
    s.p2p.symmetric = false;   // mixed provenance: use the IPC path
  }
  CUDA_OK(cudaMemsetAsync(s.gbuf.p, 0, bytes, stream_));
}

void Store::setup_p2p(Segment& s) {
  s.p2p.tried = true;
  if (s.p2p.symmetric) {
    // symmetric memory is already mapped everywhere: only the flag block is missing
    Segment::P2P& P = s.p2p;
    void* peers[kMaxCopies] = {nullptr};
    void* mc = nullptr;
    if (alloc_exchange(P.local_flags, 2u << 20, peers, &mc)) {
      CUDA_OK(cudaMemsetAsync(P.local_flags.p, 0, 256, stream_));
      for (int k = 0; k < world_; ++k) P.flags[k] = static_cast<uint32_t*>(peers[k]);
      // every rank's flags must be zero before anyone signals: a tiny all-reduce is the barrier
      DevBuf tmp;
      tmp.reserve(16);
      CUDA_OK(cudaMemsetAsync(tmp.p, 0, 16, stream_));
      NCCL_OK(NcclApi::get().AllReduce(tmp.p, tmp.p, 1, ncclFloat32, ncclSum,
                                       reinterpret_cast<ncclComm_t>(comm_), stream_));
      CUDA_OK(cudaStreamSynchronize(stream_));
      P2PFlags h;
      std::memset(&h, 0, sizeof(h));
      for (int k = 0; k < world_; ++k) h.flags[k] = P.flags[k];
      h.world = world_;
      h.rank = rank_;
      P.dev_sync.reserve(sizeof(P2PFlags));
      CUDA_OK(cudaMemcpy(P.dev_sync.p, &h, sizeof(h), cudaMemcpyHostToDevice));
      P.ok = true;
      return;
    }
    P.symmetric = false;
    P.local_flags.release();
  }
  const char* env = getenv("B200KV_P2P");
  if (env && env[0] == '0') return;
  if (world_ > kMaxCopies) return;
  Segment::P2P& P = s.p2p;
  NcclApi& nccl = NcclApi::get();
  ncclComm_t comm = reinterpret_cast<ncclComm_t>(comm_);
  struct Handles { cudaIpcMemHandle_t g, w, f, s; };
  const size_t seg_bytes = static_cast<size_t>(s.total) * dtype_size(s.dtype);
  const char* push_e = getenv("B200KV_PUSH");
  const bool want_stage = push_e && push_e[0] == '1';   // only the push variant needs W x slice of staging
  P.local_stage.reserve(want_stage ? std::max<size_t>(seg_bytes, 2u << 20) : (2u << 20));
  P.local_flags.reserve(2u << 20);   // its own 2 MiB allocation: IPC maps whole allocations
  CUDA_OK(cudaMemsetAsync(P.local_flags.p, 0, 256, stream_));
  Handles mine;
  int ok = 1;
  if (cudaIpcGetMemHandle(&mine.g, s.gbuf.p) != cudaSuccess) ok = 0;
  if (cudaIpcGetMemHandle(&mine.w, s.wbuf.p) != cudaSuccess) ok = 0;
  if (cudaIpcGetMemHandle(&mine.f, P.local_flags.p) != cudaSuccess) ok = 0;
  if (cudaIpcGetMemHandle(&mine.s, P.local_stage.p) != cudaSuccess) ok = 0;
  (void)cudaGetLastError();
  DevBuf stage;
  const size_t hb = sizeof(Handles);
  stage.reserve(hb * world_ + 64);
  char* sp = stage.as<char>();
  CUDA_OK(cudaMemcpyAsync(sp + hb * rank_, &mine, hb, cudaMemcpyHostToDevice, stream_));
  NCCL_OK(nccl.AllGather(sp + hb * rank_, sp, hb, ncclChar, comm, stream_));
  std::vector<Handles> all(world_);
  CUDA_OK(cudaMemcpyAsync(all.data(), sp, hb * world_, cudaMemcpyDeviceToHost, stream_));
  CUDA_OK(cudaStreamSynchronize(stream_));
  for (int k = 0; k < world_ && ok; ++k) {
    if (k == rank_) {
      P.g[k] = s.gbuf.p; P.w[k] = s.wbuf.p; P.flags[k] = P.local_flags.as<uint32_t>();
      P.stage[k] = P.local_stage.p;
      continue;
    }
    void* f = nullptr;
    if (cudaIpcOpenMemHandle(&P.g[k], all[k].g, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess ||
        cudaIpcOpenMemHandle(&P.w[k], all[k].w, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess ||
        cudaIpcOpenMemHandle(&f, all[k].f, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess ||
        cudaIpcOpenMemHandle(&P.stage[k], all[k].s, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess)
      ok = 0;
    P.flags[k] = static_cast<uint32_t*>(f);
  }
  (void)cudaGetLastError();
  // agreement: the peer path is used only if every rank mapped every peer
  int* flag = reinterpret_cast<int*>(sp + hb * world_);
  CUDA_OK(cudaMemcpyAsync(flag, &ok, sizeof(int), cudaMemcpyHostToDevice, stream_));
  NCCL_OK(nccl.AllReduce(flag, flag, 1, ncclInt, ncclMin, comm, stream_));
  int agreed = 0;
  CUDA_OK(cudaMemcpyAsync(&agreed, flag, sizeof(int), cudaMemcpyDeviceToHost, stream_));
  CUDA_OK(cudaStreamSynchronize(stream_));
  P.ok = agreed == 1;
  if (P.ok) {
    P2PFlags h;
    std::memset(&h, 0, sizeof(h));
    for (int k = 0; k < world_; ++k) h.flags[k] = P.flags[k];
    h.world = world_;
    h.rank = rank_;
    P.dev_sync.reserve(sizeof(P2PFlags));
    CUDA_OK(cudaMemcpy(P.dev_sync.p, &h, sizeof(h), cudaMemcpyHostToDevice));
  }
  if (!P.ok)
    for (int k = 0; k < world_; ++k) {
      if (k == rank_) continue;
      if (P.g[k]) cudaIpcCloseMemHandle(P.g[k]);
      if (P.w[k]) cudaIpcCloseMemHandle(P.w[k]);
      if (P.flags[k]) cudaIpcCloseMemHandle(P.flags[k]);
      if (P.stage[k]) cudaIpcCloseMemHandle(P.stage[k]);
      P.g[k] = P.w[k] = P.stage[k] = nullptr; P.flags[k] = nullptr;
    }
}

// One fused kernel per step does reduce-scatter + update + all-gather over peer memory: the rank's
// shard of every peer's gradient bucket is read through NVLink and summed in rank order with the
// reference's CommCPU grouping (=> same bits as the CPU reference for any number of GPUs), the
// optimizer runs on the shard, and the new weights are stored into every peer's weight buffer.
void Store::run_class_p2p(DensePlan& p, Segment& s, const std::vector<const Entry*>& ents, bool pull,
                          bool g_alias, bool o_alias) {
  const int nt = static_cast<int>(p.keys.size());
  const size_t es = dtype_size(s.dtype);
  const bool mp = has_opt() && use_mp(s);
  const int W = world_;
  Segment::P2P& P = s.p2p;
  const int ng = p.ngrad, no = pull ? p.nout : 0;
  // LAMB and global-norm clipping make several passes over the merged gradient: reduce the shard
  // into the local bucket once instead of crossing NVLink on every pass
  const bool prereduce = has_opt() && (opt_.kind == B200KV_OPT_LAMB || opt_.clip_global_norm >= 0.f);
  static const bool staged_env = [] { const char* e = getenv("B200KV_STAGED"); return e && e[0] == '1'; }();
  static const bool push_env = [] { const char* e = getenv("B200KV_PUSH"); return e && e[0] == '1'; }();
  const bool staged = staged_env;
  // Reduce-scatter half: PULL (default) — the owner loads its shard of every peer's bucket directly
  // inside the fused kernel; or PUSH (B200KV_PUSH=1) — every rank first stores the slices of its
  // bucket into their owners' staging areas (NVLink writes only), then the owner sums W local copies.
  // Measured equal at 8 GPUs for SGD (0.327 vs 0.325 ms) and slower for LAMB (extra local pass).
  const bool push_rs = push_env && !staged;
  // NVLS: with symmetric buffers and a multicast mapping the NVSwitch does the summing and the
  // replication (multimem.ld_reduce / multimem.st) — one switch-reduced load per element instead
  // of W-1 peer loads, one multicast store instead of W-1 peer stores.  Measured on ResNet-50 SGD:
  // 0.254 ms vs 0.325 ms per step at 8 GPUs, but 0.303 vs 0.193 ms at 2 GPUs (where it saves no
  // link bytes), and slower for LAMB (whose shard is first reduced into the local bucket anyway).
  // Default: on from 4 ranks for single-pass optimizers; B200KV_NVLS=1 / 0 forces it on / off.
  static const int nvls_env = [] { const char* e = getenv("B200KV_NVLS"); return e ? (e[0] == '1' ? 1 : 0) : -1; }();
  const bool nvls_want = nvls_env == 1 || (nvls_env == -1 && W >= 4 && !prereduce);
  // fp16 buckets: the switch accumulates in fp32 and rounds once, the reference rounds to half after
  // every add.  For SGD (the step is linear in the gradient) that stays inside the fp16 tolerance;
  // Adam / AdaGrad divide by |g|-like terms, so where the copies cancel the one-rounding sum moves the
  // step by up to lr (measured: 28x the tolerance on BERT-base) — those always stay on the bit-exact peer
  // path (their fp16 kernels carry no multicast code, optim.cuh Mc16).
  const bool f16_ok = s.dtype == kF32 || !has_opt() || opt_.kind == B200KV_OPT_SGD ||
                      (nvls_env == 1 && opt_.kind == B200KV_OPT_LAMB);
  const bool nvls = nvls_want && f16_ok && P.symmetric && P.mc_g != nullptr && P.mc_w != nullptr && !staged && !push_rs;
  // gradient copy k of the element at flat offset `off`, as seen by the shard owner
  auto grad_src = [&](int k, int64_t off) -> const char* {
    if (!push_rs) return static_cast<const char*>(P.g[k]) + off * es;
    if (k == rank_) return static_cast<const char*>(s.gbuf.p) + off * es;
    return static_cast<const char*>(P.local_stage.p) + (static_cast<int64_t>(k) * s.slice + off - rank_ * s.slice) * es;
  };
  last_exchange_ = nvls ? "nvls" : (push_rs ? "peer-push" : "peer");
  const int pre_ng = nvls ? 1 : W;
  const int upd_ng = (prereduce || nvls) ? 1 : W, upd_no = nvls ? 1 : (mp ? W : W - 1);
  const size_t w_pack = mta_blob_words(nt, ng, 0), w_pre = prereduce ? mta_blob_words(nt, pre_ng, 0) : 0,
               w_upd = mta_blob_words(nt, upd_ng, upd_no), w_unpack = mta_blob_words(nt, 1, no);
  uint64_t* host = ring_.begin((w_pack + w_pre + w_upd + w_unpack) * 8);
  const size_t o_pack = 0, o_pre = w_pack, o_upd = o_pre + w_pre, o_unpack = o_upd + w_upd;
  {
    Blob b(host + o_pack, nt, ng, 0);
    for (int t = 0; t < nt; ++t) {
      for (int k = 0; k < ng; ++k) b.G(k, t, ents[t]->vals[k].data);
      b.S(kSecW, t, at(s.gbuf, keys_[p.keys[t]].off, es));
    }
  }
  if (prereduce) {
    Blob b(host + o_pre, nt, pre_ng, 0);
    for (int t = 0; t < nt; ++t) {
      const int64_t off = keys_[p.keys[t]].off;
      if (nvls) b.G(0, t, static_cast<const char*>(P.mc_g) + off * es);
      else for (int k = 0; k < W; ++k) b.G(k, t, grad_src(k, off));
      b.S(kSecW, t, at(s.gbuf, off, es));
    }
  }
  {
    Blob u(host + o_upd, nt, upd_ng, upd_no);
    for (int t = 0; t < nt; ++t) {
      const KeyInfo& ki = keys_[p.keys[t]];
      if (prereduce) u.G(0, t, at(s.gbuf, ki.off, es));
      else if (nvls) u.G(0, t, static_cast<const char*>(P.mc_g) + ki.off * es);
      else for (int k = 0; k < W; ++k) u.G(k, t, grad_src(k, ki.off));
      if (nvls) {
        u.O(0, t, static_cast<char*>(P.mc_w) + ki.off * es);   // one multicast store reaches every rank
      } else {
        int o = 0;
        for (int k = 0; k < W; ++k) {
          if (!mp && k == rank_) continue;   // fp32: the W section already is the local weight buffer
          u.O(o++, t, static_cast<char*>(P.w[k]) + ki.off * es);
        }
      }
      u.S(kSecW, t, mp ? at(s.master, ki.off - s.state_begin, 4) : at(s.wbuf, ki.off, es));
      if (!has_opt()) continue;   // no updater: the merged value is assigned (kvstore_local.h:268-274)
      if (s.has_s0) u.S(kSecS0, t, at(s.s0, ki.off - s.state_begin, 4));
      if (s.has_s1) u.S(kSecS1, t, at(s.s1, ki.off - s.state_begin, 4));
      if (s.xbuf.p) u.S(kSecX, t, at(s.xbuf, ki.off - s.state_begin, 4));
      u.F2(kSecLrWd, t, ki.lr, ki.wd);
      if (opt_.kind == B200KV_OPT_LAMB) {
        const float tf = static_cast<float>(ki.step);
        u.F2(kSecAux, t, 1.0f - powf(opt_.beta1, tf), 1.0f - powf(opt_.beta2, tf));
      }
    }
  }
  if (pull) {
    Blob b(host + o_unpack, nt, 1, no);
    for (int t = 0; t < nt; ++t) {
      b.G(0, t, at(s.wbuf, keys_[p.keys[t]].off, es));
      for (int k = 0; k < no; ++k) b.O(k, t, ents[t]->outs[k].data);
    }
  }
  const uint64_t* dev = ring_.commit(stream_);
  MtaCall base;
  base.nt = nt; base.dtype = s.dtype; base.stream = stream_; base.device = device_;
  base.grid_div = grid_share_;

  MtaCall c = base;
  c.blob = dev + o_pack; c.ngrad = ng; c.nout = 0; c.plan = &p.full;
  if (!g_alias) {
    prof_begin(1);
    launches_ += launch_copy(c, true, false);
    prof_end();
  }

  P2PFlags f;
  for (int k = 0; k < kMaxCopies; ++k) f.flags[k] = P.flags[k];
  f.world = W; f.rank = rank_; f.epoch = ++P.epoch;
  if (push_rs) {
    void* dst[kMaxCopies] = {nullptr};
    for (int k = 0; k < W; ++k)
      if (k != rank_) dst[k] = static_cast<char*>(P.stage[k]) + static_cast<int64_t>(rank_) * s.slice * es;
    prof_begin(1);
    launches_ += launch_p2p_push(s.gbuf.p, dst, s.slice * static_cast<int64_t>(es), W, rank_, device_, stream_);
    prof_end();
  }
  // Barriers: folded into the first kernel that reads the exchanged gradients / the last one that
  // writes peer weights (plain kernels), or separate one-warp kernels (TMA-staged kernels).  With
  // the push reduce-scatter "ready" means "my pushes have landed" (the push kernel precedes the
  // signalling kernel in stream order).
  const bool fold = !staged;
  const P2PFlags* dsync = P.dev_sync.as<P2PFlags>();
  prof_begin(2);
  if (!fold) launches_ += launch_p2p_barrier(f, 0, stream_);       // every rank's bucket is complete
  if (prereduce) {
    MtaCall r = base;
    r.blob = dev + o_pre; r.ngrad = pre_ng; r.nout = 0; r.plan = &p.shard; r.staged = staged;
    if (fold) { r.sync = dsync; r.epoch = f.epoch; r.sync_mode = kSyncWaitReady; }
    if (nvls) r.sync_mode |= kSyncMcGrad;
    launches_ += launch_copy(r, true, false);
  }
  prof_end();

  MtaCall u = base;
  u.blob = dev + o_upd; u.ngrad = upd_ng; u.nout = upd_no; u.plan = &p.shard; u.staged = staged && !prereduce;
  if (fold) { u.sync = dsync; u.epoch = f.epoch; u.sync_mode = (prereduce ? 0 : kSyncWaitReady) | kSyncSignalDone; }
  if (nvls) u.sync_mode |= kSyncMcOut | (prereduce ? 0 : kSyncMcGrad);
  // CTAs per SM of the exchange kernel (0 = what fits).  A bare switch-reduce + multicast loop is fastest
  // with ONE CTA per SM (tools/nvls_probe.cu, profiles/r2_nvls_probe.txt: 0.216 ms against 0.256 ms with
  // four at 8 GPUs), but the update kernel measured the same with 1, 2 or all resident CTAs
  // (profiles/r2_sweep_ctas_n4.txt), so the default stays "what fits"; the knobs remain for tuning.
  static const int nvls_ctas = [] { const char* e = getenv("B200KV_NVLS_CTAS"); return e ? atoi(e) : 0; }();
  static const int peer_ctas = [] { const char* e = getenv("B200KV_PEER_CTAS"); return e ? atoi(e) : 0; }();
  u.max_per_sm = nvls ? nvls_ctas : peer_ctas;
  prof_begin(0);
  if (has_opt()) {
    const float* gscale = opt_.clip_global_norm >= 0.f ? global_norm_scale(p, p.shard, u) : nullptr;
    update_chunks(p, p.shard, u, gscale);
  } else {
    u.plan = &p.shard;
    launches_ += launch_copy(u, true, true);   // merged shard -> the stored value on every rank
  }
  prof_end();

  if (!fold) {
    prof_begin(3);
    launches_ += launch_p2p_barrier(f, 1, stream_);       // every rank's shard has landed everywhere
    prof_end();
  }

  if (pull && !o_alias) {
    MtaCall q = base;
    q.blob = dev + o_unpack; q.ngrad = 1; q.nout = no; q.plan = &p.full;
    prof_begin(4);
    launches_ += launch_copy(q, false, true);
    prof_end();
  }
  ring_.end(stream_);
}

void Store::pull_staged(const Entry& e) {
  const KeyInfo& ki = keys_[e.key];
  const size_t bytes = static_cast<size_t>(ki.size) * dtype_size(ki.dtype);
  for (const auto& o : e.outs)
    if (bytes) CUDA_OK(cudaMemcpyAsync(o.data, ki.staged->p, bytes, cudaMemcpyDeviceToDevice, stream_));
}

void Store::prepare(Mode mode, int vnum, const int* vkeys, DLManagedTensor* const* vals, int onum,
                    const int* okeys, DLManagedTensor* const* outs, BoundCall* bc, bool allow_staged) {
  // GroupKVPairs (kvstore_local.h:417-447): entries with the same key are device copies of it
  std::vector<Entry>& ents = bc->ents;
  std::unordered_map<int, size_t> where;
  auto entry = [&](int key) -> Entry& {
    auto it = where.find(key);
    if (it == where.end()) {
      it = where.emplace(key, ents.size()).first;
      ents.push_back(Entry{key, {}, {}});
    }
    return ents[it->second];
  };
  auto check = [&](const TensorView& v, int key, const char* what) {
    KV_CHECK(rsp_.find(key) == rsp_.end(), what << ": key " << key
                                                << " is row_sparse; use the row-sparse entry points");
    const KeyInfo& ki = info(key);
    KV_CHECK(v.dtype == ki.dtype, what << ": dtype mismatch for key " << key);
    KV_CHECK(v.numel == ki.size, what << ": key " << key << " holds " << ki.size << " elements, got "
                                      << v.numel);
  };
  if (mode & kPush)
    for (int i = 0; i < vnum; ++i) {
      const TensorView v = view_of(vals[i], device_, "push value");
      check(v, vkeys[i], "push");
      entry(vkeys[i]).vals.push_back(v);
    }
  if (mode & kPull)
    for (int i = 0; i < onum; ++i) {
      const TensorView v = view_of(outs[i], device_, "pull out");
      check(v, okeys[i], "pull");
      entry(okeys[i]).outs.push_back(v);
    }
  bool need_seal = false;
  for (const Entry& e : ents)
    if (!e.vals.empty() && keys_[e.key].seg < 0) need_seal = true;
  if (need_seal) seal();

  // classes: (pushes?, segment, #vals, #outs) -> entries sorted by flat offset
  std::vector<Cls>& classes = bc->classes;
  for (const Entry& e : ents) {
    const KeyInfo& ki = keys_[e.key];
    if (ki.seg < 0) {  // only pulled, never pushed yet: serve from the staged init value
      KV_CHECK(allow_staged, "cannot bind a pull of key " << e.key << " before its first push");
      pull_staged(e);
      continue;
    }
    const int m = (e.vals.empty() ? 0 : kPush) | (e.outs.empty() ? 0 : kPull);
    const int ng = static_cast<int>(e.vals.size()), no = static_cast<int>(e.outs.size());
    KV_CHECK(ng <= kMaxCopies && no <= kMaxCopies, "at most " << kMaxCopies << " device-local copies per key");
    Cls* c = nullptr;
    for (auto& x : classes)
      if (x.mode == m && x.seg == ki.seg && x.ng == ng && x.no == no) c = &x;
    if (!c) {
      classes.push_back(Cls{m, ki.seg, ng, no, {}, nullptr});
      c = &classes.back();
    }
    c->ents.push_back(&e);
  }
  // pushes first, then pure pulls (PushPullImpl = PushImpl; PullImpl, kvstore_local.h:336-343)
  std::stable_sort(classes.begin(), classes.end(), [](const Cls& a, const Cls& b) {
    return ((a.mode & kPush) != 0) > ((b.mode & kPush) != 0);
  });
  for (auto& c : classes) {
    std::sort(c.ents.begin(), c.ents.end(), [&](const Entry* a, const Entry* b) {
      return keys_[a->key].off < keys_[b->key].off;
    });
    std::vector<int> ks;
    ks.reserve(c.ents.size());
    for (const Entry* e : c.ents) ks.push_back(e->key);
    c.plan = &plan_for(c.seg, ks, c.ng, c.no);
  }
}

void Store::preallocate(BoundCall& bc) {
  size_t table_bytes = 0;
  for (auto& c : bc.classes)   // upper bound of the tables run_class / run_class_p2p carve out of a ring slot
    table_bytes = std::max(table_bytes, 8 * 4 * mta_blob_words(static_cast<int>(c.plan->keys.size()),
                                                                std::max(c.ng, world_), std::max(c.no, world_)));
  ring_.reserve_all(table_bytes);
  for (auto& c : bc.classes) {
    if (!(c.mode & kPush) || !has_opt()) continue;
    Segment& s = *segs_[c.plan->seg];
    if (s.dtype == kF16 && !opt_.multi_precision) continue;   // rejected with a message in run_class
    ensure_states(s);
  }
}

void Store::execute(BoundCall& bc) {
  for (auto& c : bc.classes) run_class(static_cast<Mode>(c.mode), *c.plan, c.ents);
}

void Store::dense(Mode mode, int vnum, const int* vkeys, DLManagedTensor* const* vals, int onum,
                  const int* okeys, DLManagedTensor* const* outs) {
  use_device();
  BoundCall bc;
  bc.ents.reserve(static_cast<size_t>(vnum + onum));
  prepare(mode, vnum, vkeys, vals, onum, okeys, outs, &bc, true);
  execute(bc);
}

int Store::bind(Mode mode, int vnum, const int* vkeys, DLManagedTensor* const* vals, int onum,
                const int* okeys, DLManagedTensor* const* outs) {
  use_device();
  std::unique_ptr<BoundCall> bc(new BoundCall());
  bc->ents.reserve(static_cast<size_t>(vnum + onum));   // Entry addresses must stay put
  prepare(mode, vnum, vkeys, vals, onum, okeys, outs, bc.get(), false);
  const int id = next_bound_++;
  bound_.emplace(id, std::move(bc));
  return id;
}

void Store::run_bound(int id) {
  use_device();
  auto it = bound_.find(id);
  KV_CHECK(it != bound_.end(), "unknown bound call " << id);
  execute(*it->second);
}

void Store::unbind(int id) { bound_.erase(id); }

// ---- kernel timing ---------------------------------------------------------------------------------
void Store::prof_begin(int phase) {
  if (!profiling_) return;
  if (prof_used_ == prof_events_.size()) {
    cudaEvent_t a, b;
    CUDA_OK(cudaEventCreate(&a));
    CUDA_OK(cudaEventCreate(&b));
    prof_events_.emplace_back(a, b);
    prof_phase_.push_back(0);
  }
  prof_phase_[prof_used_] = phase;
  CUDA_OK(cudaEventRecord(prof_events_[prof_used_].first, stream_));
}

void Store::prof_end() {
  if (!profiling_) return;
  CUDA_OK(cudaEventRecord(prof_events_[prof_used_].second, stream_));
  ++prof_used_;
}

void Store::profile_read_phases(double ms[8], int64_t count[8]) {
  use_device();
  CUDA_OK(cudaStreamSynchronize(stream_));
  for (int i = 0; i < 8; ++i) { ms[i] = 0; count[i] = 0; }
  for (size_t i = 0; i < prof_used_; ++i) {
    float t = 0;
    CUDA_OK(cudaEventElapsedTime(&t, prof_events_[i].first, prof_events_[i].second));
    ms[prof_phase_[i] & 7] += t;
    count[prof_phase_[i] & 7] += 1;
  }
  prof_used_ = 0;
}

void Store::profile_read(double* ms, int64_t* count) {
  double m[8];
  int64_t c[8];
  profile_read_phases(m, c);
  *ms = m[0];
  *count = c[0];
}

// ---- zero-copy views --------------------------------------------------------------------------------
namespace {
struct ViewBox {
  DLManagedTensor t;
  int64_t shape[1];
};
void view_deleter(DLManagedTensor* self) { delete reinterpret_cast<ViewBox*>(self->manager_ctx); }
}  // namespace

DLManagedTensor* Store::view(int key, int which) {
  use_device();
  KeyInfo& ki = info(key);
  if (ki.seg < 0) seal();
  Segment& s = *segs_[keys_[key].seg];
  KV_CHECK(which >= 0 && which <= 3, "view selector must be 0..3");
  const size_t es = dtype_size(s.dtype);
  const bool grad = which == 1 || which == 3;
  if (grad && s.gbuf.p == nullptr) alloc_gbuf(s);
  KV_CHECK(!(which == 0 || which == 2) || !s.w_stale,
           "weight views must be taken before the first multi-precision update");
  ViewBox* b = new ViewBox();
  const bool whole = which >= 2;
  b->shape[0] = whole ? s.total : keys_[key].size;
  DLTensor& d = b->t.dl_tensor;
  d.data = at(grad ? s.gbuf : s.wbuf, whole ? 0 : keys_[key].off, es);
  d.device.device_type = kDLCUDA;
  d.device.device_id = device_;
  d.ndim = 1;
  d.dtype.code = kDLFloat;
  d.dtype.bits = static_cast<uint8_t>(es * 8);
  d.dtype.lanes = 1;
  d.shape = b->shape;
  d.strides = nullptr;
  d.byte_offset = 0;
  b->t.manager_ctx = b;
  b->t.deleter = view_deleter;
  return &b->t;
}

// ---- optimizer state access (checkpointing) --------------------------------------------------------
void Store::get_state(int key, int which, DLManagedTensor* out, int* present) {
  use_device();
  const KeyInfo& ki = info(key);
  *present = 0;
  if (ki.seg < 0) return;
  Segment& s = *segs_[ki.seg];
  const DevBuf* src = which == 0 ? (s.has_s0 ? &s.s0 : nullptr)
                    : which == 1 ? (s.has_s1 ? &s.s1 : nullptr)
                    : which == 2 ? (s.has_master ? &s.master : nullptr) : nullptr;
  KV_CHECK(which >= 0 && which <= 2, "state selector must be 0, 1 or 2");
  if (!src) return;
  const TensorView v = view_of(out, device_, "state out");
  KV_CHECK(v.dtype == kF32 && v.numel == ki.size, "state out must be float32 with " << ki.size << " elements");
  const int64_t lo = std::max(ki.off, s.state_begin), hi = std::min(ki.off + ki.size, s.state_begin + s.state_len);
  if (world_ > 1) CUDA_OK(cudaMemsetAsync(v.data, 0, static_cast<size_t>(ki.size) * 4, stream_));
  if (lo < hi)
    CUDA_OK(cudaMemcpyAsync(static_cast<float*>(v.data) + (lo - ki.off), at(*src, lo - s.state_begin, 4),
                            static_cast<size_t>(hi - lo) * 4, cudaMemcpyDeviceToDevice, stream_));
  if (world_ > 1)  // every rank contributes its shard; the others hold zeros
    NCCL_OK(NcclApi::get().AllReduce(v.data, v.data, ki.size, ncclFloat32, ncclSum,
                                     reinterpret_cast<ncclComm_t>(comm_), stream_));
  *present = 1;
}

void Store::set_state(int key, int which, DLManagedTensor* in) {
  use_device();
  const KeyInfo& ki = info(key);
  KV_CHECK(which >= 0 && which <= 2, "state selector must be 0, 1 or 2");
  KV_CHECK(has_opt(), "set_state needs an optimizer");
  if (ki.seg < 0) seal();
  Segment& s = *segs_[keys_[key].seg];
  ensure_states(s);
  DevBuf* dst = which == 0 ? (s.has_s0 ? &s.s0 : nullptr)
              : which == 1 ? (s.has_s1 ? &s.s1 : nullptr) : (s.has_master ? &s.master : nullptr);
  KV_CHECK(dst, "the current optimizer keeps no such state");
  const TensorView v = view_of(in, device_, "state in");
  KV_CHECK(v.dtype == kF32 && v.numel == ki.size, "state must be float32 with " << ki.size << " elements");
  const int64_t lo = std::max(ki.off, s.state_begin), hi = std::min(ki.off + ki.size, s.state_begin + s.state_len);
  if (lo < hi)
    CUDA_OK(cudaMemcpyAsync(at(*dst, lo - s.state_begin, 4), static_cast<const float*>(v.data) + (lo - ki.off),
                            static_cast<size_t>(hi - lo) * 4, cudaMemcpyDeviceToDevice, stream_));
  if (which == 2) {
    // the fp16 stored value follows its master
    if (world_ == 1) {
      s.w_stale = true;
    } else {
      if (lo < hi)
        launches_ += launch_cast(at(*dst, lo - s.state_begin, 4), kF32, at(s.wbuf, lo, 2), kF16, hi - lo, stream_);
      // every rank refreshed only its own shard of the fp16 value: exchange the pieces so that a
      // pull before the next push returns the loaded weights everywhere
      KV_CHECK(comm_, "b200kv_comm_init has not been called");
      NcclApi& nccl = NcclApi::get();
      NCCL_OK(nccl.GroupStart());
      for (int r = 0; r < world_; ++r) {
        const int64_t plo = std::max(ki.off, r * s.slice), phi = std::min(ki.off + ki.size, (r + 1) * s.slice);
        if (plo < phi)
          NCCL_OK(nccl.Broadcast(at(s.wbuf, plo, 2), at(s.wbuf, plo, 2), phi - plo, ncclFloat16, r,
                                 reinterpret_cast<ncclComm_t>(comm_), stream_));
      }
      NCCL_OK(nccl.GroupEnd());
    }
  }
}

// ---- members of a Group (one process driving several GPUs) ------------------------------------------
void Store::init_any_device(int key, DLManagedTensor* val) {
  use_device();
  if (key_type_ < 0) key_type_ = 0;
  KV_CHECK(keys_.find(key) == keys_.end() && rsp_.find(key) == rsp_.end(), "duplicate init of key " << key);
  KV_CHECK(val != nullptr, "init value: null DLManagedTensor");
  const int src_dev = val->dl_tensor.device.device_id;
  const TensorView v = view_of(val, src_dev, "init value");
  KV_CHECK(v.dtype == kF32 || v.dtype == kF16, "init: only float32/float16 values are supported");
  KeyInfo k;
  k.key = key;
  k.dtype = v.dtype;
  k.size = v.numel;
  k.staged.reset(new DevBuf());
  const size_t bytes = static_cast<size_t>(v.numel) * dtype_size(v.dtype);
  k.staged->reserve(bytes ? bytes : 16);
  if (bytes) {
    if (src_dev == device_)
      CUDA_OK(cudaMemcpyAsync(k.staged->p, v.data, bytes, cudaMemcpyDeviceToDevice, stream_));
    else
      CUDA_OK(cudaMemcpyPeerAsync(k.staged->p, device_, v.data, src_dev, bytes, stream_));
  }
  keys_.emplace(key, std::move(k));
  unsealed_.push_back(key);
}

void Store::link_peers(int seg, const std::vector<Store*>& all) {
  use_device();
  Segment& s = *segs_[seg];
  Segment::P2P& P = s.p2p;
  if (P.ok) return;
  P.tried = true;
  // the switch exchange only when EVERY member's bucket and weight segment sit behind the same multicast
  // objects (Group's McArena); one member on plain memory puts all of them on the peer load/store path
  bool all_mc = true;
  for (int k = 0; k < world_; ++k) {
    const Segment::P2P& o = all[k]->segs_[seg]->p2p;
    all_mc = all_mc && o.mc_g != nullptr && o.mc_w != nullptr && o.mc_g == all[0]->segs_[seg]->p2p.mc_g &&
             o.mc_w == all[0]->segs_[seg]->p2p.mc_w;
  }
  P.symmetric = all_mc;
  for (int k = 0; k < world_; ++k) {
    Segment& o = *all[k]->segs_[seg];
    KV_CHECK(o.total == s.total && o.dtype == s.dtype, "internal: members disagree on the segment layout");
    P.g[k] = o.gbuf.p;
    P.w[k] = o.wbuf.p;
    P.flags[k] = o.p2p.local_flags.as<uint32_t>();
    P.stage[k] = nullptr;
  }
  P2PFlags h;
  std::memset(&h, 0, sizeof(h));
  for (int k = 0; k < world_; ++k) h.flags[k] = P.flags[k];
  h.world = world_;
  h.rank = rank_;
  P.dev_sync.reserve(sizeof(P2PFlags));
  CUDA_OK(cudaMemcpy(P.dev_sync.p, &h, sizeof(h), cudaMemcpyHostToDevice));
  P.ok = true;
}

bool Store::copy_state_shard(int key, int which, float* dst, int dst_device, int64_t numel) {
  use_device();
  const KeyInfo& ki = info(key);
  KV_CHECK(which >= 0 && which <= 2, "state selector must be 0, 1 or 2");
  KV_CHECK(numel == ki.size, "state out must hold " << ki.size << " elements");
  if (ki.seg < 0) return false;
  Segment& s = *segs_[ki.seg];
  const DevBuf* src = which == 0 ? (s.has_s0 ? &s.s0 : nullptr)
                    : which == 1 ? (s.has_s1 ? &s.s1 : nullptr) : (s.has_master ? &s.master : nullptr);
  if (!src) return false;
  const int64_t lo = std::max(ki.off, s.state_begin), hi = std::min(ki.off + ki.size, s.state_begin + s.state_len);
  if (lo < hi)
    CUDA_OK(cudaMemcpyPeerAsync(dst + (lo - ki.off), dst_device, at(*src, lo - s.state_begin, 4), device_,
                                static_cast<size_t>(hi - lo) * 4, stream_));
  return true;
}

void Store::load_state_shard(int key, int which, const float* src, int src_device, int64_t numel) {
  use_device();
  const KeyInfo& ki0 = info(key);
  KV_CHECK(which >= 0 && which <= 2, "state selector must be 0, 1 or 2");
  KV_CHECK(has_opt(), "set_state needs an optimizer");
  KV_CHECK(numel == ki0.size, "state must hold " << ki0.size << " elements");
  if (ki0.seg < 0) seal();
  const KeyInfo& ki = keys_[key];
  Segment& s = *segs_[ki.seg];
  ensure_states(s);
  DevBuf* dst = which == 0 ? (s.has_s0 ? &s.s0 : nullptr)
              : which == 1 ? (s.has_s1 ? &s.s1 : nullptr) : (s.has_master ? &s.master : nullptr);
  KV_CHECK(dst, "the current optimizer keeps no such state");
  const int64_t lo = std::max(ki.off, s.state_begin), hi = std::min(ki.off + ki.size, s.state_begin + s.state_len);
  if (lo < hi) {
    CUDA_OK(cudaMemcpyPeerAsync(at(*dst, lo - s.state_begin, 4), device_, src + (lo - ki.off), src_device,
                                static_cast<size_t>(hi - lo) * 4, stream_));
    if (which == 2)   // the fp16 stored value follows its master
      launches_ += launch_cast(at(*dst, lo - s.state_begin, 4), kF32, at(s.wbuf, lo, 2), kF16, hi - lo, stream_);
  }
}

void Store::spread_fp16_shard(int key, const std::vector<Store*>& all) {
  use_device();
  const KeyInfo& ki = info(key);
  Segment& s = *segs_[ki.seg];
  const int64_t lo = std::max(ki.off, s.state_begin), hi = std::min(ki.off + ki.size, s.state_begin + s.state_len);
  if (lo >= hi) return;
  for (Store* o : all) {
    if (o == this) continue;
    Segment& os = *o->segs_[ki.seg];
    CUDA_OK(cudaMemcpyPeerAsync(at(os.wbuf, lo, 2), o->device_, at(s.wbuf, lo, 2), device_,
                                static_cast<size_t>(hi - lo) * 2, stream_));
  }
}

}  // namespace b200kv
