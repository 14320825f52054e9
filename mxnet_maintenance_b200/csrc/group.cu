// Group: one process, several GPUs (see group.h).  Reference semantics: KVStoreLocal with CommDevice
// (src/kvstore/kvstore_local.h:226-343, src/kvstore/comm.h:451-797) and KVStoreNCCL
// (src/kvstore/kvstore_nccl.h:89-175).
#include "group.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace b200kv {

namespace {
struct DeviceGuard {
  int prev = 0;
  explicit DeviceGuard(int d) {
    cudaGetDevice(&prev);
    CUDA_OK(cudaSetDevice(d));
  }
  ~DeviceGuard() { cudaSetDevice(prev); }
};
}  // namespace

Group::Group(int ndev, const int* dev_ids) {
  is_group = true;
  KV_CHECK(ndev >= 1 && ndev <= kMaxCopies, "a group drives between 1 and " << kMaxCopies << " devices, got " << ndev);
  int visible = 0;
  CUDA_OK(cudaGetDeviceCount(&visible));
  for (int i = 0; i < ndev; ++i)
    KV_CHECK(dev_ids[i] >= 0 && dev_ids[i] < visible, "invalid CUDA device " << dev_ids[i] << " (" << visible << " visible)");
  // CommDevice::EnableP2P (comm.h:715-757): every pair of distinct devices must reach each other
  for (int i = 0; i < ndev; ++i)
    for (int j = 0; j < ndev; ++j) {
      if (dev_ids[i] == dev_ids[j]) continue;
      int can = 0;
      CUDA_OK(cudaDeviceCanAccessPeer(&can, dev_ids[i], dev_ids[j]));
      KV_CHECK(can, "GPU " << dev_ids[i] << " cannot access GPU " << dev_ids[j]
                           << " (peer access is required; use one process per GPU with NCCL instead)");
      DeviceGuard g(dev_ids[i]);
      const cudaError_t e = cudaDeviceEnablePeerAccess(dev_ids[j], 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled) (void)cudaGetLastError();
      else CUDA_OK(e);
    }
  if (ndev > 1)
    for (int i = 0; i < ndev; ++i) {
      if (std::find(dev_ids, dev_ids + i, dev_ids[i]) != dev_ids + i) continue;
      DeviceGuard g(dev_ids[i]);
      preload_kernels();   // no lazy kernel load may fall between two members' launches (common.h)
    }
  {
    // same rule as Store::run_class_p2p: the switch exchange pays from 4 GPUs, B200KV_NVLS=1 / 0 forces it on / off
    const char* e = getenv("B200KV_NVLS");
    const bool want = e ? e[0] == '1' : ndev >= 4;
    if (ndev > 1 && want) arena_ = McArena::create(std::vector<int>(dev_ids, dev_ids + ndev));
    mc_ctx_.resize(ndev);
  }
  for (int i = 0; i < ndev; ++i) {
    const int share = static_cast<int>(std::count(dev_ids, dev_ids + ndev, dev_ids[i]));
    m_.emplace_back(new Store(dev_ids[i], i, ndev, /*in_process=*/ndev > 1, share));
    raw_.push_back(m_.back().get());
    if (arena_) {
      mc_ctx_[i] = McCtx{arena_.get(), i};
      m_.back()->set_symmetric_allocator(&Group::mc_alloc, &mc_ctx_[i]);
    }
    DeviceGuard g(dev_ids[i]);
    cudaStream_t s = nullptr;
    cudaEvent_t a = nullptr, b = nullptr;
    if (ndev > 1) {
      // the members' kernels wait for each other on the GPUs, so they need streams of their own:
      // on the legacy default stream two members sharing a device would serialise and dead-lock
      CUDA_OK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
      CUDA_OK(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
      CUDA_OK(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
      m_.back()->set_stream(s);
    }
    own_.push_back(s);
    caller_.push_back(nullptr);
    ev_in_.push_back(a);
    ev_out_.push_back(b);
  }
}

int Group::mc_alloc(void* ctx, size_t bytes, void** local, void** peers, void** mc) {
  McCtx* c = static_cast<McCtx*>(ctx);
  return c->arena->request(c->member, bytes, local, peers, mc) ? 0 : 1;
}

Group::~Group() {
  int prev = 0;
  cudaGetDevice(&prev);
  struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};   // the caller's current device stays
  for (size_t i = 0; i < m_.size(); ++i) {
    cudaSetDevice(m_[i]->device());
    if (own_[i]) cudaStreamSynchronize(own_[i]);
  }
  m_.clear();   // members free their buffers (nothing of a sibling is unmapped: plain peer pointers)
  for (size_t i = 0; i < own_.size(); ++i) {
    if (ev_in_[i]) cudaEventDestroy(ev_in_[i]);
    if (ev_out_[i]) cudaEventDestroy(ev_out_[i]);
    if (own_[i]) cudaStreamDestroy(own_[i]);
  }
}

void Group::set_streams(int n, void* const* streams) {
  KV_CHECK(n == size(), "set_streams: expected one stream per member (" << size() << "), got " << n);
  for (int i = 0; i < n; ++i) {
    caller_[i] = static_cast<cudaStream_t>(streams[i]);
    if (size() == 1) m_[0]->set_stream(caller_[0]);
  }
}

// the members' streams pick up after the caller's work on their device ...
void Group::fence_in(bool all_to_all) {
  if (size() == 1) return;
  for (int i = 0; i < size(); ++i) {
    DeviceGuard g(m_[i]->device());
    CUDA_OK(cudaEventRecord(ev_in_[i], caller_[i]));
  }
  for (int i = 0; i < size(); ++i) {
    DeviceGuard g(m_[i]->device());
    if (all_to_all) {   // this member reads arrays that live on its siblings' devices
      for (int j = 0; j < size(); ++j) CUDA_OK(cudaStreamWaitEvent(own_[i], ev_in_[j], 0));
    } else {
      CUDA_OK(cudaStreamWaitEvent(own_[i], ev_in_[i], 0));
    }
  }
}

// ... and the caller's streams pick up after the members' work
void Group::fence_out() {
  if (size() == 1) return;
  for (int i = 0; i < size(); ++i) {
    DeviceGuard g(m_[i]->device());
    CUDA_OK(cudaEventRecord(ev_out_[i], own_[i]));
    CUDA_OK(cudaStreamWaitEvent(caller_[i], ev_out_[i], 0));
  }
}

void Group::sync_callers() {
  for (int i = 0; i < size(); ++i) {
    DeviceGuard g(m_[i]->device());
    CUDA_OK(cudaStreamSynchronize(caller_[i]));
  }
}

void Group::sync_members() {
  for (int i = 0; i < size(); ++i) {
    DeviceGuard g(m_[i]->device());
    CUDA_OK(cudaStreamSynchronize(m_[i]->stream()));
  }
}

void Group::wait() {
  for (auto& m : m_) {
    DeviceGuard g(m->device());
    m->wait();
  }
}

void Group::barrier() { sync_members(); }

void Group::lookup_keys(int num, const char* const* names, bool create, int* out) {
  for (auto& m : m_) m->lookup_keys(num, names, create, out);   // same names, same order => same ints
}

int Group::member_of_device(int device, int nth) const {
  int cnt = 0, first = -1;
  for (int i = 0; i < size(); ++i)
    if (m_[i]->device() == device) {
      if (first < 0) first = i;
      ++cnt;
    }
  KV_CHECK(cnt > 0, "array on GPU " << device << ", which is not a device of this kvstore group");
  int k = nth % cnt;
  for (int i = 0; i < size(); ++i)
    if (m_[i]->device() == device && k-- == 0) return i;
  return first;
}

Group::Split Group::split(int num, const int* keys, DLManagedTensor* const* arrs, const char* what) const {
  Split sp;
  sp.keys.resize(size());
  sp.arrs.resize(size());
  // several members may drive one physical GPU (virtual ranks): the copies of a key that live on
  // that GPU are dealt to them in order of appearance
  std::unordered_map<int64_t, int> seen;   // (key, device) -> copies so far
  for (int i = 0; i < num; ++i) {
    KV_CHECK(arrs[i] != nullptr, what << ": null DLManagedTensor");
    KV_CHECK(arrs[i]->dl_tensor.device.device_type == kDLCUDA, what << ": tensor must live in CUDA device memory");
    const int dev = arrs[i]->dl_tensor.device.device_id;
    const int64_t id = (static_cast<int64_t>(keys[i]) << 8) | static_cast<int64_t>(dev & 0xff);
    const int nth = seen[id]++;
    const int m = member_of_device(dev, nth);
    sp.keys[m].push_back(keys[i]);
    sp.arrs[m].push_back(arrs[i]);
  }
  return sp;
}

void Group::init(int num, const int* keys, DLManagedTensor* const* vals) {
  // KVStoreLocal::InitImpl (kvstore_local.h:226-237): ONE value per key, copied into the store
  if (size() == 1) return m_[0]->init(num, keys, vals);
  sync_callers();   // the value may live on any device of the group
  for (int i = 0; i < num; ++i)
    for (auto& m : m_) {
      DeviceGuard g(m->device());
      m->init_any_device(keys[i], vals[i]);
    }
}

void Group::link_all() {
  if (size() == 1) return;
  const int nseg = m_[0]->num_segments();
  for (auto& m : m_) KV_CHECK(m->num_segments() == nseg, "internal: members disagree on the number of segments");
  bool fresh = false;
  for (int sgi = 0; sgi < nseg; ++sgi) {
    if (m_[0]->segment_linked(sgi)) continue;
    fresh = true;
    for (auto& m : m_) {   // CommDevice::InitMergeBuffer (comm.h:678-712): every member's bucket + flag block
      DeviceGuard g(m->device());
      Segment& s = *m->segs_[sgi];
      if (s.gbuf.p == nullptr) m->alloc_gbuf(s);
      s.p2p.local_flags.reserve(2u << 20);
      CUDA_OK(cudaMemsetAsync(s.p2p.local_flags.p, 0, 256, m->stream()));
    }
    for (auto& m : m_) m->link_peers(sgi, raw_);
  }
  // every flag block is zero before any sibling may signal into it
  if (fresh) sync_members();
}

void Group::dense(Store::Mode mode, int vnum, const int* vkeys, DLManagedTensor* const* vals, int onum,
                  const int* okeys, DLManagedTensor* const* outs) {
  if (size() == 1) return m_[0]->dense(mode, vnum, vkeys, vals, onum, okeys, outs);
  const Split v = split(vnum, vkeys, vals, "push value"), o = split(onum, okeys, outs, "pull out");
  if (mode & Store::kPush)   // same constraint as KVStoreNCCL (kvstore_nccl.h:244-250): every device takes part
    for (int i = 0; i < size(); ++i)
      KV_CHECK(v.keys[i].size() == v.keys[0].size() && !v.keys[i].empty(),
               "every key must be pushed with one value per device of the group (device "
                   << m_[i]->device() << " got " << v.keys[i].size() << " values, device " << m_[0]->device()
                   << " got " << v.keys[0].size() << ")");
  fence_in();
  std::vector<Store::BoundCall> bcs(size());
  for (int i = 0; i < size(); ++i) {
    DeviceGuard g(m_[i]->device());
    bcs[i].ents.reserve(v.keys[i].size() + o.keys[i].size());
    m_[i]->prepare(mode, static_cast<int>(v.keys[i].size()), v.keys[i].data(), v.arrs[i].data(),
                   static_cast<int>(o.keys[i].size()), o.keys[i].data(), o.arrs[i].data(), &bcs[i], true);
  }
  for (int i = 0; i < size(); ++i) {
    DeviceGuard g(m_[i]->device());
    m_[i]->preallocate(bcs[i]);
  }
  link_all();
  for (int i = 0; i < size(); ++i) {
    DeviceGuard g(m_[i]->device());
    m_[i]->execute(bcs[i]);
  }
  fence_out();
}

int Group::bind(Store::Mode mode, int vnum, const int* vkeys, DLManagedTensor* const* vals, int onum,
                const int* okeys, DLManagedTensor* const* outs) {
  if (size() == 1) return m_[0]->bind(mode, vnum, vkeys, vals, onum, okeys, outs);
  const Split v = split(vnum, vkeys, vals, "push value"), o = split(onum, okeys, outs, "pull out");
  for (int i = 0; i < size(); ++i)
    KV_CHECK(v.keys[i].size() == v.keys[0].size() && !v.keys[i].empty(),
             "every key must be pushed with one value per device of the group");
  std::vector<int> ids;
  for (int i = 0; i < size(); ++i) {
    DeviceGuard g(m_[i]->device());
    ids.push_back(m_[i]->bind(mode, static_cast<int>(v.keys[i].size()), v.keys[i].data(), v.arrs[i].data(),
                              static_cast<int>(o.keys[i].size()), o.keys[i].data(), o.arrs[i].data()));
  }
  link_all();
  const int id = next_bound_++;
  bound_.emplace(id, std::move(ids));
  return id;
}

void Group::run_bound(int id) {
  if (size() == 1) return m_[0]->run_bound(id);
  auto it = bound_.find(id);
  KV_CHECK(it != bound_.end(), "unknown bound call " << id);
  fence_in();
  for (int i = 0; i < size(); ++i) {
    DeviceGuard g(m_[i]->device());
    m_[i]->preallocate(*m_[i]->bound_.at(it->second[i]));
  }
  for (int i = 0; i < size(); ++i) {
    DeviceGuard g(m_[i]->device());
    m_[i]->run_bound(it->second[i]);
  }
  fence_out();
}

void Group::unbind(int id) {
  if (size() == 1) return m_[0]->unbind(id);
  auto it = bound_.find(id);
  if (it == bound_.end()) return;
  for (int i = 0; i < size(); ++i) m_[i]->unbind(it->second[i]);
  bound_.erase(it);
}

void Group::set_optimizer(const b200kv_optimizer& o) {
  for (auto& m : m_) {
    DeviceGuard g(m->device());
    m->set_optimizer(o);
  }
}

void Group::set_hyper(int num, const int* keys, const float* lrs, const float* wds, const int* steps) {
  for (auto& m : m_) m->set_hyper(num, keys, lrs, wds, steps);
}

float Group::last_global_norm() { return m_[0]->last_global_norm(); }

DLManagedTensor* Group::view(int member, int key, int which) {
  KV_CHECK(member >= 0 && member < size(), "view: member " << member << " out of range");
  DeviceGuard g(m_[member]->device());
  return m_[member]->view(key, which);
}

void Group::get_state(int key, int which, DLManagedTensor* out, int* present) {
  if (size() == 1) return m_[0]->get_state(key, which, out, present);
  KV_CHECK(out != nullptr, "state out: null DLManagedTensor");
  const int dev = out->dl_tensor.device.device_id;
  const TensorView v = view_of(out, dev, "state out");
  KV_CHECK(v.dtype == kF32, "state out must be float32");
  sync_callers();
  bool any = false;
  for (auto& m : m_) {
    DeviceGuard g(m->device());
    any = m->copy_state_shard(key, which, static_cast<float*>(v.data), dev, v.numel) || any;
  }
  *present = any ? 1 : 0;
  sync_members();
}

void Group::set_state(int key, int which, DLManagedTensor* in) {
  if (size() == 1) return m_[0]->set_state(key, which, in);
  KV_CHECK(in != nullptr, "state in: null DLManagedTensor");
  const int dev = in->dl_tensor.device.device_id;
  const TensorView v = view_of(in, dev, "state in");
  KV_CHECK(v.dtype == kF32, "state must be float32");
  sync_callers();
  for (auto& m : m_) {
    DeviceGuard g(m->device());
    m->load_state_shard(key, which, static_cast<const float*>(v.data), dev, v.numel);
  }
  if (which == 2) {
    sync_members();
    for (auto& m : m_) {
      DeviceGuard g(m->device());
      m->spread_fp16_shard(key, raw_);
    }
  }
  sync_members();
}

void Group::set_profiling(bool on) {
  for (auto& m : m_) m->set_profiling(on);
}

void Group::profile_read_phases(double ms[8], int64_t count[8]) {
  double t[8];
  int64_t c[8];
  for (int i = size() - 1; i >= 0; --i) {   // member 0's numbers are reported; the others are drained
    DeviceGuard g(m_[i]->device());
    m_[i]->profile_read_phases(i == 0 ? ms : t, i == 0 ? count : c);
  }
}

int64_t Group::launches() const {
  int64_t n = 0;
  for (auto& m : m_) n += m->launches();
  return n;
}

// ---- row sparse ---------------------------------------------------------------------------------------
void Group::init_rsp(int key, DLManagedTensor* dense_value) {
  if (size() == 1) return m_[0]->init_rsp(key, dense_value);
  sync_callers();
  for (auto& m : m_) {
    DeviceGuard g(m->device());
    m->init_rsp(key, dense_value);
  }
}

void Group::push_rsp(int key, int ncopies, DLManagedTensor* const* idx, DLManagedTensor* const* val) {
  if (size() == 1) return m_[0]->push_rsp(key, ncopies, idx, val);
  // CommDevice::ReduceRowSparse (comm.h:478-501) gathers every copy on one device; here every member
  // reads ALL copies in list order (= the reference's device order), its siblings' through peer
  // memory.  With an optimizer each member reduces and updates only the rows it owns (dealt cyclically (row id % members)
  // table and state); without one every member stores the identical merged value
  fence_in(true);
  for (auto& m : m_) {
    DeviceGuard g(m->device());
    m->push_rsp(key, ncopies, idx, val);
  }
  if (m_[0]->rsp_is_sharded(key)) {
    // with an optimizer the table and its state are dealt cyclically (row id % members) over the members (each one
    // reduced and updated only the rows it owns); pulls gather from every member's shard
    const float* tabs[kMaxCopies] = {nullptr};
    for (int i = 0; i < size(); ++i) tabs[i] = m_[i]->rsp_table_ptr(key);
    for (auto& m : m_) m->rsp_set_peer_tables(key, tabs);
  }
  fence_out();
}

int64_t Group::rsp_merged_size(int key) { return m_[0]->rsp_merged_size(key); }

void Group::rsp_merged_copy(int key, DLManagedTensor* out_idx, DLManagedTensor* out_val) {
  KV_CHECK(out_idx != nullptr, "merged indices out: null DLManagedTensor");
  const int m = member_of_device(out_idx->dl_tensor.device.device_id, 0);
  DeviceGuard g(m_[m]->device());
  m_[m]->rsp_merged_copy(key, out_idx, out_val);
  CUDA_OK(cudaStreamSynchronize(m_[m]->stream()));
}

int64_t Group::row_sparse_pull(int key, DLManagedTensor* row_ids, DLManagedTensor* out_idx, DLManagedTensor* out_val,
                               int64_t* nnr_dev) {
  if (size() == 1) return m_[0]->row_sparse_pull(key, row_ids, out_idx, out_val, nnr_dev);
  KV_CHECK(out_val != nullptr, "row_sparse_pull values out: null DLManagedTensor");
  const int m = member_of_device(out_val->dl_tensor.device.device_id, 0);
  fence_in(true);
  DeviceGuard g(m_[m]->device());
  const int64_t n = m_[m]->row_sparse_pull(key, row_ids, out_idx, out_val, nnr_dev);   // sync form blocks for the count
  // the host only waited for the count (if at all): the caller's stream picks up after the row gather
  CUDA_OK(cudaEventRecord(ev_out_[m], own_[m]));
  CUDA_OK(cudaStreamWaitEvent(caller_[m], ev_out_[m], 0));
  return n;
}

}  // namespace b200kv
