// extern "C" boundary: every entry point returns 0 / -1 and records the failure text for
// b200kv_last_error(), like MXNet's API_BEGIN/API_END (include/mxnet/c_api_error.h:36-55).
#include <cmath>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <vector>

#include "group.h"
#include "nccl_dl.h"
#include "store.h"

namespace b200kv {
const char* get_last_error();
}

using namespace b200kv;

#define API_BEGIN() try {
#define API_END()                                   \
  }                                                 \
  catch (const std::exception& e) {                 \
    b200kv::set_last_error(e.what());               \
    return -1;                                      \
  }                                                 \
  catch (...) {                                     \
    b200kv::set_last_error("unknown C++ exception");\
    return -1;                                      \
  }                                                 \
  return 0;

static KVHandle* H(b200kv_handle h) {
  KV_CHECK(h != nullptr, "null store handle");
  return static_cast<KVHandle*>(h);
}
static Store* S(b200kv_handle h) {
  KVHandle* b = H(h);
  KV_CHECK(!b->is_group, "this entry point needs a single-GPU store handle, got a group handle");
  return static_cast<Store*>(b);
}
static Group* G(b200kv_handle h) {
  KVHandle* b = H(h);
  KV_CHECK(b->is_group, "this entry point needs a handle made by b200kv_create_group");
  return static_cast<Group*>(b);
}
// a group works on several devices in turn: the caller's current device is put back afterwards
struct KeepDevice {
  int d = -1;
  explicit KeepDevice(bool on) { if (on) cudaGetDevice(&d); }
  ~KeepDevice() { if (d >= 0) cudaSetDevice(d); }
};
// the same call on whichever object the handle is
#define KV(h, call)                                      \
  do {                                                   \
    KeepDevice keep_(H(h)->is_group);                    \
    if (H(h)->is_group) static_cast<Group*>(H(h))->call; \
    else static_cast<Store*>(H(h))->call;                \
  } while (0)
#define KVR(h, lhs, call)                                      \
  do {                                                         \
    KeepDevice keep_(H(h)->is_group);                          \
    if (H(h)->is_group) lhs = static_cast<Group*>(H(h))->call; \
    else lhs = static_cast<Store*>(H(h))->call;                \
  } while (0)

extern "C" {

const char* b200kv_last_error(void) { return b200kv::get_last_error(); }

int b200kv_version(int* out) {
  API_BEGIN();
  *out = 100;  // 0.1.0
  API_END();
}

int b200kv_create(int device_id, int rank, int world_size, b200kv_handle* out) {
  API_BEGIN();
  *out = static_cast<KVHandle*>(new Store(device_id, rank, world_size));
  API_END();
}

int b200kv_create_group(int ndev, const int* dev_ids, b200kv_handle* out) {
  API_BEGIN();
  KV_CHECK(dev_ids != nullptr && out != nullptr, "null argument");
  KeepDevice keep_(true);
  *out = static_cast<KVHandle*>(new Group(ndev, dev_ids));
  API_END();
}

int b200kv_group_set_streams(b200kv_handle h, int n, void* const* streams) {
  API_BEGIN();
  G(h)->set_streams(n, streams);
  API_END();
}

int b200kv_group_view(b200kv_handle h, int member, int key, int which, DLManagedTensor** out) {
  API_BEGIN();
  KeepDevice keep_(true);
  *out = G(h)->view(member, key, which);
  API_END();
}

int b200kv_group_size(b200kv_handle h, int* out) {
  API_BEGIN();
  *out = H(h)->is_group ? G(h)->size() : 1;
  API_END();
}

int b200kv_free(b200kv_handle h) {
  API_BEGIN();
  delete H(h);
  API_END();
}

int b200kv_set_stream(b200kv_handle h, void* cuda_stream) {
  API_BEGIN();
  S(h)->set_stream(static_cast<cudaStream_t>(cuda_stream));
  API_END();
}

int b200kv_set_symmetric_allocator(b200kv_handle h, b200kv_symm_alloc_fn fn, void* ctx) {
  API_BEGIN();
  S(h)->set_symmetric_allocator(fn, ctx);
  API_END();
}

int b200kv_wait(b200kv_handle h) {
  API_BEGIN();
  KV(h, wait());
  API_END();
}

int b200kv_comm_unique_id(char out_id[128]) {
  API_BEGIN();
  ncclUniqueId id;
  NCCL_OK(NcclApi::get().GetUniqueId(&id));
  std::memcpy(out_id, &id, 128);
  API_END();
}

int b200kv_comm_init(b200kv_handle h, const char id[128]) {
  API_BEGIN();
  S(h)->comm_init(id);
  API_END();
}

int b200kv_lookup_keys(b200kv_handle h, int num, const char* const* str_keys, int create,
                       int* out_keys) {
  API_BEGIN();
  KV(h, lookup_keys(num, str_keys, create != 0, out_keys));
  API_END();
}

int b200kv_init(b200kv_handle h, int num, const int* keys, DLManagedTensor* const* vals) {
  API_BEGIN();
  KV(h, init(num, keys, vals));
  API_END();
}

int b200kv_push(b200kv_handle h, int num, const int* keys, DLManagedTensor* const* vals, int) {
  API_BEGIN();
  KV(h, dense(Store::kPush, num, keys, vals, 0, nullptr, nullptr));
  API_END();
}

int b200kv_pull(b200kv_handle h, int num, const int* keys, DLManagedTensor* const* outs, int) {
  API_BEGIN();
  KV(h, dense(Store::kPull, 0, nullptr, nullptr, num, keys, outs));
  API_END();
}

int b200kv_pushpull(b200kv_handle h, int vnum, const int* vkeys, DLManagedTensor* const* vals,
                    int onum, const int* okeys, DLManagedTensor* const* outs, int) {
  API_BEGIN();
  KV(h, dense(Store::kPushPull, vnum, vkeys, vals, onum, okeys, outs));
  API_END();
}

int b200kv_bind_pushpull(b200kv_handle h, int vnum, const int* vkeys, DLManagedTensor* const* vals,
                         int onum, const int* okeys, DLManagedTensor* const* outs, int* out_id) {
  API_BEGIN();
  KVR(h, *out_id, bind(Store::kPushPull, vnum, vkeys, vals, onum, okeys, outs));
  API_END();
}

int b200kv_run_bound(b200kv_handle h, int id) {
  API_BEGIN();
  KV(h, run_bound(id));
  API_END();
}

int b200kv_unbind(b200kv_handle h, int id) {
  API_BEGIN();
  KV(h, unbind(id));
  API_END();
}

int b200kv_set_profiling(b200kv_handle h, int enable) {
  API_BEGIN();
  KV(h, set_profiling(enable != 0));
  API_END();
}

int b200kv_profile_read(b200kv_handle h, double* kernel_ms, int64_t* count) {
  API_BEGIN();
  double m[8];
  int64_t c[8];
  KV(h, profile_read_phases(m, c));
  *kernel_ms = m[0];
  *count = c[0];
  API_END();
}

int b200kv_profile_read_phases(b200kv_handle h, double ms[8], int64_t count[8]) {
  API_BEGIN();
  KV(h, profile_read_phases(ms, count));
  API_END();
}

int b200kv_broadcast(b200kv_handle h, int vnum, const int* vkeys, DLManagedTensor* const* vals,
                     int onum, const int* okeys, DLManagedTensor* const* outs, int) {
  API_BEGIN();
  // BroadcastImpl = InitImpl; PullImpl (kvstore_local.h:327-334)
  KV(h, init(vnum, vkeys, vals));
  KV(h, dense(Store::kPull, 0, nullptr, nullptr, onum, okeys, outs));
  API_END();
}

int b200kv_set_optimizer(b200kv_handle h, const b200kv_optimizer* opt) {
  API_BEGIN();
  KV_CHECK(opt != nullptr, "null optimizer description");
  KV(h, set_optimizer(*opt));
  API_END();
}

int b200kv_set_hyper(b200kv_handle h, int num, const int* keys, const float* lrs, const float* wds,
                     const int* steps) {
  API_BEGIN();
  KV(h, set_hyper(num, keys, lrs, wds, steps));
  API_END();
}

int b200kv_last_global_norm(b200kv_handle h, float* out) {
  API_BEGIN();
  KVR(h, *out, last_global_norm());
  API_END();
}

int b200kv_view(b200kv_handle h, int key, int which, DLManagedTensor** out) {
  API_BEGIN();
  KeepDevice keep_(H(h)->is_group);
  *out = H(h)->is_group ? G(h)->view(0, key, which) : S(h)->view(key, which);
  API_END();
}

int b200kv_get_state(b200kv_handle h, int key, int which, DLManagedTensor* out, int* present) {
  API_BEGIN();
  KV(h, get_state(key, which, out, present));
  API_END();
}

int b200kv_set_state(b200kv_handle h, int key, int which, DLManagedTensor* in) {
  API_BEGIN();
  KV(h, set_state(key, which, in));
  API_END();
}

int b200kv_init_rsp(b200kv_handle h, int key, DLManagedTensor* dense_value) {
  API_BEGIN();
  KV(h, init_rsp(key, dense_value));
  API_END();
}

int b200kv_push_rsp(b200kv_handle h, int key, int ncopies, DLManagedTensor* const* indices,
                    DLManagedTensor* const* values, int) {
  API_BEGIN();
  KV(h, push_rsp(key, ncopies, indices, values));
  API_END();
}

int b200kv_rsp_merged_size(b200kv_handle h, int key, int64_t* nnr) {
  API_BEGIN();
  KVR(h, *nnr, rsp_merged_size(key));
  API_END();
}

int b200kv_rsp_merged_copy(b200kv_handle h, int key, DLManagedTensor* out_indices,
                           DLManagedTensor* out_values) {
  API_BEGIN();
  KV(h, rsp_merged_copy(key, out_indices, out_values));
  API_END();
}

int b200kv_row_sparse_pull(b200kv_handle h, int key, DLManagedTensor* row_ids,
                           DLManagedTensor* out_indices, DLManagedTensor* out_values,
                           int64_t* out_nnr, int) {
  API_BEGIN();
  KVR(h, *out_nnr, row_sparse_pull(key, row_ids, out_indices, out_values));
  API_END();
}

int b200kv_row_sparse_pull_async(b200kv_handle h, int key, DLManagedTensor* row_ids,
                                 DLManagedTensor* out_indices, DLManagedTensor* out_values,
                                 int64_t* out_nnr_dev, int) {
  API_BEGIN();
  KV_CHECK(out_nnr_dev != nullptr, "out_nnr_dev must be a device pointer to one int64");
  int64_t r = 0;
  KVR(h, r, row_sparse_pull(key, row_ids, out_indices, out_values, out_nnr_dev));
  (void)r;
  API_END();
}

int b200kv_get_type(b200kv_handle h, const char** out) {
  API_BEGIN();
  *out = "b200";
  (void)H(h);
  API_END();
}

int b200kv_get_rank(b200kv_handle h, int* out) {
  API_BEGIN();
  *out = H(h)->is_group ? 0 : S(h)->rank();
  API_END();
}

int b200kv_get_group_size(b200kv_handle h, int* out) {
  API_BEGIN();
  *out = H(h)->is_group ? 1 : S(h)->world();   // a group is ONE worker driving several devices
  API_END();
}

int b200kv_barrier(b200kv_handle h) {
  API_BEGIN();
  KV(h, barrier());
  API_END();
}

int b200kv_last_exchange(b200kv_handle h, const char** out) {
  API_BEGIN();
  KVR(h, *out, last_exchange());
  API_END();
}

int b200kv_launch_count(b200kv_handle h, int64_t* out) {
  API_BEGIN();
  KVR(h, *out, launches());
  API_END();
}

}  // extern "C"

// =================================================================================================
// Operator entry points: the fused multi-tensor ops on caller-owned arrays (no store).
// =================================================================================================
namespace {

// Workspace of the operator calls, one per (device, stream): the tensor table ring, chunk plans cached
// by size list, and the small reduction scratch.  Everything a launch reads is either immutable once
// built (a cached plan), event-guarded (the table ring) or only rewritten by later work on the SAME
// stream (partials, sums, LAMB's scratch), so calls on different streams never share mutable state
// and a call with a new size list never rewrites a chunk list an earlier launch may still be reading.
struct OpWorkspace {
  TableRing ring;
  std::map<std::vector<int64_t>, std::unique_ptr<ChunkPlan>> plans;
  ChunkPlan* plan = nullptr;   // the plan of the current call
  DevBuf partials, sums;
  DevBuf xbuf;   // fp32 scratch for LAMB's update direction (one slot per tensor, 64-element aligned)
  std::mutex mu;
};

OpWorkspace& workspace(int device, void* stream) {
  static std::mutex mu;
  static std::map<std::pair<int, void*>, std::unique_ptr<OpWorkspace>> ws;
  std::lock_guard<std::mutex> lk(mu);
  KV_CHECK(device >= 0 && device < 64, "device ordinal out of range");
  auto& slot = ws[std::make_pair(device, stream)];
  if (!slot) slot.reset(new OpWorkspace());
  return *slot;
}

int current_device() {
  int d = 0;
  CUDA_OK(cudaGetDevice(&d));
  return d;
}

void prepare_plan(OpWorkspace& w, int num, const int64_t* sizes, cudaStream_t stream) {
  std::vector<int64_t> sz(sizes, sizes + num);
  auto it = w.plans.find(sz);
  if (it == w.plans.end()) {
    if (w.plans.size() >= 64) {   // bound the cache; nothing may still read the plans being dropped
      CUDA_OK(cudaStreamSynchronize(stream));
      w.plans.clear();
    }
    std::unique_ptr<ChunkPlan> p(new ChunkPlan());
    std::vector<int64_t> b(num, 0);
    p->build(b, sz);
    it = w.plans.emplace(std::move(sz), std::move(p)).first;
  }
  w.plan = it->second.get();
  // grow-only; DevBuf::reserve releases with cudaFree, which waits for the device to drain first
  w.partials.reserve(sizeof(float) * 2 * static_cast<size_t>(std::max(w.plan->nchunks, 1)));
  w.sums.reserve(sizeof(float) * (3 * static_cast<size_t>(num) + 4));
}

struct Table {
  uint64_t* p;
  int nt, ng, no;
  void G(int k, int t, const void* x) { p[static_cast<size_t>(k) * nt + t] = reinterpret_cast<uint64_t>(x); }
  void Sx(int sec, int t, const void* x) { p[static_cast<size_t>(ng + no + sec) * nt + t] = reinterpret_cast<uint64_t>(x); }
  void F2(int sec, int t, float a, float b) {
    float v[2] = {a, b};
    std::memcpy(&p[static_cast<size_t>(ng + no + sec) * nt + t], v, 8);
  }
};

// Fills the table for an in-place update of caller arrays.  For dtype fp16 (multi-precision) the
// W section is the fp32 master and the fp16 weight is the single out copy.
MtaCall make_call(OpWorkspace& w, int num, int dtype, void* const* weights, const void* const* grads,
                  float* const* s0, float* const* s1, float* const* w32, const float* lrs,
                  const float* wds, const float* aux0, const float* aux1, cudaStream_t stream,
                  int device, const int64_t* x_offsets = nullptr) {
  KV_CHECK(dtype == kF32 || dtype == kF16, "dtype must be 0 (float32) or 2 (float16 + fp32 master)");
  const bool mp = dtype == kF16;
  KV_CHECK(!mp || w32 != nullptr, "float16 weights need weights32 master copies");
  const int ng = grads ? 1 : 0, no = mp ? 1 : 0;
  const int ngl = ng ? ng : 1;   // the table always carries one gradient section
  Table t{w.ring.begin(mta_blob_words(num, ngl, no) * 8), num, ngl, no};
  std::memset(t.p, 0, mta_blob_words(num, ngl, no) * 8);
  for (int i = 0; i < num; ++i) {
    if (grads) t.G(0, i, grads[i]);
    if (mp) { t.G(ngl, i, weights[i]); t.Sx(kSecW, i, w32[i]); }   // out section follows the grads
    else t.Sx(kSecW, i, weights[i]);
    if (s0) t.Sx(kSecS0, i, s0[i]);
    if (s1) t.Sx(kSecS1, i, s1[i]);
    if (x_offsets) t.Sx(kSecX, i, w.xbuf.as<float>() + x_offsets[i]);
    t.F2(kSecLrWd, i, lrs ? lrs[i] : 0.f, wds ? wds[i] : 0.f);
    t.F2(kSecAux, i, aux0 ? aux0[i] : 0.f, aux1 ? aux1[i] : 0.f);
  }
  MtaCall c;
  c.blob = w.ring.commit(stream);
  c.plan = w.plan;
  c.nt = num; c.ngrad = ngl; c.nout = no; c.dtype = dtype; c.stream = stream; c.device = device;
  return c;
}

}  // namespace

extern "C" {

int b200kv_multi_sgd_update(int num, int dtype, void* const* weights, const void* const* grads,
                            float* const* moms, float* const* weights32, const int64_t* sizes,
                            const float* lrs, const float* wds, float momentum, float rescale_grad,
                            float clip_gradient, void* stream) {
  API_BEGIN();
  if (num <= 0) return 0;
  const int dev = current_device();
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  OpWorkspace& w = workspace(dev, stream);
  std::lock_guard<std::mutex> lk(w.mu);
  prepare_plan(w, num, sizes, st);
  MtaCall c = make_call(w, num, dtype, weights, grads, moms, nullptr, weights32, lrs, wds, nullptr,
                        nullptr, st, dev);
  OptScalars s;
  s.momentum = momentum; s.rescale = rescale_grad; s.clip = clip_gradient;
  launch_sgd(c, moms != nullptr, s, nullptr);
  w.ring.end(st);
  API_END();
}

int b200kv_sgd_update(int dtype, void* weight, const void* grad, float* weight32, int64_t size, float lr,
                      float wd, float rescale_grad, float clip_gradient, void* stream) {
  API_BEGIN();
  if (size <= 0) return 0;
  const int dev = current_device();
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  OpWorkspace& w = workspace(dev, stream);
  std::lock_guard<std::mutex> lk(w.mu);
  prepare_plan(w, 1, &size, st);
  void* ws[1] = {weight};
  const void* gs[1] = {grad};
  float* w32[1] = {weight32};
  MtaCall c = make_call(w, 1, dtype, ws, gs, nullptr, nullptr, dtype == kF16 ? w32 : nullptr, &lr, &wd, nullptr,
                        nullptr, st, dev);
  OptScalars s;
  s.rescale = rescale_grad; s.clip = clip_gradient;
  launch_sgd_plain(c, s);
  w.ring.end(st);
  API_END();
}

int b200kv_lamb_update_phase1(float* g_out, const float* weight, const float* grad, float* mean, float* var,
                              int64_t size, float beta1, float beta2, float epsilon, int t, int bias_correction,
                              float wd, float rescale_grad, float clip_gradient, void* stream) {
  API_BEGIN();
  if (size <= 0) return 0;
  KV_CHECK(g_out && weight && grad && mean && var, "null array");
  const int dev = current_device();
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  OpWorkspace& w = workspace(dev, stream);
  std::lock_guard<std::mutex> lk(w.mu);
  prepare_plan(w, 1, &size, st);
  Table tb{w.ring.begin(mta_blob_words(1, 1, 0) * 8), 1, 1, 0};
  std::memset(tb.p, 0, mta_blob_words(1, 1, 0) * 8);
  tb.G(0, 0, grad);
  tb.Sx(kSecW, 0, weight);
  tb.Sx(kSecS0, 0, mean);
  tb.Sx(kSecS1, 0, var);
  tb.Sx(kSecX, 0, g_out);
  tb.F2(kSecLrWd, 0, 0.f, wd);
  // DType beta1_t = std::pow(param.beta1, param.t) (optimizer_op-inl.h:1661-1662): double pow, rounded to float
  tb.F2(kSecAux, 0, static_cast<float>(std::pow(static_cast<double>(beta1), static_cast<double>(t))),
        static_cast<float>(std::pow(static_cast<double>(beta2), static_cast<double>(t))));
  MtaCall c;
  c.blob = w.ring.commit(st);
  c.plan = w.plan;
  c.nt = 1; c.ngrad = 1; c.nout = 0; c.dtype = kF32; c.stream = st; c.device = dev;
  OptScalars s;
  s.beta1 = beta1; s.beta2 = beta2; s.eps = epsilon; s.rescale = rescale_grad; s.clip = clip_gradient;
  s.bias_correction = bias_correction;
  launch_lamb_phase1(c, s);
  w.ring.end(st);
  API_END();
}

int b200kv_lamb_update_phase2(float* weight, const float* g, const float* r1, const float* r2, int64_t size,
                              float lr, float lower_bound, float upper_bound, void* stream) {
  API_BEGIN();
  if (size <= 0) return 0;
  KV_CHECK(weight && g && r1 && r2, "null array");
  const int dev = current_device();
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  OpWorkspace& w = workspace(dev, stream);
  std::lock_guard<std::mutex> lk(w.mu);
  prepare_plan(w, 1, &size, st);
  Table tb{w.ring.begin(mta_blob_words(1, 1, 0) * 8), 1, 1, 0};
  std::memset(tb.p, 0, mta_blob_words(1, 1, 0) * 8);
  tb.Sx(kSecW, 0, weight);
  tb.Sx(kSecX, 0, g);
  tb.F2(kSecLrWd, 0, lr, 0.f);
  MtaCall c;
  c.blob = w.ring.commit(st);
  c.plan = w.plan;
  c.nt = 1; c.ngrad = 1; c.nout = 0; c.dtype = kF32; c.stream = st; c.device = dev;
  launch_lamb_phase2(c, r1, r2, lower_bound, upper_bound);
  w.ring.end(st);
  API_END();
}

int b200kv_multi_adam_update(int num, int dtype, void* const* weights, const void* const* grads,
                             float* const* means, float* const* vars, float* const* weights32,
                             const int64_t* sizes, const float* lrs, const float* wds, float beta1,
                             float beta2, float epsilon, float rescale_grad, float clip_gradient,
                             void* stream) {
  API_BEGIN();
  if (num <= 0) return 0;
  const int dev = current_device();
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  OpWorkspace& w = workspace(dev, stream);
  std::lock_guard<std::mutex> lk(w.mu);
  prepare_plan(w, num, sizes, st);
  MtaCall c = make_call(w, num, dtype, weights, grads, means, vars, weights32, lrs, wds, nullptr,
                        nullptr, st, dev);
  OptScalars s;
  s.beta1 = beta1; s.beta2 = beta2; s.eps = epsilon; s.rescale = rescale_grad; s.clip = clip_gradient;
  launch_adam(c, s, nullptr);
  w.ring.end(st);
  API_END();
}

int b200kv_multi_lamb_update(int num, int dtype, void* const* weights, const void* const* grads,
                             float* const* means, float* const* vars, float* const* weights32,
                             const int64_t* sizes, const float* lrs, const float* wds,
                             const int* step_counts, float beta1, float beta2, float epsilon,
                             float rescale_grad, float clip_gradient, float lower_bound,
                             float upper_bound, int bias_correction, void* stream) {
  API_BEGIN();
  if (num <= 0) return 0;
  const int dev = current_device();
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  OpWorkspace& w = workspace(dev, stream);
  std::lock_guard<std::mutex> lk(w.mu);
  prepare_plan(w, num, sizes, st);
  std::vector<float> c1(num), c2(num);
  for (int i = 0; i < num; ++i) {
    const float tf = static_cast<float>(step_counts[i]);
    c1[i] = 1.0f - powf(beta1, tf);
    c2[i] = 1.0f - powf(beta2, tf);
  }
  // pass 1 leaves the update direction u in a scratch slot per tensor; pass 2 streams w, u -> w
  std::vector<int64_t> xoff(num);
  int64_t xtotal = 0;
  for (int i = 0; i < num; ++i) {
    xoff[i] = xtotal;
    xtotal += (sizes[i] + 63) / 64 * 64;
  }
  w.xbuf.reserve(sizeof(float) * static_cast<size_t>(std::max<int64_t>(xtotal, 64)));
  MtaCall c = make_call(w, num, dtype, weights, grads, means, vars, weights32, lrs, wds, c1.data(),
                        c2.data(), st, dev, xoff.data());
  OptScalars s;
  s.beta1 = beta1; s.beta2 = beta2; s.eps = epsilon; s.rescale = rescale_grad; s.clip = clip_gradient;
  s.lower = lower_bound; s.upper = upper_bound; s.bias_correction = bias_correction;
  float* partials = w.partials.as<float>();
  float* sums = w.sums.as<float>();
  float* ratio = sums + 2 * static_cast<size_t>(num);
  launch_lamb1(c, s, nullptr, partials, true);
  launch_reduce_partials(partials, *w.plan, 2, sums, st);
  launch_lamb_ratio(sums, num, lower_bound, upper_bound, ratio, st);
  launch_lamb2u(c, ratio);
  w.ring.end(st);
  API_END();
}

int b200kv_multi_lars(int num, const float* lrs, const float* weights_sum_sq, const float* grads_sum_sq,
                      const float* wds, float eta, float eps, float rescale_grad, float* out_lrs,
                      void* stream) {
  API_BEGIN();
  launch_multi_lars(num, lrs, weights_sum_sq, grads_sum_sq, wds, eta, eps, rescale_grad, out_lrs,
                    static_cast<cudaStream_t>(stream));
  API_END();
}

int b200kv_preloaded_multi_sgd_update(int num, int dtype, void* const* weights, const void* const* grads,
                                      float* const* moms, float* const* weights32, const int64_t* sizes,
                                      const float* lrs_dev, const float* wds_dev, float momentum,
                                      float rescale_grad, float clip_gradient, void* stream) {
  API_BEGIN();
  if (num <= 0) return 0;
  KV_CHECK(lrs_dev != nullptr && wds_dev != nullptr, "lrs / wds must be device arrays of `num` floats");
  const int dev = current_device();
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  OpWorkspace& w = workspace(dev, stream);
  std::lock_guard<std::mutex> lk(w.mu);
  prepare_plan(w, num, sizes, st);
  MtaCall c = make_call(w, num, dtype, weights, grads, moms, nullptr, weights32, nullptr, nullptr, nullptr,
                        nullptr, st, dev);
  launch_fill_lrwd(c, lrs_dev, wds_dev);
  w.ring.invalidate();   // the uploaded table was patched on the device
  OptScalars s;
  s.momentum = momentum; s.rescale = rescale_grad; s.clip = clip_gradient;
  launch_sgd(c, moms != nullptr, s, nullptr);
  w.ring.end(st);
  API_END();
}

int b200kv_multi_adamw_update(int num, int dtype, void* const* weights, const void* const* grads,
                              float* const* means, float* const* vars, float* const* weights32,
                              const int64_t* sizes, const float* lrs, const float* wds, const float* etas,
                              float beta1, float beta2, float epsilon, const float* rescale_grad_dev,
                              float clip_gradient, void* stream) {
  API_BEGIN();
  if (num <= 0) return 0;
  KV_CHECK(rescale_grad_dev != nullptr, "rescale_grad must be a device pointer to one float");
  const int dev = current_device();
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  OpWorkspace& w = workspace(dev, stream);
  std::lock_guard<std::mutex> lk(w.mu);
  prepare_plan(w, num, sizes, st);
  MtaCall c = make_call(w, num, dtype, weights, grads, means, vars, weights32, lrs, wds, etas, nullptr, st, dev);
  OptScalars s;
  s.beta1 = beta1; s.beta2 = beta2; s.eps = epsilon; s.clip = clip_gradient;
  launch_adamw(c, s, rescale_grad_dev);
  w.ring.end(st);
  API_END();
}

int b200kv_multi_all_finite(int num, int dtype, const void* const* arrays, const int64_t* sizes,
                            int init_output, float* out, void* stream) {
  API_BEGIN();
  if (num <= 0) return 0;
  const int dev = current_device();
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  OpWorkspace& w = workspace(dev, stream);
  std::lock_guard<std::mutex> lk(w.mu);
  prepare_plan(w, num, sizes, st);
  KV_CHECK(dtype == kF32 || dtype == kF16, "dtype must be 0 (float32) or 2 (float16)");
  Table t{w.ring.begin(mta_blob_words(num, 1, 0) * 8), num, 1, 0};
  std::memset(t.p, 0, mta_blob_words(num, 1, 0) * 8);
  for (int i = 0; i < num; ++i) t.G(0, i, arrays[i]);
  MtaCall c;
  c.blob = w.ring.commit(st);
  c.plan = w.plan;
  c.nt = num; c.ngrad = 1; c.nout = 0; c.dtype = dtype; c.stream = st; c.device = dev;
  launch_nonfinite(c, w.partials.as<float>());
  launch_reduce_partials(w.partials.as<float>(), *w.plan, 1, w.sums.as<float>(), st);
  launch_all_finite_final(w.sums.as<float>(), num, init_output, out, st);
  w.ring.end(st);
  API_END();
}

int b200kv_reset_arrays(int num, int dtype, void* const* arrays, const int64_t* sizes, void* stream) {
  API_BEGIN();
  if (num <= 0) return 0;
  const int dev = current_device();
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  OpWorkspace& w = workspace(dev, stream);
  std::lock_guard<std::mutex> lk(w.mu);
  prepare_plan(w, num, sizes, st);
  KV_CHECK(dtype == kF32 || dtype == kF16, "dtype must be 0 (float32) or 2 (float16)");
  Table t{w.ring.begin(mta_blob_words(num, 1, 1) * 8), num, 1, 1};
  std::memset(t.p, 0, mta_blob_words(num, 1, 1) * 8);
  for (int i = 0; i < num; ++i) t.G(1, i, arrays[i]);   // the out section
  MtaCall c;
  c.blob = w.ring.commit(st);
  c.plan = w.plan;
  c.nt = num; c.ngrad = 1; c.nout = 1; c.dtype = dtype; c.stream = st; c.device = dev;
  launch_zero(c);
  w.ring.end(st);
  API_END();
}

int b200kv_clip_global_norm(int num, int dtype, void* const* arrays, const int64_t* sizes,
                            float max_norm, float* out_scale_norm, void* stream) {
  API_BEGIN();
  if (num <= 0) return 0;
  const int dev = current_device();
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  OpWorkspace& w = workspace(dev, stream);
  std::lock_guard<std::mutex> lk(w.mu);
  prepare_plan(w, num, sizes, st);
  KV_CHECK(dtype == kF32 || dtype == kF16, "dtype must be 0 (float32) or 2 (float16)");
  KV_CHECK(out_scale_norm != nullptr, "out_scale_norm must point to two device floats");
  Table t{w.ring.begin(mta_blob_words(num, 1, 1) * 8), num, 1, 1};
  std::memset(t.p, 0, mta_blob_words(num, 1, 1) * 8);
  for (int i = 0; i < num; ++i) {
    t.G(0, i, arrays[i]);
    t.G(1, i, arrays[i]);   // the single out section follows the gradient section: scale in place
  }
  MtaCall c;
  c.blob = w.ring.commit(st);
  c.plan = w.plan;
  c.nt = num; c.ngrad = 1; c.nout = 1; c.dtype = dtype; c.stream = st; c.device = dev;
  float* sums = w.sums.as<float>();
  launch_sumsq(c, w.partials.as<float>());
  launch_reduce_partials(w.partials.as<float>(), *w.plan, 1, sums, st);
  launch_global_norm(sums, num, max_norm, out_scale_norm, st);
  launch_scale(c, out_scale_norm);
  w.ring.end(st);
  API_END();
}

int b200kv_multi_sum_sq(int num, int dtype, const void* const* arrays, const int64_t* sizes,
                        float* out, void* stream) {
  API_BEGIN();
  if (num <= 0) return 0;
  const int dev = current_device();
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  OpWorkspace& w = workspace(dev, stream);
  std::lock_guard<std::mutex> lk(w.mu);
  prepare_plan(w, num, sizes, st);
  KV_CHECK(dtype == kF32 || dtype == kF16, "dtype must be 0 (float32) or 2 (float16)");
  Table t{w.ring.begin(mta_blob_words(num, 1, 0) * 8), num, 1, 0};
  std::memset(t.p, 0, mta_blob_words(num, 1, 0) * 8);
  for (int i = 0; i < num; ++i) t.G(0, i, arrays[i]);
  MtaCall c;
  c.blob = w.ring.commit(st);
  c.plan = w.plan;
  c.nt = num; c.ngrad = 1; c.nout = 0; c.dtype = dtype; c.stream = st; c.device = dev;
  launch_sumsq(c, w.partials.as<float>());
  launch_reduce_partials(w.partials.as<float>(), *w.plan, 1, out, st);
  w.ring.end(st);
  API_END();
}

}  // extern "C"
