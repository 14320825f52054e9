// The store object behind the C ABI: KVStoreLocal's bookkeeping (src/kvstore/kvstore_local.h:94-343)
// re-designed around flat per-dtype segments in HBM so that one push of a whole model is a handful
// of launches:  [pack -> reduce-scatter ->] fused update on the rank's shard [-> all-gather -> unpack].
#pragma once
#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "common.h"
#include "launch.h"

struct ncclComm;

namespace b200kv {

constexpr int64_t kAlignElems = 64;  // every key starts on a 64-element (>=128 B) boundary

struct KeyInfo {
  int key = 0;
  bool rsp = false;
  int dtype = kF32;
  int64_t size = 0;
  int seg = -1;        // segment index once sealed into a flat segment
  int64_t off = 0;     // element offset inside the segment
  std::unique_ptr<DevBuf> staged;  // value between init and the first push
  float lr = 0.01f, wd = 0.f;
  int step = 0;
};

// A flat, per-dtype slab of keys.  Element offsets are shared by every parallel array.
struct Segment {
  int dtype = kF32;
  int64_t total = 0;   // elements, padded to world * kAlignElems
  int64_t slice = 0;   // total / world: rank r owns [r*slice, (r+1)*slice)
  std::vector<int> keys;
  DevBuf wbuf;         // stored values, `dtype`, [total]
  DevBuf gbuf;         // gradient fusion bucket, `dtype`, [total] (world > 1)
  DevBuf master;       // fp32 master weights (multi-precision), [state_len]
  DevBuf s0, s1;       // fp32 optimizer state, [state_len]
  DevBuf xbuf;         // fp32 scratch, [state_len]: LAMB's update direction between its two passes
  int64_t state_begin = 0, state_len = 0;  // element range the state arrays cover
  bool has_s0 = false, has_s1 = false, has_master = false;
  bool w_stale = false;  // world == 1 && master: wbuf lags behind master (pull casts the master)
  // peer-memory (NVLink) path: every rank's gbuf / wbuf / flag block mapped into this process
  struct P2P {
    bool tried = false, ok = false;
    void* g[kMaxCopies] = {nullptr};
    void* w[kMaxCopies] = {nullptr};
    uint32_t* flags[kMaxCopies] = {nullptr};
    void* mc_g = nullptr;                  // NVLS multicast addresses of the bucket / weight segment
    void* mc_w = nullptr;
    bool symmetric = false;                // buffers came from the host's symmetric allocator
    void* stage[kMaxCopies] = {nullptr};   // every rank's staging area ([world][slice], push reduce-scatter)
    DevBuf local_stage;
    DevBuf local_flags;
    DevBuf dev_sync;     // P2PFlags resident on the device for the in-kernel barriers
    uint32_t epoch = 0;
    uint32_t xepoch = 0;   // round counter of the partial-sum exchange (xsum_kernel)
  } p2p;
};

struct DensePlan {
  std::vector<int> keys;   // sorted by flat offset
  int seg = -1, ngrad = 0, nout = 0;
  bool full_segment = false;
  ChunkPlan full, shard;
  DevBuf partials;         // 2 floats per chunk
  DevBuf sums;             // sums[2*nt] | ratio[nt] | gnorm[2]
  struct Piece { int64_t off, len; int root; };
  std::vector<Piece> pieces;   // (key range) x (owner slice), world > 1 partial path
};

struct RspKey;  // rsp.cu

class Group;

class Store : public KVHandle {
 public:
  // in_process: this store is one member of a Group (group.cu) — W ranks inside ONE process, the
  // reference's own usage model.  No NCCL communicator exists then: every exchange goes through
  // peer memory.  grid_share: number of members driving this same physical GPU.
  Store(int device, int rank, int world, bool in_process = false, int grid_share = 1);
  ~Store() override;

  void set_stream(cudaStream_t s) { stream_ = s; }
  void set_symmetric_allocator(b200kv_symm_alloc_fn fn, void* ctx) { symm_alloc_ = fn; symm_ctx_ = ctx; }
  void wait();
  void comm_init(const char id[128]);
  void lookup_keys(int num, const char* const* names, bool create, int* out);

  void init(int num, const int* keys, DLManagedTensor* const* vals);
  enum Mode { kPush = 1, kPull = 2, kPushPull = 3 };
  void dense(Mode mode, int vnum, const int* vkeys, DLManagedTensor* const* vals, int onum,
             const int* okeys, DLManagedTensor* const* outs);
  // a pushpull whose arguments were validated, grouped and planned once (same arrays every step)
  int bind(Mode mode, int vnum, const int* vkeys, DLManagedTensor* const* vals, int onum,
           const int* okeys, DLManagedTensor* const* outs);
  void run_bound(int id);
  void unbind(int id);
  // CUDA-event timing of the update kernels (bench.py's roofline leg)
  void set_profiling(bool on) { profiling_ = on; }
  void profile_read(double* ms, int64_t* count);
  // per phase: 0 update, 1 pack, 2 reduce(-scatter), 3 all-gather / broadcast, 4 unpack
  void profile_read_phases(double ms[8], int64_t count[8]);
  void set_optimizer(const b200kv_optimizer& opt);
  void set_hyper(int num, const int* keys, const float* lrs, const float* wds, const int* steps);
  float last_global_norm();
  // zero-copy views into the store's own buffers (which: 0 weight of `key`, 1 gradient bucket slot
  // of `key`, 2 the whole weight segment holding `key`, 3 the whole gradient bucket)
  DLManagedTensor* view(int key, int which);
  void get_state(int key, int which, DLManagedTensor* out, int* present);
  void set_state(int key, int which, DLManagedTensor* in);
  void barrier();

  // row sparse (rsp.cu)
  void init_rsp(int key, DLManagedTensor* dense_value);
  void push_rsp(int key, int ncopies, DLManagedTensor* const* idx, DLManagedTensor* const* val);
  int64_t rsp_merged_size(int key);
  void rsp_merged_copy(int key, DLManagedTensor* out_idx, DLManagedTensor* out_val);
  // nnr_dev != nullptr: asynchronous — the unique-row count is left in that device int64, -1 returned
  int64_t row_sparse_pull(int key, DLManagedTensor* row_ids, DLManagedTensor* out_idx,
                          DLManagedTensor* out_val, int64_t* nnr_dev = nullptr);
  // owner-sharded row-sparse keys (rsp.cu)
  void shard_rsp_table(RspKey& k);
  void setup_rsp_exchange(RspKey& k, int64_t rows_now, int ncopies);
  void exchange_rsp_copies(RspKey& k, const struct RspInputs& in, int ncopies, struct RspSources* src, P2PFlags* bar);
  // Group wiring: this member's shard, and its siblings' shards as seen from here
  const float* rsp_table_ptr(int key);
  bool rsp_is_sharded(int key);
  void rsp_set_peer_tables(int key, const float* const* tables);
  void check_rsp_error(RspKey& k, uint32_t bits);
  void check_rsp_errors();

  // ---- used by Group (single process, several GPUs) -------------------------------------------------
  friend class Group;
  int device() const { return device_; }
  cudaStream_t stream() const { return stream_; }
  // value may live on any GPU of the process (peer copy)
  void init_any_device(int key, DLManagedTensor* val);
  // wire this member's segment `seg` to its siblings' buffers (all members hold the same layout)
  void link_peers(int seg, const std::vector<Store*>& all);
  int num_segments() const { return static_cast<int>(segs_.size()); }
  bool segment_linked(int seg) const { return segs_[seg]->p2p.ok; }
  // copy this member's shard of a state array of `key` into dst (full-key fp32 array on dst_device)
  bool copy_state_shard(int key, int which, float* dst, int dst_device, int64_t numel);
  void load_state_shard(int key, int which, const float* src, int src_device, int64_t numel);
  // after load_state_shard(which = 2) on every member: refresh every member's fp16 value from the owners
  void spread_fp16_shard(int key, const std::vector<Store*>& all);
  int64_t key_size(int key) { return info(key).size; }

  int rank() const { return rank_; }
  int world() const { return world_; }
  int64_t launches() const { return launches_; }
  const char* type() const { return "b200"; }
  const char* last_exchange() const { return last_exchange_; }

 private:
  friend struct RspKey;
  struct Entry {
    int key;
    std::vector<TensorView> vals, outs;
  };
  struct Cls {
    int mode, seg, ng, no;
    std::vector<const Entry*> ents;
    DensePlan* plan = nullptr;
  };
  struct BoundCall {
    std::vector<Entry> ents;
    std::vector<Cls> classes;
  };
  void prepare(Mode mode, int vnum, const int* vkeys, DLManagedTensor* const* vals, int onum,
               const int* okeys, DLManagedTensor* const* outs, BoundCall* bc, bool allow_staged);
  void execute(BoundCall& bc);
  // allocate whatever the call's classes will need (optimizer state): a Group does this for every
  // member before any member launches a kernel that waits for its siblings
  void preallocate(BoundCall& bc);
  void prof_begin(int phase = 0);
  void prof_end();
  KeyInfo& info(int key);
  void use_device();
  void check_key_type(bool is_str);
  void seal();
  void ensure_states(Segment& s);
  DensePlan& plan_for(int seg, const std::vector<int>& keys, int ngrad, int nout);
  void run_class(Mode mode, DensePlan& p, const std::vector<const Entry*>& ents);
  void pull_staged(const Entry& e);
  void update_chunks(DensePlan& p, const ChunkPlan& cp, const MtaCall& call, const float* gscale);
  const float* global_norm_scale(DensePlan& p, const ChunkPlan& cp, const MtaCall& gcall);
  void set_hyper_rsp(int key, float lr, float wd, int step);
  void setup_p2p(Segment& s);
  // allocate `bytes` for `buf`: through the symmetric allocator when one is set (filling peers / mc),
  // else with cudaMalloc.  Returns true when the memory is symmetric.
  bool alloc_exchange(DevBuf& buf, size_t bytes, void** peers, void** mc);
  void alloc_gbuf(Segment& s);
  void run_class_p2p(DensePlan& p, Segment& s, const std::vector<const Entry*>& ents, bool pull,
                     bool g_alias, bool o_alias);
  bool has_opt() const { return opt_.kind != B200KV_OPT_NONE; }
  bool use_mp(const Segment& s) const { return s.dtype == kF16 && opt_.multi_precision != 0; }

  // all-reduce of a few partial sums between kernels: peer-memory exchange when the segment's flag
  // blocks are mapped (always inside a Group), ncclAllReduce otherwise
  void allreduce_sums(Segment& s, float* sums, int n);

  int device_, rank_, world_;
  bool in_process_ = false;
  int grid_share_ = 1;
  cudaStream_t stream_ = nullptr;
  ncclComm* comm_ = nullptr;
  int key_type_ = -1;  // -1 unknown, 0 int, 1 str
  std::unordered_map<std::string, int> str_keys_;
  std::unordered_map<int, KeyInfo> keys_;
  std::vector<int> unsealed_;
  std::vector<std::unique_ptr<Segment>> segs_;
  std::map<std::string, std::unique_ptr<DensePlan>> plans_;
  std::unordered_map<int, std::shared_ptr<RspKey>> rsp_;
  b200kv_optimizer opt_;
  OptScalars sc_;
  TableRing ring_;
  DevBuf gnorm_;           // float[2]: scale, norm of the last clipped push
  bool gnorm_valid_ = false;
  int64_t launches_ = 0;
  const char* last_exchange_ = "local";
  b200kv_symm_alloc_fn symm_alloc_ = nullptr;
  void* symm_ctx_ = nullptr;
  std::unordered_map<int, std::unique_ptr<BoundCall>> bound_;
  int next_bound_ = 1;
  bool profiling_ = false;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events_;
  std::vector<int> prof_phase_;
  size_t prof_used_ = 0;
};

}  // namespace b200kv
