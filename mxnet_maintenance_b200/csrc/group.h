// One process driving several GPUs — the reference's own usage model for kvstore 'device' / 'nccl'
// (KVStoreLocal + CommDevice hold one NDArray per device, src/kvstore/comm.h:503-616,678-757;
// KVStoreNCCL::InitNCCL uses ncclCommInitAll, src/kvstore/kvstore_nccl.h:475-493; gluon.Trainer passes
// param.list_grad() / list_data(), python/mxnet/gluon/trainer.py:392-395).
//
// A Group owns one Store per device ("member", rank i of W) and drives all of them from the caller's
// thread.  Members exchange through peer memory (cudaDeviceEnablePeerAccess; no IPC, no NCCL; from 4
// distinct devices the buffers come from one NVSwitch multicast arena, mcast.h):
// the SAME fused reduce-scatter + update + all-gather kernel as in the one-process-per-GPU mode is
// launched once per member on that member's own stream, and its in-kernel flag barriers order the
// members on the GPUs, never on the host.
#pragma once
#include <memory>
#include <unordered_map>
#include <vector>

#include "mcast.h"
#include "store.h"

namespace b200kv {

class Group : public KVHandle {
 public:
  Group(int ndev, const int* dev_ids);
  ~Group() override;

  int size() const { return static_cast<int>(m_.size()); }
  Store& member(int i) { return *m_.at(i); }
  // the caller's stream on each member's device (default: the legacy default stream)
  void set_streams(int n, void* const* streams);
  void wait();
  void barrier();
  void lookup_keys(int num, const char* const* names, bool create, int* out);
  void init(int num, const int* keys, DLManagedTensor* const* vals);
  void dense(Store::Mode mode, int vnum, const int* vkeys, DLManagedTensor* const* vals, int onum,
             const int* okeys, DLManagedTensor* const* outs);
  int bind(Store::Mode mode, int vnum, const int* vkeys, DLManagedTensor* const* vals, int onum,
           const int* okeys, DLManagedTensor* const* outs);
  void run_bound(int id);
  void unbind(int id);
  void set_optimizer(const b200kv_optimizer& o);
  void set_hyper(int num, const int* keys, const float* lrs, const float* wds, const int* steps);
  float last_global_norm();
  DLManagedTensor* view(int member, int key, int which);
  void get_state(int key, int which, DLManagedTensor* out, int* present);
  void set_state(int key, int which, DLManagedTensor* in);
  void set_profiling(bool on);
  void profile_read_phases(double ms[8], int64_t count[8]);
  int64_t launches() const;
  const char* last_exchange() const { return m_[0]->last_exchange(); }
  // row sparse: every member reads all copies (its siblings' through peer memory); with an optimizer
  // the table is dealt cyclically (row id % members) over the members, each reducing / updating the rows it owns
  void init_rsp(int key, DLManagedTensor* dense_value);
  void push_rsp(int key, int ncopies, DLManagedTensor* const* idx, DLManagedTensor* const* val);
  int64_t rsp_merged_size(int key);
  void rsp_merged_copy(int key, DLManagedTensor* out_idx, DLManagedTensor* out_val);
  int64_t row_sparse_pull(int key, DLManagedTensor* row_ids, DLManagedTensor* out_idx, DLManagedTensor* out_val,
                          int64_t* nnr_dev = nullptr);

 private:
  struct Split {
    std::vector<std::vector<int>> keys;
    std::vector<std::vector<DLManagedTensor*>> arrs;
  };
  // GroupKVPairs across devices (kvstore_local.h:417-447): entry i goes to the member that drives
  // the device the array lives on
  Split split(int num, const int* keys, DLManagedTensor* const* arrs, const char* what) const;
  int member_of_device(int device, int nth) const;
  void fence_in(bool all_to_all = false);
  void fence_out();
  void sync_callers();
  void sync_members();
  void link_all();

  // NVSwitch multicast memory for the members' weight segments and gradient buckets (from 4 distinct
  // devices, or B200KV_NVLS=1); declared before the members so that it outlives them
  struct McCtx {
    McArena* arena;
    int member;
  };
  static int mc_alloc(void* ctx, size_t bytes, void** local, void** peers, void** mc);
  std::unique_ptr<McArena> arena_;
  std::vector<McCtx> mc_ctx_;
  std::vector<std::unique_ptr<Store>> m_;
  std::vector<Store*> raw_;
  std::vector<cudaStream_t> own_, caller_;
  std::vector<cudaEvent_t> ev_in_, ev_out_;
  std::unordered_map<int, std::vector<int>> bound_;
  int next_bound_ = 1;
};

}  // namespace b200kv
