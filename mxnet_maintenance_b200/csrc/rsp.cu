// Row-sparse keys of the store: ReduceRowSparse + lazy row update on push, unique + row gather on
// row_sparse_pull.  Reference: CommDevice::ReduceRowSparse (src/kvstore/comm.h:478-501),
// ElementwiseSumRsp (src/ndarray/ndarray_function.cc:59-176, .cu:104-190),
// KVStoreLocal::PullRowSparseImpl / Unique (src/kvstore/kvstore_local.h:294-314,468-512),
// UniqueImpl (src/kvstore/kvstore_utils.cu:44-110), BroadcastRowSparse (comm.h:618-674),
// SparseRetainOpForwardRspImpl (src/operator/tensor/sparse_retain-inl.h:262-323).
#include "rsp.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "nccl_dl.h"
#include "store.h"

namespace b200kv {

struct RspKey {
  int key = 0;
  int64_t num_rows = 0, row_len = 0;
  int bits = 0;             // significant bits of a row id
  // this rank holds the rows row_lo + j*stride, j in [0, nlocal): everything (0, 1, num_rows) unless the
  // key is sharded — then the rows with id % world == rank (rows_per_rank = ceil(num_rows / world))
  int64_t row_lo = 0, stride = 1, nlocal = 0, rows_per_rank = 0;
  DevBuf table, s0, s1;     // dense [num_rows, row_len] value and optimizer state
  bool has_s0 = false, has_s1 = false;
  bool full = true;         // the table holds every row; false: compact (c_idx, c_val) holds the value
  DevBuf c_idx, c_val, c_nnr;   // compact value after an updater-less push; c_nnr: int64 on device
  int64_t c_cap = 0;
  // bitmap path (rsp.cuh): one bit per row, padded to whole compaction CTAs; chain = the scan's
  // per-CTA hand-over words; count = {int64 row count, error bits (RspErr)}
  DevBuf bitmap, chain;
  int nwords = 0;
  // the row count row_sparse_pull has to hand back travels to this pinned word right after the scan,
  // so the host waits for the scan only — the row gather keeps running behind the returning call
  int64_t* host_count = nullptr;
  cudaEvent_t count_ready = nullptr;
  ~RspKey() {
    if (host_count) cudaFreeHost(host_count);
    if (count_ready) cudaEventDestroy(count_ready);
  }
  uint32_t scan_epoch = 0;
  bool use_bitmap() const { return num_rows <= (int64_t(1) << 28); }
  // sort / unique scratch, grow-only (count and uniq are shared with the bitmap path)
  DevBuf keys_a, keys_b, pay_a, pay_b, counts, bsum, seg_start, uniq, count;
  int64_t cap = 0, ucap = 0;
  uint32_t* err() const { return reinterpret_cast<uint32_t*>(count.as<int64_t>() + 1); }
  // room for n unique ids (bitmap path)
  void reserve_uniq(int64_t n) {
    if (n <= ucap && count.p != nullptr) return;
    ucap = std::max<int64_t>(std::max(n, ucap), 1024);
    uniq.reserve(sizeof(int64_t) * static_cast<size_t>(ucap));
    if (count.p == nullptr) {
      count.reserve(64);
      CUDA_OK(cudaMemset(count.p, 0, 64));
    }
  }
  void ensure_bitmap() {
    if (bitmap.p != nullptr) return;
    nwords = static_cast<int>((num_rows + 31) / 32);
    const int ncta = (nwords + kCompactWords - 1) / kCompactWords;
    bitmap.reserve(sizeof(uint32_t) * static_cast<size_t>(ncta) * kCompactWords);
    chain.reserve(sizeof(uint32_t) * 2 * static_cast<size_t>(ncta));
    CUDA_OK(cudaMemset(bitmap.p, 0, bitmap.bytes));
    CUDA_OK(cudaMemset(chain.p, 0, chain.bytes));
  }
  bool merged_valid = false;
  // multi-rank exchange buffers: this rank's rows packed contiguously, and every rank's rows
  DevBuf send_idx, send_val, recv_idx, recv_val, counts_dev;
  // Owner-sharded layout (several ranks + an optimizer): rank r keeps the rows with id % world == r of
  // the table and of the optimizer state only; peer_table[o] is rank o's shard as seen from here
  bool sharded = false;
  const float* peer_table[kMaxCopies] = {nullptr};
  // one-process-per-GPU exchange: every rank's (ids, rows) staging area and header / flag block are
  // mapped into every other rank (CUDA IPC); owners read the rows of their range straight from the
  // senders' staging areas over NVLink
  struct Exchange {
    bool ok = false;
    int64_t cap = 0;        // rows per copy slot
    int slots = 0;          // copy slots per rank
    DevBuf stage_idx, stage_val, flags;
    void* p_idx[kMaxCopies] = {nullptr};
    void* p_val[kMaxCopies] = {nullptr};
    uint32_t* p_flags[kMaxCopies] = {nullptr};
    uint32_t epoch = 0;
  } x;
  float lr = 0.01f, wd = 0.f;
  int step = 0;

  void reserve(int64_t n) {
    if (n <= cap) return;
    cap = std::max<int64_t>(n, 1024);
    const int nblocks = static_cast<int>((cap + kSortTile - 1) / kSortTile);
    const int nb = static_cast<int>((cap + kScanTile - 1) / kScanTile);
    keys_a.reserve(4 * cap); keys_b.reserve(4 * cap);
    pay_a.reserve(4 * cap);  pay_b.reserve(4 * cap);
    counts.reserve(sizeof(uint32_t) * 256 * static_cast<size_t>(nblocks));
    bsum.reserve(sizeof(uint32_t) * static_cast<size_t>(nb + 1));
    seg_start.reserve(sizeof(int32_t) * static_cast<size_t>(cap + 1));
    reserve_uniq(cap);
  }
};

namespace {

int bits_for(int64_t num_rows) {
  int b = 1;
  while (b < 32 && (int64_t(1) << b) < num_rows) ++b;
  return b;
}

// sorts (keys_a, pay_a)[0,n) by key, stable; returns true if the result is in the *_b buffers
int sort_pairs(RspKey& k, int n, cudaStream_t st, int64_t* launches) {
  const int nblocks = (n + kSortTile - 1) / kSortTile;
  uint32_t *ka = k.keys_a.as<uint32_t>(), *kb = k.keys_b.as<uint32_t>();
  uint32_t *pa = k.pay_a.as<uint32_t>(), *pb = k.pay_b.as<uint32_t>();
  int flips = 0;
  for (int shift = 0; shift < k.bits; shift += 8) {
    radix_hist_kernel<<<nblocks, kSortThreads, 0, st>>>(ka, n, shift, k.counts.as<uint32_t>(), nblocks);
    if (nblocks <= 256) {   // each scatter CTA derives its offsets from the raw counts itself
      radix_scatter_kernel<false><<<nblocks, kSortThreads, 0, st>>>(ka, pa, kb, pb, n, shift,
                                                                   k.counts.as<uint32_t>(), nblocks);
      *launches += 2;
    } else {
      scan_small_kernel<<<1, 1024, 0, st>>>(k.counts.as<uint32_t>(), 256 * nblocks);
      radix_scatter_kernel<true><<<nblocks, kSortThreads, 0, st>>>(ka, pa, kb, pb, n, shift,
                                                                  k.counts.as<uint32_t>(), nblocks);
      *launches += 3;
    }
    CUDA_OK(cudaGetLastError());
    std::swap(ka, kb);
    std::swap(pa, pb);
    ++flips;
  }
  return flips & 1;
}

// sorted keys -> uniq (int64), seg_start, count[0]
void unique_heads(RspKey& k, const uint32_t* keys, int n, cudaStream_t st, int64_t* launches) {
  const int nb = (n + kScanTile - 1) / kScanTile;
  head_count_kernel<<<nb, 256, 0, st>>>(keys, n, k.bsum.as<uint32_t>());
  if (nb <= 2048) {
    head_write_kernel<false><<<nb, 256, 0, st>>>(keys, n, k.bsum.as<uint32_t>(), k.uniq.as<int64_t>(),
                                                 k.seg_start.as<int32_t>(), k.count.as<int64_t>(), nb);
    *launches += 2;
  } else {
    scan_small_kernel<<<1, 1024, 0, st>>>(k.bsum.as<uint32_t>(), nb);
    head_write_kernel<true><<<nb, 256, 0, st>>>(keys, n, k.bsum.as<uint32_t>(), k.uniq.as<int64_t>(),
                                                k.seg_start.as<int32_t>(), k.count.as<int64_t>(), nb);
    *launches += 3;
  }
  CUDA_OK(cudaGetLastError());
}

}  // namespace

namespace {

// sorted unique ids of the marked rows (ids row_lo + j*stride, j < nlocal) -> k.uniq / k.count[0]; clears
// the bitmap
void compact_bitmap(RspKey& k, int64_t row_lo, int64_t stride, int64_t nlocal, cudaStream_t st, int64_t* launches) {
  const int nw = static_cast<int>((nlocal + 31) / 32);
  const int ncta = std::max(1, (nw + kCompactWords - 1) / kCompactWords);
  rsp_compact_kernel<<<ncta, 256, 0, st>>>(k.bitmap.as<uint32_t>(), ncta * kCompactWords, row_lo, stride,
                                           k.uniq.as<int64_t>(), k.count.as<int64_t>(),
                                           k.chain.as<uint32_t>(), ++k.scan_epoch);
  CUDA_OK(cudaGetLastError());
  ++*launches;
}

template <int MODE>
void launch_reduce2(bool vec, bool direct, int grid, cudaStream_t st, const RspSources& src, RspKey& k,
                    int64_t row_lo, int64_t stride, float* out_val, int64_t* out_idx, int64_t* out_nnr,
                    const RspUpdate& up) {
#define GO(V, D)                                                                                        \
  rsp_reduce2_kernel<MODE, V, D><<<grid, 128, 0, st>>>(src, k.uniq.as<int64_t>(), k.count.as<int64_t>(), \
                                                       k.row_len, k.num_rows, row_lo, stride, out_val,  \
                                                       out_idx, out_nnr, up, k.err())
  if (vec && direct) GO(true, true);
  else if (vec) GO(true, false);
  else if (direct) GO(false, true);
  else GO(false, false);
#undef GO
  CUDA_OK(cudaGetLastError());
}

}  // namespace

// ---- owner-sharded layout ---------------------------------------------------------------------------------
// With several ranks and an optimizer every rank keeps only the rows with id % world == rank of the table
// and of the optimizer state.  SURVEY §8(e) proposed owner = row_id / rows_per_rank; embedding ids are
// heavily skewed towards small ids (Zipf), which would put most of every push on rank 0, so the rows
// are dealt cyclically instead — every owner gets the same share of the hot rows.  The full table the
// key was initialised with is cut down at the first such push.
void Store::shard_rsp_table(RspKey& k) {
  KV_CHECK(k.use_bitmap(), "sharded row-sparse keys are limited to 2^28 rows");
  k.rows_per_rank = (k.num_rows + world_ - 1) / world_;
  k.row_lo = rank_;
  k.stride = world_;
  k.nlocal = k.num_rows > rank_ ? (k.num_rows - rank_ + world_ - 1) / world_ : 0;
  const size_t rowb = 4 * static_cast<size_t>(k.row_len);
  DevBuf shard;
  // at least 2 MiB: an allocation of its own, so that its IPC handle maps exactly this shard
  shard.reserve(std::max<size_t>(rowb * static_cast<size_t>(k.rows_per_rank), 2u << 20));
  CUDA_OK(cudaMemsetAsync(shard.p, 0, shard.bytes, stream_));
  if (k.nlocal > 0)   // every world-th row, starting at row `rank`
    CUDA_OK(cudaMemcpy2DAsync(shard.p, rowb, k.table.as<char>() + rowb * static_cast<size_t>(rank_),
                              rowb * static_cast<size_t>(world_), rowb, static_cast<size_t>(k.nlocal),
                              cudaMemcpyDeviceToDevice, stream_));
  CUDA_OK(cudaStreamSynchronize(stream_));
  k.table.release();
  std::swap(k.table.p, shard.p);
  std::swap(k.table.bytes, shard.bytes);
  for (int r = 0; r < kMaxCopies; ++r) k.peer_table[r] = nullptr;
  k.peer_table[rank_] = k.table.as<float>();
  k.sharded = true;
}

// One process per GPU: map every rank's staging areas, header / flag block and table shard into this
// process (CUDA IPC over NVLink peer access).  Collective; runs once per key, at its first sharded push.
void Store::setup_rsp_exchange(RspKey& k, int64_t rows_now, int ncopies) {
  KV_CHECK(comm_, "b200kv_comm_init has not been called");
  KV_CHECK(world_ <= kMaxCopies, "at most " << kMaxCopies << " ranks");
  NcclApi& nccl = NcclApi::get();
  ncclComm_t comm = reinterpret_cast<ncclComm_t>(comm_);
  RspKey::Exchange& x = k.x;
  // capacity of a copy slot and the number of slots, agreed over the ranks
  int64_t want[2] = {std::max<int64_t>(rows_now, 1), ncopies};
  if (const char* e = getenv("B200KV_RSP_CAPACITY")) want[0] = std::max<int64_t>(want[0], atoll(e));
  DevBuf tmp;
  tmp.reserve(64);
  CUDA_OK(cudaMemcpyAsync(tmp.p, want, 16, cudaMemcpyHostToDevice, stream_));
  NCCL_OK(nccl.AllReduce(tmp.p, tmp.p, 2, ncclInt64, ncclMax, comm, stream_));
  CUDA_OK(cudaMemcpyAsync(want, tmp.p, 16, cudaMemcpyDeviceToHost, stream_));
  CUDA_OK(cudaStreamSynchronize(stream_));
  x.cap = getenv("B200KV_RSP_CAPACITY") ? want[0] : std::max<int64_t>(2 * want[0], 65536);
  x.slots = static_cast<int>(want[1]);
  KV_CHECK(x.slots * world_ <= kMaxSources, "too many row-sparse copies per push across the ranks");
  const size_t rowb = 4 * static_cast<size_t>(k.row_len);
  x.stage_idx.reserve(std::max<size_t>(8 * static_cast<size_t>(x.cap) * x.slots, 2u << 20));
  x.stage_val.reserve(std::max<size_t>(rowb * static_cast<size_t>(x.cap) * x.slots, 2u << 20));
  x.flags.reserve(2u << 20);
  CUDA_OK(cudaMemsetAsync(x.flags.p, 0, 4096, stream_));
  struct Handles { cudaIpcMemHandle_t i, v, f, t; };
  Handles mine;
  int ok = 1;
  if (cudaIpcGetMemHandle(&mine.i, x.stage_idx.p) != cudaSuccess) ok = 0;
  if (cudaIpcGetMemHandle(&mine.v, x.stage_val.p) != cudaSuccess) ok = 0;
  if (cudaIpcGetMemHandle(&mine.f, x.flags.p) != cudaSuccess) ok = 0;
  if (cudaIpcGetMemHandle(&mine.t, k.table.p) != cudaSuccess) ok = 0;
  (void)cudaGetLastError();
  const size_t hb = sizeof(Handles);
  DevBuf hx;
  hx.reserve(hb * world_ + 64);
  char* sp = hx.as<char>();
  CUDA_OK(cudaMemcpyAsync(sp + hb * rank_, &mine, hb, cudaMemcpyHostToDevice, stream_));
  NCCL_OK(nccl.AllGather(sp + hb * rank_, sp, hb, ncclChar, comm, stream_));
  std::vector<Handles> all(world_);
  CUDA_OK(cudaMemcpyAsync(all.data(), sp, hb * world_, cudaMemcpyDeviceToHost, stream_));
  CUDA_OK(cudaStreamSynchronize(stream_));
  for (int r = 0; r < world_ && ok; ++r) {
    if (r == rank_) {
      x.p_idx[r] = x.stage_idx.p; x.p_val[r] = x.stage_val.p; x.p_flags[r] = x.flags.as<uint32_t>();
      k.peer_table[r] = k.table.as<float>();
      continue;
    }
    void *f = nullptr, *t = nullptr;
    if (cudaIpcOpenMemHandle(&x.p_idx[r], all[r].i, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess ||
        cudaIpcOpenMemHandle(&x.p_val[r], all[r].v, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess ||
        cudaIpcOpenMemHandle(&f, all[r].f, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess ||
        cudaIpcOpenMemHandle(&t, all[r].t, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess)
      ok = 0;
    x.p_flags[r] = static_cast<uint32_t*>(f);
    k.peer_table[r] = static_cast<const float*>(t);
  }
  (void)cudaGetLastError();
  int* flag = reinterpret_cast<int*>(sp + hb * world_);
  CUDA_OK(cudaMemcpyAsync(flag, &ok, sizeof(int), cudaMemcpyHostToDevice, stream_));
  NCCL_OK(nccl.AllReduce(flag, flag, 1, ncclInt, ncclMin, comm, stream_));   // also: every flag block is zero by now
  int agreed = 0;
  CUDA_OK(cudaMemcpyAsync(&agreed, flag, sizeof(int), cudaMemcpyDeviceToHost, stream_));
  CUDA_OK(cudaStreamSynchronize(stream_));
  KV_CHECK(agreed == 1, "row-sparse exchange: the ranks cannot map each other's memory (CUDA IPC / NVLink peer "
                        "access is required for sharded row-sparse keys)");
  x.ok = true;
}

// Stage this rank's copies (slot c = copy c), publish their lengths in the header, arrive at the
// "ready" barrier; *src then names EVERY rank's slots in (rank, copy) order.
void Store::exchange_rsp_copies(RspKey& k, const RspInputs& in, int ncopies, RspSources* src, P2PFlags* bar) {
  int64_t most = 1;
  for (int c = 0; c < in.n; ++c) most = std::max(most, in.start[c + 1] - in.start[c]);
  if (!k.x.ok) setup_rsp_exchange(k, most, ncopies);
  RspKey::Exchange& x = k.x;
  KV_CHECK(in.n <= x.slots, "row-sparse push: " << in.n << " copies, but the exchange was set up for " << x.slots
                                                 << " per rank (every rank must push the same number of copies)");
  KV_CHECK(most <= x.cap, "row-sparse push of " << most << " rows exceeds the exchange capacity of " << x.cap
                              << " rows per copy fixed at the first push; set B200KV_RSP_CAPACITY to the largest "
                                 "number of rows a push may carry");
  const size_t rowb = 4 * static_cast<size_t>(k.row_len);
  int64_t lens[kMaxCopies] = {0};
  for (int c = 0; c < in.n; ++c) {
    lens[c] = in.start[c + 1] - in.start[c];
    CUDA_OK(cudaMemcpyAsync(x.stage_idx.as<int64_t>() + static_cast<int64_t>(c) * x.cap, in.idx[c],
                            8 * static_cast<size_t>(lens[c]), cudaMemcpyDeviceToDevice, stream_));
    CUDA_OK(cudaMemcpyAsync(x.stage_val.as<char>() + rowb * static_cast<size_t>(c) * x.cap, in.val[c],
                            rowb * static_cast<size_t>(lens[c]), cudaMemcpyDeviceToDevice, stream_));
  }
  CUDA_OK(cudaMemcpyAsync(x.flags.as<char>() + 1024, lens, sizeof(lens), cudaMemcpyHostToDevice, stream_));
  for (int r = 0; r < kMaxCopies; ++r) bar->flags[r] = x.p_flags[r];
  bar->world = world_;
  bar->rank = rank_;
  bar->epoch = ++x.epoch;
  launches_ += launch_p2p_barrier(*bar, 0, stream_);   // every rank's slots and header are in place
  std::memset(src, 0, sizeof(*src));
  int n = 0;
  for (int r = 0; r < world_; ++r)
    for (int c = 0; c < x.slots; ++c) {
      src->idx[n] = static_cast<const int64_t*>(x.p_idx[r]) + static_cast<int64_t>(c) * x.cap;
      src->val[n] = reinterpret_cast<const float*>(static_cast<const char*>(x.p_val[r]) + rowb * static_cast<size_t>(c) * x.cap);
      src->len_ptr[n] = reinterpret_cast<const int64_t*>(reinterpret_cast<const char*>(x.p_flags[r]) + 1024) + c;
      ++n;
    }
  src->n = n;
}

const float* Store::rsp_table_ptr(int key) { return rsp_.at(key)->table.as<float>(); }
bool Store::rsp_is_sharded(int key) { return rsp_.at(key)->sharded; }
void Store::rsp_set_peer_tables(int key, const float* const* tables) {
  RspKey& k = *rsp_.at(key);
  for (int r = 0; r < world_; ++r) k.peer_table[r] = tables[r];
}

// inside a Group the arrays of a call may live on any device of the process (peer memory)
static int dev_of(const DLManagedTensor* t, bool in_process, int own) {
  return (in_process && t != nullptr) ? t->dl_tensor.device.device_id : own;
}

void Store::init_rsp(int key, DLManagedTensor* dense_value) {
  use_device();
  if (key_type_ < 0) key_type_ = 0;
  KV_CHECK(keys_.find(key) == keys_.end() && rsp_.find(key) == rsp_.end(), "duplicate init of key " << key);
  const TensorView v = view_of(dense_value, dev_of(dense_value, in_process_, device_), "row-sparse init value");
  KV_CHECK(v.dtype == kF32, "row-sparse keys are float32");
  KV_CHECK(v.ndim >= 2, "row-sparse init needs a [num_rows, ...] table");
  std::shared_ptr<RspKey> k(new RspKey());
  k->key = key;
  k->num_rows = v.shape[0];
  k->row_len = v.numel / std::max<int64_t>(v.shape[0], 1);
  KV_CHECK(k->num_rows < (int64_t(1) << 32), "row-sparse tables are limited to 2^32 rows");
  k->bits = bits_for(k->num_rows);
  k->row_lo = 0;
  k->stride = 1;
  k->nlocal = k->rows_per_rank = k->num_rows;
  k->table.reserve(static_cast<size_t>(v.numel) * 4 + 16);
  CUDA_OK(cudaMemcpyAsync(k->table.p, v.data, static_cast<size_t>(v.numel) * 4, cudaMemcpyDefault, stream_));
  if (world_ > 1 && !in_process_) {
    KV_CHECK(comm_, "b200kv_comm_init has not been called");
    NCCL_OK(NcclApi::get().Broadcast(k->table.p, k->table.p, v.numel, ncclFloat32, 0,
                                     reinterpret_cast<ncclComm_t>(comm_), stream_));
  }
  rsp_.emplace(key, k);
}

void Store::set_hyper_rsp(int key, float lr, float wd, int step) {
  RspKey& k = *rsp_.at(key);
  if (!std::isnan(lr)) k.lr = lr;
  if (!std::isnan(wd)) k.wd = wd;
  if (step >= 0) k.step = step;
}

void Store::push_rsp(int key, int ncopies, DLManagedTensor* const* idx, DLManagedTensor* const* val) {
  use_device();
  auto it = rsp_.find(key);
  KV_CHECK(it != rsp_.end(), "key " << key << " has not been initialized as row_sparse");
  RspKey& k = *it->second;
  KV_CHECK(ncopies >= 1 && ncopies <= kMaxCopies, "between 1 and " << kMaxCopies << " copies per push");
  RspInputs in;
  std::memset(&in, 0, sizeof(in));
  int64_t total = 0;
  int n = 0;
  for (int c = 0; c < ncopies; ++c) {
    const TensorView vi = view_of(idx[c], dev_of(idx[c], in_process_, device_), "row-sparse indices");
    const TensorView vv = view_of(val[c], dev_of(val[c], in_process_, device_), "row-sparse values");
    KV_CHECK(vi.dtype == kI64, "row-sparse indices must be int64 (src/ndarray/ndarray.cc:65)");
    KV_CHECK(vv.dtype == kF32, "row-sparse values must be float32");
    KV_CHECK(vv.numel == vi.numel * k.row_len, "values must be [nnr, " << k.row_len << "]");
    if (vi.numel == 0) continue;   // storage not initialised: skipped (ndarray_function.cc:81,129)
    in.idx[n] = vi.data;
    in.val[n] = static_cast<const float*>(vv.data);
    in.start[n] = total;
    total += vi.numel;
    ++n;
  }
  in.n = n;
  in.start[n] = total;
  const bool upd = has_opt();
  bool gathered = false;
  if (world_ > 1 && !in_process_ && !upd) {
    gathered = true;
    KV_CHECK(!k.sharded, "row-sparse key " << key << " is sharded across the ranks for optimizer updates; "
                                                   "an updater-less push cannot replace its value any more");
    // Every rank holds the whole table (replicated), so the exchange is an all-gather of the
    // ranks' (ids, rows); afterwards each rank reduces and updates identically.  Rank r's rows keep
    // their device order and ranks are concatenated in rank order = the reference's device order.
    KV_CHECK(comm_, "b200kv_comm_init has not been called");
    NcclApi& nccl = NcclApi::get();
    ncclComm_t comm = reinterpret_cast<ncclComm_t>(comm_);
    k.counts_dev.reserve(8 * static_cast<size_t>(world_));
    std::vector<int64_t> counts(world_, 0);
    CUDA_OK(cudaMemcpyAsync(k.counts_dev.as<int64_t>() + rank_, &total, 8, cudaMemcpyHostToDevice, stream_));
    NCCL_OK(nccl.AllGather(k.counts_dev.as<int64_t>() + rank_, k.counts_dev.p, 1, ncclInt64, comm, stream_));
    CUDA_OK(cudaMemcpyAsync(counts.data(), k.counts_dev.p, 8 * static_cast<size_t>(world_),
                            cudaMemcpyDeviceToHost, stream_));
    CUDA_OK(cudaStreamSynchronize(stream_));
    const int64_t maxn = std::max<int64_t>(*std::max_element(counts.begin(), counts.end()), 1);
    const size_t rowb = 4 * static_cast<size_t>(k.row_len);
    k.send_idx.reserve(8 * static_cast<size_t>(maxn));
    k.send_val.reserve(rowb * static_cast<size_t>(maxn));
    k.recv_idx.reserve(8 * static_cast<size_t>(maxn) * world_);
    k.recv_val.reserve(rowb * static_cast<size_t>(maxn) * world_);
    for (int c = 0; c < n; ++c) {
      const int64_t cnt = in.start[c + 1] - in.start[c];
      CUDA_OK(cudaMemcpyAsync(k.send_idx.as<int64_t>() + in.start[c], in.idx[c], 8 * static_cast<size_t>(cnt),
                              cudaMemcpyDeviceToDevice, stream_));
      CUDA_OK(cudaMemcpyAsync(k.send_val.as<char>() + rowb * in.start[c], in.val[c], rowb * static_cast<size_t>(cnt),
                              cudaMemcpyDeviceToDevice, stream_));
    }
    NCCL_OK(nccl.AllGather(k.send_idx.p, k.recv_idx.p, maxn, ncclInt64, comm, stream_));
    NCCL_OK(nccl.AllGather(k.send_val.p, k.recv_val.p, static_cast<size_t>(maxn) * k.row_len, ncclFloat32, comm,
                           stream_));
    std::memset(&in, 0, sizeof(in));
    total = 0;
    n = 0;
    for (int r = 0; r < world_; ++r) {
      if (counts[r] == 0) continue;
      in.idx[n] = k.recv_idx.as<int64_t>() + static_cast<int64_t>(r) * maxn;
      in.val[n] = k.recv_val.as<float>() + static_cast<int64_t>(r) * maxn * k.row_len;
      in.start[n] = total;
      total += counts[r];
      ++n;
    }
    in.n = n;
    in.start[n] = total;
  }
  KV_CHECK(total < (int64_t(1) << 31), "too many rows in one push");
  k.merged_valid = false;
  const bool exchange = world_ > 1 && upd && !in_process_;   // peers may have rows even if this rank has none
  if (total == 0 && !exchange) {
    if (!upd) {  // all inputs empty -> the stored value becomes empty (ndarray_function.cu:113-123)
      k.full = false;
      k.c_nnr.reserve(8);
      CUDA_OK(cudaMemsetAsync(k.c_nnr.p, 0, 8, stream_));
      k.merged_valid = true;
    }
    return;
  }
  RspUpdate up;
  std::memset(&up, 0, sizeof(up));
  if (!upd) {
    // no updater: the merged rows replace the stored value (kvstore_local.h:268-274)
    if (total > k.c_cap) {
      k.c_cap = total;
      k.c_idx.reserve(8 * static_cast<size_t>(total));
      k.c_val.reserve(4 * static_cast<size_t>(total) * k.row_len);
    }
    k.c_nnr.reserve(8);
    up.mode = kRspStore;
  } else {
    KV_CHECK(k.full, "lazy row update needs the dense table (the key was overwritten by an updater-less push)");
    KV_CHECK(opt_.kind == B200KV_OPT_SGD || opt_.kind == B200KV_OPT_ADAM || opt_.kind == B200KV_OPT_ADAGRAD,
             "row-sparse keys support lazy SGD(-momentum), Adam and AdaGrad updates");
    // the sparse AdaGrad op only exists for wd == 0 (AdagradStorageType, optimizer_op-inl.h:2497-2502)
    KV_CHECK(opt_.kind != B200KV_OPT_ADAGRAD || k.wd == 0.f, "sparse AdaGrad requires wd == 0");
    if (world_ > 1 && !k.sharded) shard_rsp_table(k);
    const size_t bytes = static_cast<size_t>(k.sharded ? k.rows_per_rank : k.num_rows) * k.row_len * 4;
    const bool want_s0 = opt_.kind == B200KV_OPT_ADAM || opt_.kind == B200KV_OPT_ADAGRAD || opt_.momentum != 0.f;
    if (want_s0 && !k.has_s0) {
      k.s0.reserve(bytes);
      CUDA_OK(cudaMemsetAsync(k.s0.p, 0, bytes, stream_));
      k.has_s0 = true;
    }
    if (opt_.kind == B200KV_OPT_ADAM && !k.has_s1) {
      k.s1.reserve(bytes);
      CUDA_OK(cudaMemsetAsync(k.s1.p, 0, bytes, stream_));
      k.has_s1 = true;
    }
    up.lr = k.lr; up.wd = k.wd; up.momentum = opt_.momentum; up.beta1 = opt_.beta1; up.beta2 = opt_.beta2;
    up.eps = opt_.epsilon; up.rescale = opt_.rescale_grad; up.clip = opt_.clip_gradient;
    up.has_mom = opt_.kind == B200KV_OPT_SGD && opt_.momentum != 0.f;
    up.table = k.table.as<float>(); up.s0 = k.s0.as<float>(); up.s1 = k.s1.as<float>();
    up.mode = opt_.kind == B200KV_OPT_SGD ? kRspSgd : (opt_.kind == B200KV_OPT_ADAM ? kRspAdam : kRspAdagrad);
    // lazy_update=False: every row is updated (there is no standard form of the sparse AdaGrad op)
    up.std_update = opt_.std_update != 0 && opt_.kind != B200KV_OPT_ADAGRAD;
    if (up.std_update) {
      KV_CHECK(k.use_bitmap(), "the standard (non-lazy) row-sparse update is limited to tables of 2^28 rows");
      // (between processes the peers' row counts are only known on the device: room for the whole shard)
      const int64_t room = std::max<int64_t>(1, exchange ? k.rows_per_rank
                                                         : (k.sharded ? std::min<int64_t>(k.rows_per_rank, total) : total));
      if (room > k.c_cap) {   // the merged gradient is staged where an updater-less push keeps its value
        k.c_cap = room;
        k.c_idx.reserve(8 * static_cast<size_t>(room));
        k.c_val.reserve(4 * static_cast<size_t>(room) * k.row_len);
      }
      k.c_nnr.reserve(8);
    }
  }
  const bool vec = (k.row_len % 4) == 0;   // cudaMalloc'd tables and DLPack row blocks are 16-B aligned
  const int nn = static_cast<int>(total);

  KV_CHECK(!k.sharded || k.use_bitmap(), "sharded row-sparse keys are limited to 2^28 rows");
  if (k.use_bitmap() && !gathered) {
    // ---- bitmap path: the copies are sorted-unique already, nothing is sorted --------------------------
    RspSources src;
    std::memset(&src, 0, sizeof(src));
    int64_t maxlen = 0, bound = total;
    P2PFlags bar;
    if (exchange) {
      // stage this rank's copies where the owners can read them, publish their lengths, and let every
      // rank arrive: afterwards the sources are ALL ranks' copies, in (rank, copy) order = the
      // reference's device order
      exchange_rsp_copies(k, in, ncopies, &src, &bar);
      maxlen = k.x.cap;
      bound = std::min<int64_t>(k.rows_per_rank, k.x.cap * k.x.slots * world_);
    } else {
      src.n = in.n;
      for (int c = 0; c < in.n; ++c) {
        src.idx[c] = in.idx[c];
        src.val[c] = in.val[c];
        src.len[c] = in.start[c + 1] - in.start[c];
        maxlen = std::max(maxlen, src.len[c]);
      }
      if (k.sharded) bound = std::min<int64_t>(k.rows_per_rank, total);
    }
    k.reserve_uniq(std::max<int64_t>(bound, 1));
    const bool direct = src.n == 1 && !k.sharded && !up.std_update;
    if (!direct) {
      k.ensure_bitmap();
      const int g = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>((maxlen + 255) / 256, sm_count(device_) * 4)));
      rsp_mark_kernel<int64_t, true><<<g, 256, 0, stream_>>>(src, k.num_rows, k.row_lo, k.stride,
                                                           k.bitmap.as<uint32_t>(), k.err());
      CUDA_OK(cudaGetLastError());
      ++launches_;
      compact_bitmap(k, k.row_lo, k.stride, k.nlocal, stream_, &launches_);
    }
    const int grid = static_cast<int>(std::max<int64_t>(
        1, std::min<int64_t>(exchange ? bound : (direct ? maxlen : total), sm_count(device_) * 16)));
    const bool stage = !upd || up.std_update;
    float* ov = stage ? k.c_val.as<float>() : nullptr;
    int64_t* oi = stage ? k.c_idx.as<int64_t>() : nullptr;
    int64_t* on = stage ? k.c_nnr.as<int64_t>() : nullptr;
    if (up.std_update) {
      // merged rows first (sum in device order), then one pass over EVERY row of this rank's shard
      if (exchange) KV_CHECK(bound <= k.c_cap, "internal: staging too small for the standard update");
      launch_reduce2<kRspStore>(vec, false, grid, stream_, src, k, k.row_lo, k.stride, ov, oi, on, up);
      const int g2 = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(k.nlocal, sm_count(device_) * 16)));
#define STD(MODE)                                                                                             \
  do {                                                                                                        \
    if (vec) rsp_std_update_kernel<MODE, true><<<g2, 128, 0, stream_>>>(oi, on, ov, k.row_len, k.row_lo, k.stride, k.nlocal, up);  \
    else rsp_std_update_kernel<MODE, false><<<g2, 128, 0, stream_>>>(oi, on, ov, k.row_len, k.row_lo, k.stride, k.nlocal, up);     \
  } while (0)
      if (up.mode == kRspSgd) STD(kRspSgd);
      else STD(kRspAdam);
#undef STD
      CUDA_OK(cudaGetLastError());
      launches_ += 2;
      if (exchange) {
        bar.epoch = k.x.epoch;
        launches_ += launch_p2p_barrier(bar, 1, stream_);
      }
      return;
    }
    switch (up.mode) {
      case kRspStore: launch_reduce2<kRspStore>(vec, direct, grid, stream_, src, k, k.row_lo, k.stride, ov, oi, on, up); break;
      case kRspSgd: launch_reduce2<kRspSgd>(vec, direct, grid, stream_, src, k, k.row_lo, k.stride, ov, oi, on, up); break;
      case kRspAdam: launch_reduce2<kRspAdam>(vec, direct, grid, stream_, src, k, k.row_lo, k.stride, ov, oi, on, up); break;
      default: launch_reduce2<kRspAdagrad>(vec, direct, grid, stream_, src, k, k.row_lo, k.stride, ov, oi, on, up); break;
    }
    ++launches_;
    if (exchange) {   // every owner has finished reading the senders' staging areas; all shards are updated
      bar.epoch = k.x.epoch;
      launches_ += launch_p2p_barrier(bar, 1, stream_);
    }
    if (!upd) {
      k.full = false;
      k.merged_valid = true;
    }
    return;
  }

  // ---- radix-sort path (tables beyond 2^28 rows) --------------------------------------------------------
  k.reserve(total);
  CUDA_OK(cudaMemsetAsync(k.count.p, 0, 8, stream_));
  make_keys_kernel<int64_t><<<std::min((nn + 255) / 256, 148 * 8), 256, 0, stream_>>>(
      in, total, k.keys_a.as<uint32_t>(), k.pay_a.as<uint32_t>());
  ++launches_;
  const int in_b = sort_pairs(k, nn, stream_, &launches_);
  const uint32_t* pay = in_b ? k.pay_b.as<uint32_t>() : k.pay_a.as<uint32_t>();
  unique_heads(k, in_b ? k.keys_b.as<uint32_t>() : k.keys_a.as<uint32_t>(), nn, stream_, &launches_);
  const int grid = std::min(nn, sm_count(device_) * 16);
#define GO(MODE, OUT)                                                                                   \
  do {                                                                                                  \
    if (vec) rsp_reduce_kernel<MODE, true><<<grid, 128, 0, stream_>>>(                                   \
        in, pay, k.seg_start.as<int32_t>(), k.uniq.as<int64_t>(), k.count.as<int64_t>(), k.row_len, OUT, up); \
    else rsp_reduce_kernel<MODE, false><<<grid, 128, 0, stream_>>>(                                      \
        in, pay, k.seg_start.as<int32_t>(), k.uniq.as<int64_t>(), k.count.as<int64_t>(), k.row_len, OUT, up); \
  } while (0)
  if (!upd) {
    GO(kRspStore, k.c_val.as<float>());
    copy_i64_kernel<<<std::min((nn + 255) / 256, 1024), 256, 0, stream_>>>(
        k.uniq.as<int64_t>(), k.c_idx.as<int64_t>(), k.count.as<int64_t>());
    CUDA_OK(cudaMemcpyAsync(k.c_nnr.p, k.count.p, 8, cudaMemcpyDeviceToDevice, stream_));
    launches_ += 2;
    k.full = false;
    k.merged_valid = true;
  } else {
    if (up.mode == kRspSgd) GO(kRspSgd, nullptr);
    else if (up.mode == kRspAdam) GO(kRspAdam, nullptr);
    else GO(kRspAdagrad, nullptr);
    ++launches_;
  }
#undef GO
  CUDA_OK(cudaGetLastError());
}

int64_t Store::rsp_merged_size(int key) {
  use_device();
  auto it = rsp_.find(key);
  KV_CHECK(it != rsp_.end(), "key " << key << " has not been initialized as row_sparse");
  RspKey& k = *it->second;
  KV_CHECK(k.merged_valid, "no merged row-sparse value: the last push ran an optimizer (rows were updated in place)");
  int64_t n = 0;
  CUDA_OK(cudaMemcpyAsync(&n, k.c_nnr.p, 8, cudaMemcpyDeviceToHost, stream_));
  CUDA_OK(cudaStreamSynchronize(stream_));
  return n;
}

void Store::rsp_merged_copy(int key, DLManagedTensor* out_idx, DLManagedTensor* out_val) {
  const int64_t n = rsp_merged_size(key);
  RspKey& k = *rsp_.at(key);
  const TensorView vi = view_of(out_idx, device_, "merged indices out");
  const TensorView vv = view_of(out_val, device_, "merged values out");
  KV_CHECK(vi.dtype == kI64 && vi.numel >= n, "indices out must be int64 with at least " << n << " entries");
  KV_CHECK(vv.dtype == kF32 && vv.numel >= n * k.row_len, "values out too small");
  if (n == 0) return;
  CUDA_OK(cudaMemcpyAsync(vi.data, k.c_idx.p, 8 * static_cast<size_t>(n), cudaMemcpyDeviceToDevice, stream_));
  CUDA_OK(cudaMemcpyAsync(vv.data, k.c_val.p, 4 * static_cast<size_t>(n) * k.row_len, cudaMemcpyDeviceToDevice,
                          stream_));
}

int64_t Store::row_sparse_pull(int key, DLManagedTensor* row_ids, DLManagedTensor* out_idx,
                               DLManagedTensor* out_val, int64_t* nnr_dev) {
  use_device();
  auto it = rsp_.find(key);
  KV_CHECK(it != rsp_.end(), "key " << key << " has not been initialized as row_sparse");
  RspKey& k = *it->second;
  const TensorView ids = view_of(row_ids, dev_of(row_ids, in_process_, device_), "row_ids");
  const TensorView vi = view_of(out_idx, device_, "row_sparse_pull indices out");
  const TensorView vv = view_of(out_val, device_, "row_sparse_pull values out");
  KV_CHECK(ids.dtype == kI64 || ids.dtype == kI32 || ids.dtype == kF32, "row_ids must be int64, int32 or float32");
  KV_CHECK(vi.dtype == kI64 && vv.dtype == kF32, "outputs must be int64 indices and float32 values");
  const int64_t total = ids.numel;
  KV_CHECK(total < (int64_t(1) << 31), "too many row ids");
  KV_CHECK(vi.numel >= total && vv.numel >= total * k.row_len,
           "outputs must have room for numel(row_ids) = " << total << " rows");
  if (total == 0) return 0;
  const int nn = static_cast<int>(total);
  const bool vec = (k.row_len % 4) == 0;
  bool early = false;
  RspTables tabs;
  std::memset(&tabs, 0, sizeof(tabs));
  if (k.sharded) {
    for (int r = 0; r < world_; ++r) {
      KV_CHECK(k.peer_table[r] != nullptr, "internal: the shard of rank " << r << " is not mapped");
      tabs.t[r] = k.peer_table[r];
    }
    tabs.nranks = world_;
  } else {
    tabs.t[0] = k.table.as<float>();
    tabs.nranks = 1;
  }
  if (k.use_bitmap()) {
    // mark -> chained popcount scan (emits the ids in ascending order) -> gather: 3 launches
    k.reserve_uniq(total);
    k.ensure_bitmap();
    RspSources src;
    std::memset(&src, 0, sizeof(src));
    src.n = 1;
    src.idx[0] = ids.data;
    src.len[0] = total;
    const int g = std::min((nn + 255) / 256, sm_count(device_) * 4);
    uint32_t* bm = k.bitmap.as<uint32_t>();
    if (ids.dtype == kI64) rsp_mark_kernel<int64_t, false><<<g, 256, 0, stream_>>>(src, k.num_rows, 0, 1, bm, k.err());
    else if (ids.dtype == kI32) rsp_mark_kernel<int32_t, false><<<g, 256, 0, stream_>>>(src, k.num_rows, 0, 1, bm, k.err());
    else rsp_mark_kernel<float, false><<<g, 256, 0, stream_>>>(src, k.num_rows, 0, 1, bm, k.err());
    CUDA_OK(cudaGetLastError());
    ++launches_;
    compact_bitmap(k, 0, 1, k.num_rows, stream_, &launches_);
    if (nnr_dev == nullptr) {
      if (k.host_count == nullptr) {
        CUDA_OK(cudaMallocHost(reinterpret_cast<void**>(&k.host_count), 16));
        CUDA_OK(cudaEventCreateWithFlags(&k.count_ready, cudaEventDisableTiming));
      }
      CUDA_OK(cudaMemcpyAsync(k.host_count, k.count.p, 16, cudaMemcpyDeviceToHost, stream_));
      CUDA_OK(cudaEventRecord(k.count_ready, stream_));
      early = true;
    }
  } else {
    k.reserve(total);
    CUDA_OK(cudaMemsetAsync(k.count.p, 0, 8, stream_));
    RspInputs in;
    std::memset(&in, 0, sizeof(in));
    in.idx[0] = ids.data;
    in.start[0] = 0;
    in.start[1] = total;
    in.n = 1;
    const int g = std::min((nn + 255) / 256, 148 * 8);
    if (ids.dtype == kI64) make_keys_kernel<int64_t><<<g, 256, 0, stream_>>>(in, total, k.keys_a.as<uint32_t>(), k.pay_a.as<uint32_t>());
    else if (ids.dtype == kI32) make_keys_kernel<int32_t><<<g, 256, 0, stream_>>>(in, total, k.keys_a.as<uint32_t>(), k.pay_a.as<uint32_t>());
    else make_keys_kernel<float><<<g, 256, 0, stream_>>>(in, total, k.keys_a.as<uint32_t>(), k.pay_a.as<uint32_t>());
    ++launches_;
    const int in_b = sort_pairs(k, nn, stream_, &launches_);
    unique_heads(k, in_b ? k.keys_b.as<uint32_t>() : k.keys_a.as<uint32_t>(), nn, stream_, &launches_);
  }
  const int grid = std::min(nn, sm_count(device_) * 16);
  if (vec) rsp_gather_kernel<true><<<grid, 128, 0, stream_>>>(
      k.uniq.as<int64_t>(), k.count.as<int64_t>(), k.row_len, k.full ? 1 : 0, tabs, k.num_rows,
      k.c_idx.as<int64_t>(), k.c_val.as<float>(), k.c_nnr.as<int64_t>(), static_cast<int64_t*>(vi.data),
      static_cast<float*>(vv.data));
  else rsp_gather_kernel<false><<<grid, 128, 0, stream_>>>(
      k.uniq.as<int64_t>(), k.count.as<int64_t>(), k.row_len, k.full ? 1 : 0, tabs, k.num_rows,
      k.c_idx.as<int64_t>(), k.c_val.as<float>(), k.c_nnr.as<int64_t>(), static_cast<int64_t*>(vi.data),
      static_cast<float*>(vv.data));
  CUDA_OK(cudaGetLastError());
  ++launches_;
  if (nnr_dev != nullptr) {   // asynchronous form: the count stays on the device
    CUDA_OK(cudaMemcpyAsync(nnr_dev, k.count.p, 8, cudaMemcpyDeviceToDevice, stream_));
    return -1;
  }
  // the caller needs the row count to shape its output (the reference syncs here too,
  // kvstore_utils.cu:92-96); the error bits of this and earlier pushes ride along
  int64_t host[2] = {0, 0};
  if (early) {   // outputs are complete in STREAM order; the host only needs the count
    CUDA_OK(cudaEventSynchronize(k.count_ready));
    host[0] = k.host_count[0];
    host[1] = k.host_count[1];
  } else {
    CUDA_OK(cudaMemcpyAsync(host, k.count.p, 16, cudaMemcpyDeviceToHost, stream_));
    CUDA_OK(cudaStreamSynchronize(stream_));
  }
  check_rsp_error(k, static_cast<uint32_t>(host[1]));
  return host[0];
}

// row-id faults found on the device (push ids outside the table or not sorted-unique, pull ids outside
// the table) surface at the next point the host synchronises anyway
void Store::check_rsp_error(RspKey& k, uint32_t bits) {
  if (bits == 0) return;
  CUDA_OK(cudaMemsetAsync(k.err(), 0, 4, stream_));
  KV_CHECK((bits & kRspErrRange) == 0, "row-sparse key " << k.key << ": a row id lies outside [0, " << k.num_rows
                                                           << ") (the offending rows were skipped)");
  KV_CHECK((bits & kRspErrOrder) == 0, "row-sparse key " << k.key
                                           << ": pushed indices must be sorted and unique (row-sparse invariant, "
                                              "src/ndarray/ndarray.cc:65)");
}

void Store::check_rsp_errors() {
  for (auto& kv : rsp_) {
    RspKey& k = *kv.second;
    if (k.count.p == nullptr) continue;
    uint32_t bits = 0;
    CUDA_OK(cudaMemcpy(&bits, k.err(), 4, cudaMemcpyDeviceToHost));
    check_rsp_error(k, bits);
  }
}

}  // namespace b200kv
