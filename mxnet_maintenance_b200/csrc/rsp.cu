// Row-sparse keys of the store: ReduceRowSparse + lazy row update on push, unique + row gather on
// row_sparse_pull.  Reference: CommDevice::ReduceRowSparse (src/kvstore/comm.h:478-501),
// ElementwiseSumRsp (src/ndarray/ndarray_function.cc:59-176, .cu:104-190),
// KVStoreLocal::PullRowSparseImpl / Unique (src/kvstore/kvstore_local.h:294-314,468-512),
// UniqueImpl (src/kvstore/kvstore_utils.cu:44-110), BroadcastRowSparse (comm.h:618-674),
// SparseRetainOpForwardRspImpl (src/operator/tensor/sparse_retain-inl.h:262-323).
#include "rsp.cuh"

#include <algorithm>
#include <cmath>
#include <cstring>

#include "nccl_dl.h"
#include "store.h"

namespace b200kv {

struct RspKey {
  int key = 0;
  int64_t num_rows = 0, row_len = 0;
  int bits = 0;             // significant bits of a row id
  DevBuf table, s0, s1;     // dense [num_rows, row_len] value and optimizer state
  bool has_s0 = false, has_s1 = false;
  bool full = true;         // the table holds every row; false: compact (c_idx, c_val) holds the value
  DevBuf c_idx, c_val, c_nnr;   // compact value after an updater-less push; c_nnr: int64 on device
  int64_t c_cap = 0;
  // sort / unique scratch, grow-only
  DevBuf keys_a, keys_b, pay_a, pay_b, counts, bsum, seg_start, uniq, count;
  int64_t cap = 0;
  bool merged_valid = false;
  // multi-rank exchange buffers: this rank's rows packed contiguously, and every rank's rows
  DevBuf send_idx, send_val, recv_idx, recv_val, counts_dev;
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
    uniq.reserve(sizeof(int64_t) * static_cast<size_t>(cap));
    count.reserve(64);
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
  if (world_ > 1 && !in_process_) {
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
  k.reserve(total);
  k.merged_valid = false;
  CUDA_OK(cudaMemsetAsync(k.count.p, 0, 8, stream_));
  const bool upd = has_opt();
  if (total == 0) {
    if (!upd) {  // all inputs empty -> the stored value becomes empty (ndarray_function.cu:113-123)
      k.full = false;
      k.c_nnr.reserve(8);
      CUDA_OK(cudaMemsetAsync(k.c_nnr.p, 0, 8, stream_));
      k.merged_valid = true;
    }
    return;
  }
  const int nn = static_cast<int>(total);
  make_keys_kernel<int64_t><<<std::min((nn + 255) / 256, 148 * 8), 256, 0, stream_>>>(
      in, total, k.keys_a.as<uint32_t>(), k.pay_a.as<uint32_t>());
  ++launches_;
  const int in_b = sort_pairs(k, nn, stream_, &launches_);
  const uint32_t* keys = in_b ? k.keys_b.as<uint32_t>() : k.keys_a.as<uint32_t>();
  const uint32_t* pay = in_b ? k.pay_b.as<uint32_t>() : k.pay_a.as<uint32_t>();
  unique_heads(k, keys, nn, stream_, &launches_);

  RspUpdate up;
  std::memset(&up, 0, sizeof(up));
  const bool vec = (k.row_len % 4) == 0;   // cudaMalloc'd tables and DLPack row blocks are 16-B aligned
  const int grid = std::min(nn, sm_count(device_) * 16);
  if (!upd) {
    // no updater: the merged rows replace the stored value (kvstore_local.h:268-274)
    if (total > k.c_cap) {
      k.c_cap = total;
      k.c_idx.reserve(8 * static_cast<size_t>(total));
      k.c_val.reserve(4 * static_cast<size_t>(total) * k.row_len);
      k.c_nnr.reserve(8);
    }
    up.mode = kRspStore;
    if (vec) rsp_reduce_kernel<kRspStore, true><<<grid, 128, 0, stream_>>>(
        in, pay, k.seg_start.as<int32_t>(), k.uniq.as<int64_t>(), k.count.as<int64_t>(), k.row_len,
        k.c_val.as<float>(), up);
    else rsp_reduce_kernel<kRspStore, false><<<grid, 128, 0, stream_>>>(
        in, pay, k.seg_start.as<int32_t>(), k.uniq.as<int64_t>(), k.count.as<int64_t>(), k.row_len,
        k.c_val.as<float>(), up);
    copy_i64_kernel<<<std::min((nn + 255) / 256, 1024), 256, 0, stream_>>>(
        k.uniq.as<int64_t>(), k.c_idx.as<int64_t>(), k.count.as<int64_t>());
    CUDA_OK(cudaMemcpyAsync(k.c_nnr.p, k.count.p, 8, cudaMemcpyDeviceToDevice, stream_));
    launches_ += 2;
    k.full = false;
    k.merged_valid = true;
  } else {
    KV_CHECK(k.full, "lazy row update needs the dense table (the key was overwritten by an updater-less push)");
    KV_CHECK(opt_.kind == B200KV_OPT_SGD || opt_.kind == B200KV_OPT_ADAM || opt_.kind == B200KV_OPT_ADAGRAD,
             "row-sparse keys support lazy SGD(-momentum), Adam and AdaGrad updates");
    // the sparse AdaGrad op only exists for wd == 0 (AdagradStorageType, optimizer_op-inl.h:2497-2502)
    KV_CHECK(opt_.kind != B200KV_OPT_ADAGRAD || k.wd == 0.f, "sparse AdaGrad requires wd == 0");
    const size_t bytes = static_cast<size_t>(k.num_rows) * k.row_len * 4;
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
#define GO(MODE)                                                                                        \
  do {                                                                                                  \
    if (vec) rsp_reduce_kernel<MODE, true><<<grid, 128, 0, stream_>>>(                                   \
        in, pay, k.seg_start.as<int32_t>(), k.uniq.as<int64_t>(), k.count.as<int64_t>(), k.row_len, nullptr, up); \
    else rsp_reduce_kernel<MODE, false><<<grid, 128, 0, stream_>>>(                                      \
        in, pay, k.seg_start.as<int32_t>(), k.uniq.as<int64_t>(), k.count.as<int64_t>(), k.row_len, nullptr, up); \
  } while (0)
    if (opt_.kind == B200KV_OPT_SGD) GO(kRspSgd);
    else if (opt_.kind == B200KV_OPT_ADAM) GO(kRspAdam);
    else GO(kRspAdagrad);
#undef GO
    ++launches_;
  }
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
                               DLManagedTensor* out_val) {
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
  k.reserve(total);
  const int nn = static_cast<int>(total);
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
  const uint32_t* keys = in_b ? k.keys_b.as<uint32_t>() : k.keys_a.as<uint32_t>();
  unique_heads(k, keys, nn, stream_, &launches_);
  const bool vec = (k.row_len % 4) == 0;
  const int grid = std::min(nn, sm_count(device_) * 16);
  if (vec) rsp_gather_kernel<true><<<grid, 128, 0, stream_>>>(
      k.uniq.as<int64_t>(), k.count.as<int64_t>(), k.row_len, k.full ? 1 : 0, k.table.as<float>(), k.num_rows,
      k.c_idx.as<int64_t>(), k.c_val.as<float>(), k.c_nnr.as<int64_t>(), static_cast<int64_t*>(vi.data),
      static_cast<float*>(vv.data));
  else rsp_gather_kernel<false><<<grid, 128, 0, stream_>>>(
      k.uniq.as<int64_t>(), k.count.as<int64_t>(), k.row_len, k.full ? 1 : 0, k.table.as<float>(), k.num_rows,
      k.c_idx.as<int64_t>(), k.c_val.as<float>(), k.c_nnr.as<int64_t>(), static_cast<int64_t*>(vi.data),
      static_cast<float*>(vv.data));
  CUDA_OK(cudaGetLastError());
  ++launches_;
  // the caller needs the row count to shape its output (the reference syncs here too,
  // kvstore_utils.cu:92-96)
  int64_t nnr = 0;
  CUDA_OK(cudaMemcpyAsync(&nnr, k.count.p, 8, cudaMemcpyDeviceToHost, stream_));
  CUDA_OK(cudaStreamSynchronize(stream_));
  return nnr;
}

}  // namespace b200kv
