/*
 * b200kv — C ABI of the B200-native gradient-sync + update library (libb200kv.so).
 *
 * Drop-in boundary for MXNet 1.x's kvstore hot path.  Every entry point returns 0 on success and
 * -1 on failure with a thread-local message retrievable through b200kv_last_error(), the same
 * convention as the reference C API (include/mxnet/c_api_error.h:36-55, MXGetLastError in
 * src/c_api/c_api_error.cc:195-206).  C++ exceptions never cross this boundary.
 *
 * Tensors are handed over zero-copy as DLPack `DLManagedTensor*` (the struct MXNDArrayToDLPack
 * produces, include/mxnet/c_api.h:994-1038; 3rdparty/dlpack/include/dlpack/dlpack.h).  They are
 * BORROWED for the duration of the call: the library never calls `deleter` and never frees caller
 * memory.  Tensors must be compact row-major CUDA tensors on the store's device.
 *
 * Two ways to drive several GPUs: b200kv_create makes a store for ONE GPU and several processes
 * (one per GPU) form a communicator (b200kv_comm_unique_id / b200kv_comm_init); b200kv_create_group
 * makes ONE store that drives several GPUs from one process, the way the reference's 'device' and
 * 'nccl' stores do.  On one GPU a key may be pushed as several device-local copies (the reference's
 * "one value per device" lists); they are summed in the reference's CommCPU order before the
 * cross-GPU reduction.
 *
 * All work is enqueued on the stream given to b200kv_set_stream (default: the legacy default
 * stream, which orders with MXNet's blocking streams); b200kv_wait blocks until it has drained.
 */
#ifndef B200KV_H_
#define B200KV_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- DLPack (binary-identical to dlpack.h v0.2 .. v0.8 unversioned DLManagedTensor) ---------- */
#ifndef DLPACK_DLPACK_H_
#ifndef DLPACK_VERSION
typedef enum { kDLCPU = 1, kDLCUDA = 2, kDLCUDAHost = 3 } DLDeviceType;
typedef struct { int32_t device_type; int32_t device_id; } DLDevice;
typedef enum { kDLInt = 0U, kDLUInt = 1U, kDLFloat = 2U } DLDataTypeCode;
typedef struct { uint8_t code; uint8_t bits; uint16_t lanes; } DLDataType;
typedef struct {
  void* data;
  DLDevice device;
  int32_t ndim;
  DLDataType dtype;
  int64_t* shape;
  int64_t* strides;
  uint64_t byte_offset;
} DLTensor;
typedef struct DLManagedTensor {
  DLTensor dl_tensor;
  void* manager_ctx;
  void (*deleter)(struct DLManagedTensor* self);
} DLManagedTensor;
#endif
#endif

typedef void* b200kv_handle;

/* MXGetLastError (src/c_api/c_api_error.cc:195-206) */
const char* b200kv_last_error(void);
/* library version: major*10000 + minor*100 + patch */
int b200kv_version(int* out);

/* ============================ store: replaces MXKVStore* ======================================
 * Reference block: include/mxnet/c_api.h:2720-3140, impl src/c_api/c_api.cc:2534-2975,
 * semantics src/kvstore/kvstore_local.h:226-343. */

/* MXKVStoreCreate (c_api.h:2720) + KVStore::Create (src/kvstore/kvstore.cc:41-82).
 * device_id: CUDA ordinal this store drives.  rank/world_size: position among the cooperating
 * processes (MXKVStoreGetRank / MXKVStoreGetGroupSize). */
int b200kv_create(int device_id, int rank, int world_size, b200kv_handle* out);
/* The reference's own usage model: ONE process driving several GPUs — mx.kv.create('device' / 'nccl')
 * holds one NDArray per device (CommDevice::Reduce / Broadcast / InitMergeBuffer / EnableP2P,
 * src/kvstore/comm.h:503-616,678-757; KVStoreNCCL::InitNCCL = ncclCommInitAll,
 * src/kvstore/kvstore_nccl.h:475-493) and gluon.Trainer hands param.list_grad() / list_data() — one
 * array per context — to every pushpull (python/mxnet/gluon/trainer.py:392-395).
 * The handle is used with the SAME entry points as a single-GPU store: for every key the value list
 * of push / pushpull must hold (at least) one array per device of the group, outs may name any of
 * them; an array is routed by the device it lives on.  The devices must have peer access to each
 * other (NVLink / NVSwitch): the members exchange through peer memory with the same fused
 * reduce-scatter + update + all-gather kernel as the one-process-per-GPU mode, one launch per device,
 * ordered on the GPUs by flag barriers; no NCCL communicator and no IPC mapping is involved.
 * Optimizer state is sharded over the devices (state memory / ndev).  A device may be listed several
 * times ("virtual ranks": the copies of a key on that GPU are dealt to its members in list order) —
 * that runs the full multi-rank path on one GPU and is what the single-GPU tests use.
 * MXKVStoreGetRank / GetGroupSize report 0 / 1: a group is ONE worker. */
int b200kv_create_group(int ndev, const int* dev_ids, b200kv_handle* out);
/* number of members (devices) of a group handle; 1 for a single-GPU store */
int b200kv_group_size(b200kv_handle h, int* out);
/* the caller's stream on each member's device (cudaStream_t, one per member, in dev_ids order): the
 * group's work is ordered after what these streams hold at the time of a call, and they are made to
 * wait for it.  Default: the legacy default stream of each device.  (b200kv_set_stream is the
 * single-GPU form.) */
int b200kv_group_set_streams(b200kv_handle h, int n, void* const* streams);
/* b200kv_view for member `member` of a group (b200kv_view on a group handle means member 0) */
int b200kv_group_view(b200kv_handle h, int member, int key, int which, DLManagedTensor** out);
/* MXKVStoreFree (c_api.h:2740) */
int b200kv_free(b200kv_handle h);
/* stream all subsequent work is enqueued on (a cudaStream_t passed as an integer/pointer) */
int b200kv_set_stream(b200kv_handle h, void* cuda_stream);
/* block the caller until everything enqueued by this store has completed
 * (the role of NDArray::WaitToRead on the results in the reference) */
int b200kv_wait(b200kv_handle h);

/* KVStoreNCCL::InitNCCL (src/kvstore/kvstore_nccl.h:475-493).  Rank 0 obtains a 128-byte id,
 * the host side ships it to the other processes, every rank then calls b200kv_comm_init. */
int b200kv_comm_unique_id(char out_id[128]);
int b200kv_comm_init(b200kv_handle h, const char id[128]);

/* Optional: let the host side allocate the store's exchange buffers (weight segment, gradient
 * bucket, flag block) as SYMMETRIC memory — the same allocation made by every rank, mapped into
 * every rank, plus an NVLink-SHARP multicast address (e.g. torch.distributed._symmetric_memory).
 * With it the fused step uses multimem.ld_reduce / multimem.st: the NVSwitch sums the W gradient
 * copies and replicates the new weights, halving the NVLink bytes of the peer-memory path.
 * The callback is invoked collectively (every rank, same order, same sizes) from inside a store
 * call; it fills *local, peers[0..world) (peers[rank] == *local) and *multicast (NULL if the
 * fabric has no multicast) and returns 0, or non-zero to make the store fall back to its own
 * allocations.  The memory must stay valid until b200kv_free.  No counterpart in the reference
 * (its NCCL store lets NCCL own all exchange buffers, src/kvstore/kvstore_nccl.h:496-506). */
typedef int (*b200kv_symm_alloc_fn)(void* ctx, size_t bytes, void** local, void** peers, void** multicast);
int b200kv_set_symmetric_allocator(b200kv_handle h, b200kv_symm_alloc_fn fn, void* ctx);

/* KVStoreLocal::LookupKeys (kvstore_local.h:449-457): string keys -> int keys.  create != 0
 * registers unknown names (init); otherwise an unknown name is an error. Mixing int and string
 * keys in one store is an error (kvstore_local.h:322-325). */
int b200kv_lookup_keys(b200kv_handle h, int num, const char* const* str_keys, int create,
                       int* out_keys);

/* MXKVStoreInit (c_api.h:2749): copies each value into the store; duplicate key -> error
 * (kvstore_local.h:229-232).  With several processes the value of rank 0 is broadcast. */
int b200kv_init(b200kv_handle h, int num, const int* keys, DLManagedTensor* const* vals);
/* MXKVStorePush (c_api.h:2776): entries with equal keys are device copies of one key
 * (GroupKVPairs, kvstore_local.h:417-447).  Reduce, then run the optimizer on the stored value,
 * or assign the merged value when none is set (kvstore_local.h:239-276). */
int b200kv_push(b200kv_handle h, int num, const int* keys, DLManagedTensor* const* vals,
                int priority);
/* MXKVStorePullWithSparse (c_api.h:2805): copy the stored value into every out
 * (kvstore_local.h:278-292). */
int b200kv_pull(b200kv_handle h, int num, const int* keys, DLManagedTensor* const* outs,
                int priority);
/* MXKVStorePushPull (c_api.h:2947): push then pull, fused into one pass
 * (kvstore_local.h:336-343). */
int b200kv_pushpull(b200kv_handle h, int vnum, const int* vkeys, DLManagedTensor* const* vals,
                    int onum, const int* okeys, DLManagedTensor* const* outs, int priority);
/* A pushpull whose argument list is validated, grouped and planned ONCE and then replayed every
 * step with b200kv_run_bound — the arrays (MXNet's persistent gradient / weight buffers) must stay
 * alive and in place until b200kv_unbind.  Equivalent to calling b200kv_pushpull with the same
 * arguments each step; it removes the per-call handle marshalling that the reference pays per key
 * (src/c_api/c_api.cc:2707-2731 copies every NDArray handle on every call). */
int b200kv_bind_pushpull(b200kv_handle h, int vnum, const int* vkeys, DLManagedTensor* const* vals,
                         int onum, const int* okeys, DLManagedTensor* const* outs, int* out_id);
int b200kv_run_bound(b200kv_handle h, int id);
int b200kv_unbind(b200kv_handle h, int id);

/* MXKVStoreBroadcast (c_api.h:2905): init + pull (kvstore_local.h:327-334). */
int b200kv_broadcast(b200kv_handle h, int vnum, const int* vkeys, DLManagedTensor* const* vals,
                     int onum, const int* okeys, DLManagedTensor* const* outs, int priority);

/* ---- optimizer on the store: replaces MXKVStoreSetUpdaterEx + the Python updater callback ----
 * (c_api.h:2985-3027; kvstore_local.h:259-267; python/mxnet/optimizer/optimizer.py:2070-2127).
 * The host side keeps the Optimizer's scheduler / multiplier logic and ships resolved scalars,
 * exactly the attributes that reach multi_sgd_mom_update / adam_update / multi_lamb_update. */
enum { B200KV_OPT_NONE = 0, B200KV_OPT_SGD = 1, B200KV_OPT_ADAM = 2, B200KV_OPT_LAMB = 3,
       B200KV_OPT_ADAGRAD = 4 /* AdaGrad (python/mxnet/optimizer/optimizer.py:1633-1687; sparse rows:
                                 _sparse_adagrad_update, optimizer_op-inl.h:2457-2657); epsilon = eps */ };
typedef struct {
  int32_t kind;             /* B200KV_OPT_* */
  int32_t multi_precision;  /* fp16 weights keep an fp32 master copy (optimizer.py:252-304,583-600) */
  int32_t bias_correction;  /* LAMB (multi_lamb-inl.h:51-105) */
  float momentum;           /* SGD; 0 = no momentum state */
  float beta1, beta2, epsilon;
  float rescale_grad, clip_gradient;   /* clip_gradient < 0: off */
  float lower_bound, upper_bound;      /* LAMB; < 0: off */
  float clip_global_norm;              /* gluon.utils.clip_global_norm max_norm applied to the
                                          merged gradients of one push; < 0: off */
  int32_t std_update;                  /* row-sparse keys: 0 = lazy update, only the rows of the gradient
                                          (lazy_update=True, the reference's default); 1 = standard update
                                          of EVERY row, absent rows with a zero gradient (lazy_update=False:
                                          SGDMomStdDnsRspDnsKernel / AdamStdDnsRspDnsKernel,
                                          src/operator/optimizer_op.cc:108-286) */
} b200kv_optimizer;
int b200kv_set_optimizer(b200kv_handle h, const b200kv_optimizer* opt);
/* per-key scalars for the next update: learning rate (for Adam already bias-corrected, as
 * optimizer.py:1616-1619 does), weight decay, update count t (LAMB bias correction). */
int b200kv_set_hyper(b200kv_handle h, int num, const int* keys, const float* lrs, const float* wds,
                     const int* steps);
/* after a push with clip_global_norm >= 0: the total gradient norm of that push (blocks) */
int b200kv_last_global_norm(b200kv_handle h, float* out);

/* Zero-copy views INTO the store (the mirror image of the DLPack hand-over, MXNDArrayFromDLPackEx,
 * include/mxnet/c_api.h:1008-1025): which = 0 the stored weight of `key`, 1 the slot of `key` in
 * the gradient fusion bucket, 2 / 3 the whole weight / gradient segment that holds `key` (1-D).
 * A framework that makes its parameter and gradient arrays these views (mx.nd.from_dlpack) pays
 * no pack / unpack copy at all: pushpull recognises arrays that already are the store's buffers.
 * The returned DLManagedTensor is owned by the consumer (call its deleter once); the memory stays
 * valid until b200kv_free.  All keys must have been initialised before the first view is taken. */
int b200kv_view(b200kv_handle h, int key, int which, DLManagedTensor** out);

/* optimizer-state access for Trainer.save_states/load_states (python/mxnet/kvstore/kvstore.py:641-690,
 * optimizer.py:2142-2160).  which: 0 = momentum/mean, 1 = var, 2 = fp32 master weight.
 * get: copies the key's state into `out` (fp32, same element count); *present = 0 if that state
 * does not exist yet.  Collective when world_size > 1 (state is sharded across ranks). */
int b200kv_get_state(b200kv_handle h, int key, int which, DLManagedTensor* out, int* present);
int b200kv_set_state(b200kv_handle h, int key, int which, DLManagedTensor* in);

/* ---- row-sparse keys ---------------------------------------------------------------------------
 * MXKVStoreInit with a row_sparse NDArray: `dense_value` is the [num_rows, row_len] table. */
int b200kv_init_rsp(b200kv_handle h, int key, DLManagedTensor* dense_value);
/* Push of `ncopies` row-sparse gradients (indices int64 sorted unique [nnr_k], values
 * [nnr_k, row_len]; a copy with nnr 0 is skipped like an uninitialised NDArray):
 * CommDevice::ReduceRowSparse / ElementwiseSumRsp (src/kvstore/comm.h:478-501,
 * src/ndarray/ndarray_function.cc:59-176) then the lazy row update (sgd_mom_update / adam_update
 * FComputeEx, optimizer_op-inl.h:747-802,1342-1409) or assignment of the merged rows. */
int b200kv_push_rsp(b200kv_handle h, int key, int ncopies, DLManagedTensor* const* indices,
                    DLManagedTensor* const* values, int priority);
/* the merged gradient of the last b200kv_push_rsp on this key: row count, then a copy-out
 * (indices int64 [nnr], values [nnr,row_len]) for callers that need it */
int b200kv_rsp_merged_size(b200kv_handle h, int key, int64_t* nnr);
int b200kv_rsp_merged_copy(b200kv_handle h, int key, DLManagedTensor* out_indices,
                           DLManagedTensor* out_values);
/* MXKVStorePullRowSparse (c_api.h:2868; kvstore_local.h:294-314,468-512; kvstore_utils.cu:44-110;
 * sparse_retain-inl.h:262-323): row_ids may be any shape / int or float dtype, unsorted, with
 * duplicates.  out_indices (int64) and out_values must hold at least numel(row_ids) rows;
 * *out_nnr receives the number of unique ids; rows absent from the store come back as zeros.
 * The call returns as soon as the count is known; the outputs are complete in stream order (like
 * every other result of this library), not necessarily when the call returns.
 * Row ids must lie in [0, num_rows): an id outside is an error (here, or — for ids of a push, which
 * are checked on the device — at the next b200kv_row_sparse_pull / b200kv_wait). */
int b200kv_row_sparse_pull(b200kv_handle h, int key, DLManagedTensor* row_ids,
                           DLManagedTensor* out_indices, DLManagedTensor* out_values,
                           int64_t* out_nnr, int priority);

/* The same without the host synchronisation (the reference's UniqueImpl blocks on a device-to-host
 * copy of the count, src/kvstore/kvstore_utils.cu:92-96): the number of unique ids is left in
 * *out_nnr_dev (a DEVICE int64) in stream order; rows [nnr, numel(row_ids)) of the outputs are not
 * written.  Row ids outside the table are reported by the next b200kv_wait. */
int b200kv_row_sparse_pull_async(b200kv_handle h, int key, DLManagedTensor* row_ids,
                                 DLManagedTensor* out_indices, DLManagedTensor* out_values,
                                 int64_t* out_nnr_dev, int priority);

/* MXKVStoreGetType / GetRank / GetGroupSize / Barrier (c_api.h:3029-3086) */
int b200kv_get_type(b200kv_handle h, const char** out);
int b200kv_get_rank(b200kv_handle h, int* out);
int b200kv_get_group_size(b200kv_handle h, int* out);
int b200kv_barrier(b200kv_handle h);
/* CUDA-event timing of the optimizer-update kernels on the store's stream (the analogue of the
 * reference's engine-op profiler hooks, src/engine/threaded_engine.h:354-362): while enabled every
 * update section is bracketed by two events; profile_read synchronises, returns the summed
 * milliseconds and the number of sections, and resets. */
int b200kv_set_profiling(b200kv_handle h, int enable);
int b200kv_profile_read(b200kv_handle h, double* kernel_ms, int64_t* count);
/* the same, split by phase: [0] update kernels, [1] pack, [2] reduce(-scatter), [3] all-gather /
 * broadcast, [4] unpack (the reference names these engine ops "KVStoreReduce", "KVStoreBCast", …,
 * src/kvstore/comm.h:171,264; kvstore_nccl.h:316,420) */
int b200kv_profile_read_phases(b200kv_handle h, double ms[8], int64_t count[8]);
/* which mechanism carried the cross-GPU exchange of the last dense push: "local" (one GPU, nothing
 * exchanged), "peer" (fused kernel over NVLink peer memory), "peer-push", "nvls" (fused kernel over
 * the NVSwitch multicast mapping) or "nccl" (ncclReduceScatter / ncclAllGather).  Reporting only
 * (the reference exposes its choice through the store type string, src/kvstore/kvstore.cc:41-82). */
int b200kv_last_exchange(b200kv_handle h, const char** out);
/* number of kernels this store has launched so far (bench.py's gpu_launches) */
int b200kv_launch_count(b200kv_handle h, int64_t* out);

/* ============================ operators: replace the NNVM optimizer ops =======================
 * Direct, store-less forms of the fused multi-tensor ops a local Updater calls when
 * update_on_kvstore=False (python/mxnet/optimizer/optimizer.py:602-669,1316-1361).  Arrays are raw
 * device pointers; every pointer array has `num` entries; all updates are in place.
 * dtype: 0 = float32 everywhere; 2 = float16 weights/grads with float32 states and `weights32`
 * master copies (the multi_mp_* variants).  stream: cudaStream_t.
 * Scratch (tensor tables, chunk lists, reduction partials) is kept per (current device, stream):
 * calls on one stream reuse it in stream order, calls on different streams never share it, so the
 * entry points may be used concurrently from several streams / threads. */

/* multi_sgd_update / multi_sgd_mom_update / multi_mp_sgd_update / multi_mp_sgd_mom_update
 * (src/operator/optimizer_op-inl.h:207-380; registration optimizer_op.cc:322-506).
 * moms == NULL selects the momentum-free forms. */
int b200kv_multi_sgd_update(int num, int dtype, void* const* weights, const void* const* grads,
                            float* const* moms, float* const* weights32, const int64_t* sizes,
                            const float* lrs, const float* wds, float momentum, float rescale_grad,
                            float clip_gradient, void* stream);
/* sgd_update / mp_sgd_update (src/operator/optimizer_op-inl.h:382-419,656-699): the single-tensor,
 * momentum-free form (1 - lr*wd)*w - lr*clip(rescale*g) — its rounding differs from the multi-tensor
 * op's mom = 0 form, which is why it is a separate entry point.  dtype 2: fp16 weight / grad with the
 * fp32 master `weight32` (mp_sgd_update); weight32 is ignored for dtype 0. */
int b200kv_sgd_update(int dtype, void* weight, const void* grad, float* weight32, int64_t size, float lr,
                      float wd, float rescale_grad, float clip_gradient, void* stream);
/* lamb_update_phase1 / lamb_update_phase2 (src/operator/optimizer_op-inl.h:1567-1763; Python
 * optimizer.py:1275-1315), fp32: phase 1 updates mean / var and writes the update direction to g_out;
 * the caller computes r1 = ||weight||, r2 = ||g_out|| (e.g. sqrt of b200kv_multi_sum_sq) and hands
 * them to phase 2 as DEVICE scalars.  The multi-precision variants are not provided: the reference's
 * MPLambUpdatePhaseTwoKernel never writes weight32 back (:1863-1886). */
int b200kv_lamb_update_phase1(float* g_out, const float* weight, const float* grad, float* mean, float* var,
                              int64_t size, float beta1, float beta2, float epsilon, int t, int bias_correction,
                              float wd, float rescale_grad, float clip_gradient, void* stream);
int b200kv_lamb_update_phase2(float* weight, const float* g, const float* r1, const float* r2, int64_t size,
                              float lr, float lower_bound, float upper_bound, void* stream);
/* adam_update over a tensor list (optimizer_op-inl.h:1252-1340); dtype 2 is the fused form of
 * the multi-precision wrapper (optimizer.py:298-304). */
int b200kv_multi_adam_update(int num, int dtype, void* const* weights, const void* const* grads,
                             float* const* means, float* const* vars, float* const* weights32,
                             const int64_t* sizes, const float* lrs, const float* wds, float beta1,
                             float beta2, float epsilon, float rescale_grad, float clip_gradient,
                             void* stream);
/* _multi_lamb_update / _multi_mp_lamb_update (src/operator/contrib/multi_lamb-inl.h:51-340,
 * multi_lamb.cc:33-251, multi_lamb.cu:35-258); no 45-tensor limit. */
int b200kv_multi_lamb_update(int num, int dtype, void* const* weights, const void* const* grads,
                             float* const* means, float* const* vars, float* const* weights32,
                             const int64_t* sizes, const float* lrs, const float* wds,
                             const int* step_counts, float beta1, float beta2, float epsilon,
                             float rescale_grad, float clip_gradient, float lower_bound,
                             float upper_bound, int bias_correction, void* stream);
/* multi_sum_sq (src/operator/contrib/multi_sum_sq-inl.h:88-103, .cu:84-205): out[i] = sum(x_i^2),
 * fp32, deterministic (no atomics).  `out` is a device pointer to `num` floats. */
int b200kv_multi_sum_sq(int num, int dtype, const void* const* arrays, const int64_t* sizes,
                        float* out, void* stream);

/* multi_lars (src/operator/contrib/multi_lars-inl.h:60-98): layer-wise learning rates from the
 * per-tensor sums of squares that b200kv_multi_sum_sq produces.  All arrays are DEVICE arrays of
 * `num` floats. */
int b200kv_multi_lars(int num, const float* lrs, const float* weights_sum_sq, const float* grads_sum_sq,
                      const float* wds, float eta, float eps, float rescale_grad, float* out_lrs,
                      void* stream);
/* preloaded_multi_sgd_update / preloaded_multi_sgd_mom_update / preloaded_multi_mp_sgd(_mom)_update
 * (src/operator/contrib/preloaded_multi_sgd-inl.h:160-260): as b200kv_multi_sgd_update, with the
 * learning rates and weight decays read from DEVICE arrays (the outputs of multi_lars). */
int b200kv_preloaded_multi_sgd_update(int num, int dtype, void* const* weights, const void* const* grads,
                                      float* const* moms, float* const* weights32, const int64_t* sizes,
                                      const float* lrs_dev, const float* wds_dev, float momentum,
                                      float rescale_grad, float clip_gradient, void* stream);
/* _multi_adamw_update / _multi_mp_adamw_update (src/operator/contrib/adamw-inl.h:326-460,
 * adamw.cu:32-41): per-tensor lrs / wds / etas; `rescale_grad_dev` is a DEVICE pointer to the
 * loss-scale reciprocal — a non-finite or zero value skips the whole update, decided on the GPU
 * (the reference copies the scalar to the host and synchronises the stream).  No 50-tensor limit. */
int b200kv_multi_adamw_update(int num, int dtype, void* const* weights, const void* const* grads,
                              float* const* means, float* const* vars, float* const* weights32,
                              const int64_t* sizes, const float* lrs, const float* wds, const float* etas,
                              float beta1, float beta2, float epsilon, const float* rescale_grad_dev,
                              float clip_gradient, void* stream);
/* multi_all_finite (src/operator/contrib/all_finite.cc:64-95, all_finite.cu:33-93): out[0] (device
 * float) becomes 0 if any element of any array is inf / nan; it is first set to 1 when init_output. */
int b200kv_multi_all_finite(int num, int dtype, const void* const* arrays, const int64_t* sizes,
                            int init_output, float* out, void* stream);
/* reset_arrays (src/operator/contrib/reset_arrays.cc, .cu:33): zero all arrays in one launch
 * (the reference issues one cudaMemsetAsync per array). */
int b200kv_reset_arrays(int num, int dtype, void* const* arrays, const int64_t* sizes, void* stream);

/* gluon.utils.clip_global_norm (python/mxnet/gluon/utils.py:122-161) as one fused sequence:
 * total = sqrt(sum_i sum(x_i^2)); scale = min(max_norm / (total + 1e-8), 1); every array *= scale
 * in place (rounded in the array dtype).  out_scale_norm: device pointer to two floats that
 * receive {scale, total norm} — no host synchronisation (the reference blocks on .asscalar()). */
int b200kv_clip_global_norm(int num, int dtype, void* const* arrays, const int64_t* sizes,
                            float max_norm, float* out_scale_norm, void* stream);

/* b200kv_multi_sum_sq_static — no counterpart in the reference.
 * Same result as b200kv_multi_sum_sq (bit for bit), for callers that cannot allocate device memory
 * while they run: a framework whose own pool has already taken the GPU's memory (a late cudaMalloc
 * from a plugin may fail there), a trainer that cannot tolerate the device-wide synchronisation
 * cudaMalloc implies, or a process that must know a library's footprint up front.  This entry
 * point never allocates (no cudaMalloc / cudaFree / cudaMallocHost, no framework allocator): its
 * tensor table, chunk list and per-chunk partial sums live in one fixed array inside the library
 * image — every GPU that loads libb200kv.so holds those 343,296 bytes until the process exits.
 * Limits (a clear error is returned beyond them): at most B200KV_STATIC_MAX_ARRAYS arrays and
 * B200KV_STATIC_MAX_CHUNKS chunks of 4096 elements (sum over arrays of ceil(size/4096)) per call.
 * The first call on each device looks up where that GPU placed the array (a one-time runtime query);
 * make that first call before starting any stream capture.  That is all this sentence says: every
 * call also copies its tables from host memory, and no statement is made about capturing calls in
 * a CUDA graph.  All calls share the one array and it keeps whatever the
 * previous call wrote (nothing in it needs clearing between calls); two calls whose GPU work
 * overlaps — different streams, or different threads — would overwrite each other's tables, so
 * order them on one stream or with events.  `arrays`/`sizes` are host arrays; `out` is a device
 * pointer to `num` floats. */
#define B200KV_STATIC_MAX_ARRAYS 256
#define B200KV_STATIC_MAX_CHUNKS 16384
int b200kv_multi_sum_sq_static(int num, int dtype, const void* const* arrays, const int64_t* sizes,
                               float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200KV_H_ */
