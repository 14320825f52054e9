"""The B200 kvstore: a drop-in for ``mx.kv.create('device' / 'nccl')`` and
``gluon.Trainer(kvstore=...)`` registered through ``KVStoreBase.register``
(python/mxnet/kvstore/base.py:219-243), so ``mx.kv.create('b200')`` finds it.

Method names, argument meaning and error behaviour follow the native wrapper
python/mxnet/kvstore/kvstore.py:54-736 (``init/push/pull/pushpull/broadcast/row_sparse_pull/
set_optimizer/save_optimizer_states/load_optimizer_states/type/rank/num_workers``).  Every data
operation is one call into libb200kv.so with the arrays handed over zero-copy through DLPack;
nothing is computed in Python and there is no CPU fallback.

One store instance drives one GPU.  Several processes (one per GPU, e.g. under torchrun) cooperate
through NCCL: gradients are reduce-scattered, each rank updates its shard of the weights and the
optimizer state, and the new weights are all-gathered (``update_on_kvstore`` semantics preserved:
after ``pushpull`` every ``out`` holds the updated weight).
"""
import ctypes as C
import os

from . import _lib
from ._lib import B200KVError, OptimizerDesc, check_call
from ._mx_shim import KVStoreBase, HAVE_MXNET
from .dlpack import HandleCache
from . import optimizer as opt_mod

try:
    import torch
except Exception:  # pragma: no cover
    torch = None

__all__ = ["B200", "RowSparse", "B200KVError"]


class RowSparse(object):
    """Minimal row-sparse array for hosts without MXNet: ``indices`` (int64, sorted unique) and
    ``data`` ([nnr, ...row shape]) of a ``shape[0]``-row array.  An MXNet ``RowSparseNDArray`` is
    accepted wherever this class is (same attribute names, python/mxnet/ndarray/sparse.py:276-295)."""
    stype = "row_sparse"

    def __init__(self, data, indices, shape):
        self.data, self.indices, self.shape = data, indices, tuple(shape)


def _is_rsp(x):
    return getattr(x, "stype", "default") == "row_sparse"


def _is_array(x):
    return _is_rsp(x) or (torch is not None and isinstance(x, torch.Tensor)) or \
        hasattr(x, "to_dlpack_for_read") or hasattr(x, "__dlpack__")


def _empty(shape, dtype, device_id):
    """Uninitialised device array from whichever array library the host process uses: torch in the
    test / bench harness, ``mx.nd`` inside an MXNet process (dtype: 'float32' | 'int64')."""
    if torch is not None:
        return torch.empty(shape, dtype=getattr(torch, dtype), device=torch.device("cuda", device_id))
    import mxnet as mx   # pragma: no cover - only without torch
    return mx.nd.empty(shape, ctx=mx.gpu(device_id), dtype=dtype)


def _to_numpy(x):
    return x.cpu().numpy() if torch is not None and isinstance(x, torch.Tensor) else x.asnumpy()


def _from_numpy(a, device_id):
    if torch is not None:
        return torch.as_tensor(a).to(torch.device("cuda", device_id))
    import mxnet as mx   # pragma: no cover
    return mx.nd.array(a, ctx=mx.gpu(device_id), dtype=a.dtype)


def _numel(x):
    return int(x.numel()) if hasattr(x, "numel") else int(x.size)


def _dev_index(x):
    """CUDA ordinal an array lives on (torch tensor or NDArray), or None."""
    if _is_rsp(x):
        x = x.data
    if x is None:
        return None
    if torch is not None and isinstance(x, torch.Tensor):
        return x.device.index if x.is_cuda else None
    ctx = getattr(x, "context", None)
    return getattr(ctx, "device_id", None)


def _arrays_of(x):
    """Every array inside a (nested) list argument."""
    if x is None:
        return []
    if _is_array(x):
        return [x]
    out = []
    for y in x:
        out += _arrays_of(y)
    return out


def _clone(x):
    return x.clone() if hasattr(x, "clone") else x.copy()


def _flatten(keys, vals):
    """_ctype_key_value (python/mxnet/kvstore/base.py:32-64): a list under a scalar key means
    'device copies of that key'.  Returns (keys, arrays, use_str_keys)."""
    if isinstance(keys, (tuple, list)):
        assert len(keys) == len(vals)
        ks, vs, use_str = [], [], None
        for k, v in zip(keys, vals):
            k1, v1, s1 = _flatten(k, v)
            ks += k1
            vs += v1
            use_str = s1 if use_str is None else use_str
            assert use_str == s1, "inconsistent types of keys detected."
        return ks, vs, use_str
    assert isinstance(keys, (int, str)), "unexpected type for keys: " + str(type(keys))
    use_str = isinstance(keys, str)
    if _is_array(vals):
        return [keys], [vals], use_str
    for v in vals:
        assert _is_array(v)
    return [keys] * len(vals), list(vals), use_str


class BoundStep(object):
    """A replayable pushpull (see B200.bind_pushpull).  Each call ships this step's scalars and runs
    the pre-planned launch sequence.  The scalar refresh is the vectorised, value-identical form of
    ``resolve_step``: counters advance by one, the scheduler is evaluated once, and per-key
    multipliers are re-read only when they changed."""

    def __init__(self, kv, bid, keys, keep):
        self.kv, self.bid, self.keys, self._keep = kv, bid, keys, keep
        self.n = len(keys)
        self._ck = kv._keys(keys, isinstance(keys[0], str))
        self._first = True          # bind itself resolved the scalars of the first replay
        self._sig = None
        self._lr_mult = self._wd_mult = None
        self._steps = (C.c_int * self.n)()

    def _mult_signature(self, o):
        pd = o.param_dict
        return (tuple([(pd[k].lr_mult, pd[k].wd_mult) for k in self.keys if k in pd]) if pd else (),
                tuple(o.lr_mult.items()), tuple(o.wd_mult.items()), len(o.idx2name))

    def _refresh(self):
        kv = self.kv
        kv._send_optimizer()
        up = kv._updater
        if up is None:
            return
        o = up.optimizer
        import numpy as np
        sig = self._mult_signature(o)
        if sig != self._sig:
            # multipliers exactly as _get_lrs/_get_wds apply them (optimizer.py:437-494), probed
            # with unit lr / wd so that lr * mult reproduces the reference's double product
            lr0, wd0, sch = o.lr, o.wd, o.lr_scheduler
            try:
                o.lr, o.wd, o.lr_scheduler = 1.0, 1.0, None
                self._lr_mult = np.asarray(o._get_lrs(self.keys), dtype=np.float64)
                self._wd_mult = np.asarray(o._get_wds(self.keys), dtype=np.float64)
            finally:
                o.lr, o.wd, o.lr_scheduler = lr0, wd0, sch
            self._sig = sig
        if hasattr(o, "_set_current_context"):
            o._set_current_context(0)
        o._update_count(self.keys)
        cnt = o._index_update_count
        steps = [cnt[k] for k in self.keys]
        lr = o.lr_scheduler(o.num_update) if o.lr_scheduler is not None else o.lr
        lrs = lr * self._lr_mult
        wds = o.wd * self._wd_mult
        if opt_mod.kind_of(o) == "adam":
            import math
            coef = {}
            for t in set(steps):
                coef[t] = math.sqrt(1. - o.beta2 ** t) / (1. - o.beta1 ** t)
            lrs = lrs * np.asarray([coef[t] for t in steps], dtype=np.float64)
        lrs32 = np.ascontiguousarray(lrs, dtype=np.float32)
        wds32 = np.ascontiguousarray(wds, dtype=np.float32)
        self._steps[:] = steps
        check_call(kv._lib.b200kv_set_hyper(
            kv._h, self.n, self._ck, lrs32.ctypes.data_as(C.POINTER(C.c_float)),
            wds32.ctypes.data_as(C.POINTER(C.c_float)), self._steps))

    def __call__(self):
        if self._first:
            self._first = False
        else:
            self._refresh()
        check_call(self.kv._lib.b200kv_run_bound(self.kv._h, self.bid))

    def release(self):
        if self.bid:
            check_call(self.kv._lib.b200kv_unbind(self.kv._h, self.bid))
            self.bid = 0


@KVStoreBase.register
class B200(KVStoreBase):
    """``mx.kv.create('b200')``.

    Extensions over the reference API (all optional):
      * ``clip_global_norm`` attribute: when set, every push clips the merged gradients of that call
        by their global norm (gluon.utils.clip_global_norm semantics) before the update, so the
        LAMB + clipping recipe can run sharded on the store;
      * ``pushpull`` / ``push`` / ``pull`` accept the whole model in one call (lists of keys), which
        is what makes a step a handful of launches.
    """

    def __init__(self, device_id=None, rank=None, world_size=None, devices=None):
        """Three ways to come up:
          * under an initialised torch.distributed group (torchrun): one store per process / GPU,
            exchanging with its peers over NVLink (b200kv_create + communicator);
          * ``devices=[0, 1, ...]``: ONE process driving these GPUs (b200kv_create_group) — what
            ``mx.kv.create('device')`` is in the reference; a device may be listed several times
            (virtual ranks on one GPU, used by the single-GPU tests);
          * neither (``mx.kv.create('b200')`` passes no arguments): the device set is taken from the
            arrays of the first calls — ``init`` values are held back until a call names the
            per-device arrays (Trainer's ``broadcast(idx, w0, all_w)``, a ``push`` / ``pull`` with a
            list) — and a group is made when they span several GPUs.
        """
        self._lib = _lib.load()
        dist = torch is not None and torch.distributed.is_available() and torch.distributed.is_initialized()
        if rank is None:
            rank = torch.distributed.get_rank() if dist else 0
        if world_size is None:
            world_size = torch.distributed.get_world_size() if dist else 1
        self._rank, self._world = rank, world_size
        self._handle = None
        self._members = None            # device ordinal of every member of a group
        self._seen = []                 # devices named by the arrays seen so far (lazy creation)
        self._deferred = []             # (keys, cloned values) of init calls made before the handle exists
        self._deferred_keys = set()
        self._lazy = False
        if devices is not None:
            assert world_size == 1, "devices=[...] is the single-process mode"
            self._device = int(devices[0])
            self._create(list(devices))
        elif world_size > 1 or device_id is not None:
            if device_id is None:
                device_id = int(os.environ.get("LOCAL_RANK", "0"))
            self._device = device_id
            h = C.c_void_p()
            check_call(self._lib.b200kv_create(device_id, rank, world_size, C.byref(h)))
            self._handle = h
        else:
            self._lazy = True
            self._device = torch.cuda.current_device() if torch is not None and torch.cuda.is_available() else 0
        self._cache = HandleCache()
        self._updater = None
        self._opt_sent = None
        self._str_keys = None
        self._rsp_shapes = {}
        self._dense_keys = []
        self._sizes = {}
        self._shapes = {}
        self.clip_global_norm = None
        # auto_batch: gluon.Trainer issues one pushpull PER PARAMETER (trainer.py:392-395).  With
        # auto_batch the per-key calls are only recorded; once every dense key of the store has
        # been pushed the whole model runs as ONE fused launch sequence (replayed through a bound
        # step while the same arrays keep arriving).  Anything else that touches the store
        # (pull, push of a list, wait, checkpointing) flushes first.  Opt-in, because a caller that
        # pushes only SOME keys must call flush()/wait() before reading the results.
        self.auto_batch = os.environ.get("B200KV_AUTO_BATCH", "0") == "1"
        self._pending = []
        self._auto_steps = {}
        # auto_buckets > 1: the keys of a step are split into that many buckets (in the order they were
        # first pushed, balanced by size); a bucket is launched as soon as its last key has been
        # pushed, so the GPU works on early buckets while the caller is still handing over the rest
        # (the reference's priority = -index asks for exactly this order, trainer.py:392-395)
        self.auto_buckets = int(os.environ.get("B200KV_AUTO_BUCKETS", "1"))
        self._bucket_of = None          # key -> bucket, fixed after the first complete step
        self._bucket_size = []
        self._custom_updater = None     # _set_updater
        self._reducer = None
        if self._handle is not None:
            self._bind_stream()
        if world_size > 1:
            self._init_comm()

    # ------------------------------------------------------------------ plumbing
    def __del__(self):
        h, self._handle = getattr(self, "_handle", None), None
        if h:
            try:
                self._lib.b200kv_free(h)
            except Exception:
                pass

    def _create(self, devices):
        """One store for one device, a group for several (or for a device listed more than once)."""
        h = C.c_void_p()
        if len(devices) == 1:
            check_call(self._lib.b200kv_create(devices[0], 0, 1, C.byref(h)))
            self._members = None
        else:
            check_call(self._lib.b200kv_create_group(len(devices), (C.c_int * len(devices))(*devices), C.byref(h)))
            self._members = list(devices)
        self._device = devices[0]
        self._handle = h

    def _note(self, arrays):
        """Lazy mode: remember which GPUs the caller's arrays live on (until the handle exists)."""
        if self._handle is None:
            for a in arrays:
                d = _dev_index(a)
                if d is not None and d not in self._seen:
                    self._seen.append(d)

    @property
    def _h(self):
        if self._handle is None:
            self._create(self._seen or [self._device])
            self._bind_stream()
            if self._opt_sent is not None:     # set_optimizer came first (trainer.py:254-256)
                self._opt_sent = None
                self._send_optimizer()
            deferred, self._deferred = self._deferred, []
            for key, value in deferred:
                self.init(key, value)
        return self._handle

    @property
    def devices(self):
        """GPUs this store drives (one entry per member of a group)."""
        self._h
        return list(self._members) if self._members else [self._device]

    def _bind_stream(self):
        """Run on the host framework's current stream so ordering with its kernels is implicit."""
        if torch is None or not torch.cuda.is_available() or self._handle is None:
            return
        if self._members:
            n = len(self._members)
            ss = (C.c_void_p * n)(*[torch.cuda.current_stream(d).cuda_stream for d in self._members])
            check_call(self._lib.b200kv_group_set_streams(self._handle, n, ss))
        else:
            s = torch.cuda.current_stream(self._device).cuda_stream
            check_call(self._lib.b200kv_set_stream(self._handle, C.c_void_p(s)))

    def _init_comm(self):
        """KVStoreNCCL::InitNCCL (kvstore_nccl.h:475-493), one process per GPU: rank 0 creates the
        NCCL id and ships it through the already-initialised torch.distributed group."""
        buf = C.create_string_buffer(128)
        if self._rank == 0:
            check_call(self._lib.b200kv_comm_unique_id(buf))
        obj = [buf.raw]
        torch.distributed.broadcast_object_list(obj, src=0)
        check_call(self._lib.b200kv_comm_init(self._h, C.create_string_buffer(obj[0], 128)))
        self._install_symmetric_allocator()

    def _install_symmetric_allocator(self):
        """Plumbing for the NVLS path: let the store take its exchange buffers from
        torch.distributed._symmetric_memory, which maps the same allocation into every rank and binds
        it to an NVSwitch multicast object.  The library only receives raw addresses (local, peers,
        multicast); all arithmetic stays in its own kernels.  Skipped (the store then uses CUDA IPC
        peer mappings) when B200KV_NVLS=0, below 4 ranks (unless B200KV_NVLS=1) or when the torch
        build has no symmetric memory."""
        env = os.environ.get("B200KV_NVLS")
        if env == "0" or (env != "1" and self._world < 4):
            return
        try:
            import torch.distributed._symmetric_memory as symm_mem
            group = torch.distributed.group.WORLD
            gname = group.group_name
        except Exception:
            return
        dev = torch.device("cuda", self._device)
        world, keep = self._world, []
        CB = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                         C.POINTER(C.c_void_p))

        def alloc(_ctx, nbytes, local, peers, mc):
            try:
                t = symm_mem.empty(int(nbytes), dtype=torch.uint8, device=dev)
                hdl = symm_mem.rendezvous(t, gname)
                ptrs = list(hdl.buffer_ptrs)
                if len(ptrs) != world or ptrs[hdl.rank] != t.data_ptr():
                    return 1
                keep.append((t, hdl))
                local[0] = t.data_ptr()
                for k in range(world):
                    peers[k] = ptrs[k]
                mc[0] = hdl.multicast_ptr if getattr(hdl, "multicast_ptr", 0) else None
                return 0
            except Exception:
                return 1
        self._symm_cb, self._symm_keep = CB(alloc), keep
        check_call(self._lib.b200kv_set_symmetric_allocator(self._h, C.cast(self._symm_cb, C.c_void_p), None))

    def _keys(self, keys, use_str, create=False):
        """int keys for the C ABI; string keys are mapped by the store (LookupKeys)."""
        if self._str_keys is None:
            self._str_keys = use_str
        if self._str_keys != use_str:
            raise B200KVError("inconsistent key types: mixing int and string keys is not allowed")
        n = len(keys)
        if not use_str:
            return (C.c_int * n)(*keys)
        out = (C.c_int * n)()
        names = (C.c_char_p * n)(*[k.encode() for k in keys])
        check_call(self._lib.b200kv_lookup_keys(self._h, n, names, int(create), out))
        return out

    def _handles(self, arrays, for_write):
        n = len(arrays)
        ptrs = (C.c_void_p * n)()
        keep = [None] * n
        get = self._cache.get
        for i, a in enumerate(arrays):
            ptrs[i], keep[i] = get(a, for_write)
        return ptrs, keep

    # ------------------------------------------------------------------ optimizer
    def _send_optimizer(self):
        o = self._updater.optimizer if self._updater is not None else None
        std = int(o is not None and not getattr(o, "lazy_update", True))   # optimizer.py:577-580,1595-1600
        if o is None:
            desc = (0,) * 3 + (0.0,) * 6 + (-1.0, -1.0, -1.0)
        else:
            kind = opt_mod.kind_of(o)
            clip = o.clip_gradient if o.clip_gradient else -1.0     # `if self.clip_gradient:` in the reference
            gclip = self.clip_global_norm if self.clip_global_norm is not None else -1.0
            if kind == "sgd":
                desc = (_lib.OPT_SGD, int(bool(o.multi_precision)), 0,
                        o.momentum if o.momentum > 0 else 0.0, 0.0, 0.0, 0.0, o.rescale_grad, clip,
                        -1.0, -1.0, gclip)
            elif kind == "adam":
                desc = (_lib.OPT_ADAM, int(bool(o.multi_precision)), 0, 0.0, o.beta1, o.beta2,
                        o.epsilon, o.rescale_grad, clip, -1.0, -1.0, gclip)
            elif kind == "adagrad":
                desc = (_lib.OPT_ADAGRAD, int(bool(o.multi_precision)), 0, 0.0, 0.0, 0.0,
                        o.float_stable_eps, o.rescale_grad, clip, -1.0, -1.0, gclip)
            else:
                desc = (_lib.OPT_LAMB, int(bool(o.multi_precision)), int(bool(o.bias_correction)), 0.0,
                        o.beta1, o.beta2, o.epsilon, o.rescale_grad, clip,
                        o.lower_bound if o.lower_bound else -1.0,
                        o.upper_bound if o.upper_bound else -1.0, gclip)
        desc = tuple(desc) + (std,)
        if desc != self._opt_sent:
            if self._handle is None:        # lazy mode: sent when the handle comes up
                self._opt_sent = desc
                return
            d = OptimizerDesc(*desc)
            check_call(self._lib.b200kv_set_optimizer(self._h, C.byref(d)))
            self._opt_sent = desc

    def _before_update(self, keys):
        """Ship this step's scalars: the Python side of the updater callback (kvstore.py:34-41,
        optimizer.py:2078-2127) without any tensor work."""
        self._send_optimizer()
        if self._updater is None:
            return
        uniq = list(dict.fromkeys(keys))
        lrs, wds, steps = opt_mod.resolve_step(self._updater.optimizer, uniq)
        n = len(uniq)
        ck = self._keys(uniq, isinstance(uniq[0], str)) if uniq else (C.c_int * 0)()
        check_call(self._lib.b200kv_set_hyper(self._h, n, ck, (C.c_float * n)(*lrs),
                                              (C.c_float * n)(*wds), (C.c_int * n)(*steps)))

    # ------------------------------------------------------------------ KVStore API
    def init(self, key, value):
        """kvstore.py:120-176.  Row-sparse values initialise a row-sparse key."""
        keys, vals, use_str = _flatten(key, value)
        if self._handle is None and self._lazy:
            # which GPUs take part is not known yet (KVStoreLocal::InitImpl gets ONE value per key,
            # kvstore_local.h:226-237): keep a private copy, like the store's own copy on init
            self._note(vals)
            if self._str_keys is None:
                self._str_keys = use_str
            if self._str_keys != use_str:
                raise B200KVError("inconsistent key types: mixing int and string keys is not allowed")
            for k in keys:
                if k in self._deferred_keys:
                    raise B200KVError("duplicate init of key %s" % str(k))
                self._deferred_keys.add(k)
            self._deferred.append((keys, [v if _is_rsp(v) else _clone(v) for v in vals]))
            return
        ck = self._keys(keys, use_str, create=True)
        self._bind_stream()
        dense_idx = [i for i, v in enumerate(vals) if not _is_rsp(v)]
        for i, v in enumerate(vals):
            if _is_rsp(v):
                self._init_rsp(ck[i], v, keys[i])
        if dense_idx:
            dk = (C.c_int * len(dense_idx))(*[ck[i] for i in dense_idx])
            ptrs, keep = self._handles([vals[i] for i in dense_idx], False)
            check_call(self._lib.b200kv_init(self._h, len(dense_idx), dk, ptrs))
            for i in dense_idx:
                self._dense_keys.append(keys[i])
                self._sizes[keys[i]] = _numel(vals[i])
                self._shapes[keys[i]] = tuple(vals[i].shape)

    def push(self, key, value, priority=0):
        """kvstore.py:178-270."""
        self.flush()
        keys, vals, use_str = _flatten(key, value)
        self._note(vals)
        if vals and _is_rsp(vals[0]):
            return self._push_rsp(keys, vals, use_str, priority)
        if self._custom_updater is not None:
            return self._push_custom(keys, vals)
        ck = self._keys(keys, use_str)
        self._bind_stream()
        self._before_update(keys)
        ptrs, keep = self._handles(vals, False)
        check_call(self._lib.b200kv_push(self._h, len(keys), ck, ptrs, priority))

    def pull(self, key, out=None, priority=0, ignore_sparse=True):
        """kvstore.py:272-336.  Row-sparse outs are skipped when ignore_sparse (kvstore_local.h:374-385)."""
        assert out is not None
        self.flush()
        keys, outs, use_str = _flatten(key, out)
        self._note(outs)
        if ignore_sparse:
            sel = [i for i, o in enumerate(outs) if not _is_rsp(o)]
            keys, outs = [keys[i] for i in sel], [outs[i] for i in sel]
            if not keys:
                return
        elif any(_is_rsp(o) for o in outs):
            raise B200KVError("pull with row_sparse outputs is not supported; use row_sparse_pull")
        ck = self._keys(keys, use_str)
        self._bind_stream()
        ptrs, keep = self._handles(outs, True)
        check_call(self._lib.b200kv_pull(self._h, len(keys), ck, ptrs, priority))

    def flush(self):
        """Run the per-key pushpull calls recorded under auto_batch as one batched step."""
        if not self._pending:
            return
        pend, self._pending = self._pending, []
        if self._bucket_of is None and self.auto_buckets > 1 and len(pend) == len(set(self._dense_keys)):
            self._plan_buckets([e[0] for e in pend])
        self._run_pending(pend)

    def _plan_buckets(self, keys):
        """Contiguous buckets over the push order of the first complete step, balanced by elements."""
        total = float(sum(self._sizes[k] for k in keys)) or 1.0
        nb = max(1, min(self.auto_buckets, len(keys)))
        self._bucket_of, acc, b = {}, 0.0, 0
        for k in keys:
            self._bucket_of[k] = b
            acc += self._sizes[k]
            if acc >= total * (b + 1) / nb and b < nb - 1:
                b += 1
        self._bucket_size = [sum(1 for k in keys if self._bucket_of[k] == i) for i in range(nb)]

    def _run_pending(self, pend):
        # higher priority first (the reference's engine runs higher-priority pushes first)
        pend = sorted(pend, key=lambda e: -e[3]) if len(set(e[3] for e in pend)) > 1 else pend
        pend = [(k, v, o) for k, v, o, _ in pend]
        keys = [k for k, _, _ in pend]
        vals = [v for _, v, _ in pend]
        outs = [(o if o is not None else v) for _, v, o in pend]
        flat = []
        for v, o in zip(vals, outs):
            flat += list(v) if isinstance(v, (list, tuple)) else [v]
            flat += list(o) if isinstance(o, (list, tuple)) else [o]
        sig = (tuple(keys), tuple(map(id, flat)))
        ptrs = tuple(a.data_ptr() for a in flat) if torch is not None and isinstance(flat[0], torch.Tensor) else None
        hit = self._auto_steps.get(sig)
        if hit is not None and hit[1] == ptrs:
            hit[0]()
            return
        if len(self._auto_steps) >= max(4, 2 * self.auto_buckets):   # stale bound steps (arrays were replaced)
            for old in self._auto_steps.values():
                old[0].release()
            self._auto_steps.clear()
        auto, self.auto_batch = self.auto_batch, False
        try:
            step = self.bind_pushpull(keys, vals, out=outs)     # runs the first replay itself
            step()
        finally:
            self.auto_batch = auto
        self._auto_steps[sig] = (step, ptrs, flat)               # `flat` keeps the arrays (and their ids) alive

    def pushpull(self, key, value, out=None, priority=0):
        """kvstore.py:338-415: push then pull; with out=None the result lands in `value`."""
        if self._handle is None:
            self._note(_arrays_of(value))
            self._note(_arrays_of(out))
            self._h
        if self.auto_batch and isinstance(key, (int, str)) and key in self._sizes \
                and not _is_rsp(value if _is_array(value) else value[0]):
            if self._custom_updater is not None:
                pass        # falls through to the immediate path below
            else:
                if any(e[0] == key for e in self._pending):
                    self.flush()
                self._pending.append((key, value, out, priority))
                if self._bucket_of is not None and key in self._bucket_of:
                    b = self._bucket_of[key]
                    mine = [e for e in self._pending if self._bucket_of.get(e[0]) == b]
                    if len(mine) == self._bucket_size[b]:        # the bucket is complete: launch it now
                        self._pending = [e for e in self._pending if self._bucket_of.get(e[0]) != b]
                        self._run_pending(mine)
                elif len(self._pending) == len(set(self._dense_keys)):
                    self.flush()
                return
        self.flush()
        if self._custom_updater is not None:
            vkeys, vals, use_str = _flatten(key, value)
            self._push_custom(vkeys, vals)
            if out is None:
                out = value
            return self.pull(key, out=out, priority=priority)
        vkeys, vals, use_str = _flatten(key, value)
        if out is None:
            okeys, outs = vkeys, vals
        else:
            okeys, outs, use_str_o = _flatten(key, out)
            assert use_str == use_str_o
        self._note(vals)
        self._note(outs)
        cv = self._keys(vkeys, use_str)
        co = cv if okeys is vkeys else self._keys(okeys, use_str)
        self._bind_stream()
        self._before_update(vkeys)
        pv, keep_v = self._handles(vals, False)
        po, keep_o = (pv, keep_v) if outs is vals else self._handles(outs, True)
        check_call(self._lib.b200kv_pushpull(self._h, len(vkeys), cv, pv, len(okeys), co, po, priority))

    def bind_pushpull(self, key, value, out=None):
        """Returns a callable that replays ``pushpull(key, value, out)`` on the SAME arrays every
        step (MXNet's gradient and weight buffers are persistent, python/mxnet/gluon/parameter.py
        ``_init_grad``): handles are exported, validated and planned once, so a step costs one
        scalar refresh plus one C call instead of marshalling hundreds of handles."""
        vkeys, vals, use_str = _flatten(key, value)
        if out is None:
            okeys, outs = vkeys, vals
        else:
            okeys, outs, _ = _flatten(key, out)
        self._note(vals)
        self._note(outs)
        cv = self._keys(vkeys, use_str)
        co = cv if okeys is vkeys else self._keys(okeys, use_str)
        self._bind_stream()
        self._before_update(vkeys)   # the first replay is a normal step
        pv, keep_v = self._handles(vals, False)
        po, keep_o = (pv, keep_v) if outs is vals else self._handles(outs, True)
        bid = C.c_int()
        check_call(self._lib.b200kv_bind_pushpull(self._h, len(vkeys), cv, pv, len(okeys), co, po,
                                                  C.byref(bid)))
        return BoundStep(self, bid.value, list(dict.fromkeys(vkeys)), (vals, outs, keep_v, keep_o))

    def _view(self, key, which, shape=None, member=0):
        ck = self._keys([key], isinstance(key, str))
        p = C.c_void_p()
        if self._members:
            check_call(self._lib.b200kv_group_view(self._h, member, ck[0], which, C.byref(p)))
        else:
            check_call(self._lib.b200kv_view(self._h, ck[0], which, C.byref(p)))
        new = C.pythonapi.PyCapsule_New
        new.restype, new.argtypes = C.py_object, [C.c_void_p, C.c_char_p, C.c_void_p]
        t = torch.utils.dlpack.from_dlpack(new(p.value, b"dltensor", None))
        return t.view(shape) if shape is not None else t

    def _view_capsule(self, key, which):
        ck = self._keys([key], isinstance(key, str))
        p = C.c_void_p()
        check_call(self._lib.b200kv_view(self._h, ck[0], which, C.byref(p)))
        new = C.pythonapi.PyCapsule_New
        new.restype, new.argtypes = C.py_object, [C.c_void_p, C.c_char_p, C.c_void_p]
        return new(p.value, b"dltensor", None)

    def weight_view_capsule(self, key):
        """DLPack capsule of the stored weight, for ``mx.nd.from_dlpack`` (1-D; reshape after)."""
        return self._view_capsule(key, 0)

    def grad_view_capsule(self, key):
        """DLPack capsule of the key's gradient-bucket slot, for ``mx.nd.from_dlpack``."""
        return self._view_capsule(key, 1)

    def weight_view(self, key, shape=None, member=0):
        """The stored weight of `key` as a zero-copy tensor.  A framework whose parameter array IS
        this view (MXNet: ``mx.nd.from_dlpack``) needs no pull copy: pushpull updates it in place.
        `member`: which device's copy, for a store that drives several GPUs."""
        return self._view(key, 0, shape, member)

    def grad_view(self, key, shape=None, member=0):
        """The slot of `key` in the gradient fusion bucket; gradients written here need no pack copy."""
        return self._view(key, 1, shape, member)

    def segment_views(self, key, member=0):
        """(whole weight segment, whole gradient bucket) holding `key`, as flat tensors."""
        return self._view(key, 2, None, member), self._view(key, 3, None, member)

    def set_profiling(self, enable):
        check_call(self._lib.b200kv_set_profiling(self._h, int(bool(enable))))

    def profile_read(self):
        """(milliseconds spent in update kernels, number of update sections) since the last read."""
        ms, n = C.c_double(), C.c_int64()
        check_call(self._lib.b200kv_profile_read(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def profile_read_phases(self):
        """{phase: (ms, count)} since the last read; phases: update, pack, reduce, gather, unpack."""
        ms, n = (C.c_double * 8)(), (C.c_int64 * 8)()
        check_call(self._lib.b200kv_profile_read_phases(self._h, ms, n))
        names = ["update", "pack", "reduce", "gather", "unpack"]
        return {names[i]: (ms[i], n[i]) for i in range(5)}

    def broadcast(self, key, value, out, priority=0):
        """kvstore.py:98-118 / base.py:77-95: init with rank 0's value, then pull into `out`."""
        vkeys, vals, use_str = _flatten(key, value)
        okeys, outs, _ = _flatten(key, out)
        self._note(outs)
        self._note(vals)
        self._h                     # lazy mode: the out list names the devices (trainer.py:155-160)
        cv = self._keys(vkeys, use_str, create=True)
        co = self._keys(okeys, use_str)
        self._bind_stream()
        pv, keep_v = self._handles(vals, False)
        po, keep_o = self._handles(outs, True)
        check_call(self._lib.b200kv_broadcast(self._h, len(vkeys), cv, pv, len(okeys), co, po, priority))
        for k, v in zip(vkeys, vals):
            self._dense_keys.append(k)
            self._sizes[k] = _numel(v)
            self._shapes[k] = tuple(v.shape)

    # ------------------------------------------------------------------ row sparse
    def _init_rsp(self, ckey, v, key):
        """init with a row-sparse value: rows that are absent are zero rows of the table."""
        shape = tuple(v.shape)
        if torch is not None:
            dev = torch.device("cuda", self._device)
            table = torch.zeros(shape, dtype=torch.float32, device=dev)
            if v.indices is not None and _numel(v.indices) > 0:
                table[v.indices.to(dev).long()] = v.data.to(dev).reshape((-1,) + shape[1:])
        else:  # pragma: no cover - MXNet host: densify the RowSparseNDArray
            table = v.tostype("default")
        ptrs, keep = self._handles([table], False)
        check_call(self._lib.b200kv_init_rsp(self._h, ckey, ptrs[0]))
        self._rsp_shapes[key] = shape

    def init_row_sparse(self, key, table):
        """Initialise a row-sparse key from the dense [num_rows, row_len] table holding all rows
        (what `w.tostype('row_sparse')` of a freshly initialised embedding is; trainer.py:155-157)."""
        self._note([table])
        ck = self._keys([key], isinstance(key, str), create=True)
        self._bind_stream()
        ptrs, keep = self._handles([table], False)
        check_call(self._lib.b200kv_init_rsp(self._h, ck[0], ptrs[0]))
        self._rsp_shapes[key] = tuple(table.shape)

    def _push_rsp(self, keys, vals, use_str, priority):
        ck = self._keys(keys, use_str)
        self._bind_stream()
        groups = {}
        for k, c, v in zip(keys, ck, vals):
            groups.setdefault(k, (c, []))[1].append(v)
        for k, (c, vs) in groups.items():
            self._send_optimizer()
            if self._updater is not None:
                lrs, wds, steps = opt_mod.resolve_step(self._updater.optimizer, [k])
                check_call(self._lib.b200kv_set_hyper(self._h, 1, (C.c_int * 1)(c), (C.c_float * 1)(*lrs),
                                                      (C.c_float * 1)(*wds), (C.c_int * 1)(*steps)))
            pi, keep_i = self._handles([v.indices for v in vs], False)
            pd, keep_d = self._handles([v.data for v in vs], False)
            check_call(self._lib.b200kv_push_rsp(self._h, c, len(vs), pi, pd, priority))

    def rsp_merged(self, key):
        """The merged (reduced) row-sparse gradient of the last push on `key` as a RowSparse."""
        ck = self._keys([key], isinstance(key, str))
        n = C.c_int64()
        check_call(self._lib.b200kv_rsp_merged_size(self._h, ck[0], C.byref(n)))
        shape = self._rsp_shapes[key]
        idx = _empty((n.value,), "int64", self._device)
        val = _empty((n.value,) + tuple(shape[1:]), "float32", self._device)
        pi, k1 = self._handles([idx], True)
        pd, k2 = self._handles([val], True)
        check_call(self._lib.b200kv_rsp_merged_copy(self._h, ck[0], pi[0], pd[0]))
        return RowSparse(val, idx, shape)

    def row_sparse_pull(self, key, out=None, priority=0, row_ids=None, sync=True):
        """kvstore.py:417-495: pulls the rows named by `row_ids` (any shape, unsorted, duplicates
        allowed) into row-sparse `out`.  A single row_ids with a list of outs pulls once and copies.
        sync=False (extension): no host synchronisation — ``out.indices`` / ``out.data`` keep room for
        numel(row_ids) rows and ``out.nnr`` is a 1-element device array holding how many are valid."""
        assert out is not None and row_ids is not None
        if _is_array(row_ids) and not isinstance(row_ids, (list, tuple)):
            row_ids = [row_ids]
        single = _is_rsp(out)
        outs = [out] if single else list(out)
        first_out = outs[:1] if len(row_ids) == 1 and len(outs) > 1 else outs
        if isinstance(key, (list, tuple)):
            keys = list(key)
        else:
            keys = [key] * len(first_out)
        assert len(keys) == len(first_out) == len(row_ids), \
            "the number of row_ids doesn't match the number of values"
        self._note(row_ids)
        ck = self._keys(keys, isinstance(keys[0], str))
        self._bind_stream()
        for k, c, o, ids in zip(keys, ck, first_out, row_ids):
            if not _is_rsp(o):
                raise B200KVError("row_sparse_pull requires row_sparse outputs")
            shape = self._rsp_shapes[k]
            n = int(ids.numel()) if hasattr(ids, "numel") else int(ids.size)
            dev = _dev_index(ids)
            dev = self._device if dev is None else dev     # the pulled rows land where the ids live
            idx = _empty((max(n, 1),), "int64", dev)
            val = _empty((max(n, 1),) + tuple(shape[1:]), "float32", dev)
            pr, k0 = self._handles([ids], False)
            pi, k1 = self._handles([idx], True)
            pd, k2 = self._handles([val], True)
            if not sync:
                cnt = _empty((1,), "int64", dev)
                check_call(self._lib.b200kv_row_sparse_pull_async(self._h, c, pr[0], pi[0], pd[0],
                                                                  C.c_void_p(cnt.data_ptr()), priority))
                o.indices, o.data, o.shape, o.nnr = idx, val, shape, cnt
                continue
            nnr = C.c_int64()
            check_call(self._lib.b200kv_row_sparse_pull(self._h, c, pr[0], pi[0], pd[0], C.byref(nnr),
                                                        priority))
            o.indices, o.data, o.shape = idx[:nnr.value], val[:nnr.value], shape
        if first_out is not outs:
            for o in outs[1:]:
                o.indices, o.data, o.shape = outs[0].indices.copy() if not hasattr(outs[0].indices, "clone") \
                    else outs[0].indices.clone(), outs[0].data.copy() if not hasattr(outs[0].data, "clone") \
                    else outs[0].data.clone(), outs[0].shape

    # ------------------------------------------------------------------ optimizer / capabilities
    def set_optimizer(self, optimizer):
        """kvstore.py:553-600.  SGD, Adam and LAMB run on the store; anything else raises so that
        the Trainer falls back to update_on_kvstore=False."""
        opt_mod.kind_of(optimizer)
        self._updater = opt_mod.Updater(optimizer)
        self._send_optimizer()

    def _set_updater(self, updater):
        """kvstore.py:692-736 / MXKVStoreSetUpdaterEx (include/mxnet/c_api.h:2985-3027): a custom Python
        updater ``updater(key, merged, stored)`` that changes ``stored`` in place.  The reduce stays on
        the GPU (a second, optimizer-less store merges the pushed copies in CommCPU order); the callback
        then runs synchronously inside ``push`` on zero-copy views of the merged and the stored value,
        like KVStoreLocal::PushImpl does (src/kvstore/kvstore_local.h:259-267).  With several GPUs the
        callback runs once per device copy of the stored value, so it must be deterministic (the
        reference runs it once and broadcasts).  ``None`` removes it."""
        self.flush()
        self._custom_updater = updater
        if updater is not None:
            self._updater = None
            self._send_optimizer()

    def _push_custom(self, keys, vals):
        if self._reducer is None:
            if self._world > 1:
                self._reducer = B200(device_id=self._device, rank=self._rank, world_size=self._world)
            else:
                self._reducer = B200(devices=self.devices)
        r = self._reducer
        fresh = [k for k in dict.fromkeys(keys) if k not in r._sizes]
        if fresh:
            first = {}
            for k, v in zip(keys, vals):
                first.setdefault(k, v)
            r.init(fresh, [_clone(first[k]) for k in fresh])
        r.push(keys, vals)                              # no optimizer there: the merged value is stored
        nmem = len(self.devices) if self._members else 1
        for k in dict.fromkeys(keys):
            shape = self._shapes[k]
            for m in range(nmem):
                merged = r.weight_view(k, shape, member=m)
                stored = self.weight_view(k, shape, member=m)
                self._custom_updater(k, merged, stored)

    def set_gradient_compression(self, compression_params):
        """2-bit compression is lossy and useless on NVLink; refused like KVStoreNCCL does
        (src/kvstore/kvstore_nccl.h:183-186)."""
        raise B200KVError("gradient compression is not supported by the b200 kvstore")

    def is_capable(self, capability):
        if capability.lower() == KVStoreBase.OPTIMIZER:
            return True
        raise ValueError("Unknown capability: {}".format(capability))

    @property
    def type(self):
        return "b200"

    @property
    def rank(self):
        return self._rank

    @property
    def num_workers(self):
        return self._world

    def _barrier(self):
        check_call(self._lib.b200kv_barrier(self._h))

    def wait(self):
        """Block until every enqueued operation has finished (NDArray.wait_to_read on the outs)."""
        self.flush()
        check_call(self._lib.b200kv_wait(self._h))

    def launch_count(self):
        n = C.c_int64()
        check_call(self._lib.b200kv_launch_count(self._h, C.byref(n)))
        return n.value

    def last_exchange(self):
        """'local' | 'peer' | 'peer-push' | 'nvls' | 'nccl': what carried the last dense push across GPUs."""
        p = C.c_char_p()
        check_call(self._lib.b200kv_last_exchange(self._h, C.byref(p)))
        return p.value.decode()

    def last_global_norm(self):
        v = C.c_float()
        check_call(self._lib.b200kv_last_global_norm(self._h, C.byref(v)))
        return v.value

    # ------------------------------------------------------------------ checkpointing
    def _collect_states(self):
        """Gather the (possibly sharded) optimizer state back into the reference's per-index layout
        (optimizer.py:252-255,583-600,1264-1268,1602-1607): SGD `mom` | `(mom, w32)`;
        Adam/LAMB `(mean, var)` | `(w32, (mean, var))`.  Collective when num_workers > 1."""
        assert self._updater is not None
        self.flush()
        kind = opt_mod.kind_of(self._updater.optimizer)
        states = {}
        for key in dict.fromkeys(self._dense_keys):
            s0, s1, w32 = [self.get_state(key, which) for which in (0, 1, 2)]
            if s0 is None and s1 is None and w32 is None and kind != "sgd":
                continue
            if kind == "sgd":
                st = s0 if w32 is None else (s0, w32)
            elif kind == "adagrad":
                st = s0 if w32 is None else (w32, s0)
            else:
                st = (s0, s1) if w32 is None else (w32, (s0, s1))
            states[key] = st
        self.wait()
        if HAVE_MXNET:   # pragma: no cover - files interchangeable with native stores hold NDArrays
            import mxnet as mx

            def _nd(x):
                if isinstance(x, tuple):
                    return tuple(_nd(y) for y in x)
                return None if x is None else mx.nd.array(x, dtype=x.dtype)
            states = {k: _nd(v) for k, v in states.items()}
        self._updater.states = states
        return states

    def get_state(self, key, which):
        """One optimizer-state array of `key` as numpy, shaped like the weight (the reference creates
        states with zeros(weight.shape), optimizer.py:583-600,1602-1607), or None when the current
        optimizer keeps no such state.  which: 0 momentum | mean | history, 1 var, 2 fp32 master.
        Collective when num_workers > 1 (the state is sharded across ranks)."""
        ck = self._keys([key], isinstance(key, str))
        buf = _empty((self._sizes[key],), "float32", self._device)
        p, keep = self._handles([buf], True)
        present = C.c_int()
        check_call(self._lib.b200kv_get_state(self._h, ck[0], which, p[0], C.byref(present)))
        if not present.value:
            return None
        check_call(self._lib.b200kv_wait(self._h))
        return _to_numpy(buf).reshape(self._shapes[key])

    def save_optimizer_states(self, fname, dump_optimizer=False):
        """kvstore.py:641-666."""
        assert self._updater is not None, "Cannot save states for distributed training"
        self._collect_states()
        with open(fname, 'wb') as fout:
            fout.write(self._updater.get_states(dump_optimizer))

    def load_optimizer_states(self, fname):
        """kvstore.py:668-690."""
        assert self._updater is not None, "Cannot load states for distributed training"
        with open(fname, 'rb') as f:
            self._updater.set_states(f.read())
        self._opt_sent = None
        self._send_optimizer()
        kind = opt_mod.kind_of(self._updater.optimizer)
        import numpy as _np
        for key, st in self._updater.states.items():
            if kind == "sgd":
                s0, s1, w32 = (st[0], None, st[1]) if isinstance(st, tuple) else (st, None, None)
            elif kind == "adagrad":
                s0, s1, w32 = (st[1], None, st[0]) if isinstance(st, tuple) else (st, None, None)
            elif isinstance(st[1], tuple):
                w32, (s0, s1) = st
            else:
                (s0, s1), w32 = st, None
            ck = self._keys([key], isinstance(key, str))
            for which, arr in ((2, w32), (0, s0), (1, s1)):
                if arr is None:
                    continue
                host = arr.asnumpy() if hasattr(arr, "asnumpy") else _np.asarray(arr)   # NDArray or numpy
                t = _from_numpy(_np.ascontiguousarray(host.astype(_np.float32).reshape(-1)), self._device)
                p, keep = self._handles([t], False)
                check_call(self._lib.b200kv_set_state(self._h, ck[0], which, p[0]))
        self.wait()
