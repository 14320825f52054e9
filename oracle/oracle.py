"""TEST INFRASTRUCTURE ONLY — numpy/ctypes front-end of the CPU oracle.

Two interchangeable back-ends export the same ``mxref_*`` C symbols:

* ``Oracle("port")`` -> ``oracle/liboracle.so``: plain-C restatement (``oracle/oracle.c``); always
  available, travels to the GPU box, and is what ``cpu_baseline.kind == "port"`` times.
* ``Oracle("ref")``  -> ``oracle/_ref/libmxref.so``: the reference's OWN functors compiled from
  ``/root/reference`` in place (``oracle/ref_shim.cc``); built only where that tree exists.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import this module.  The product package never does.

Python-level sequencing restated here (cannot be lifted because ``import mxnet`` is impossible in
this image): ``Updater``/``SGD``/``Adam``/``LAMB`` scalar plumbing
(python/mxnet/optimizer/optimizer.py:411-508,602-669,1270-1372,1609-1628,2078-2127) and
``KVStoreLocal`` push/pull semantics (src/kvstore/kvstore_local.h:226-343).
"""
import ctypes as C
import math
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_LIB = os.path.join(HERE, "liboracle.so")
REF_LIB = os.path.join(HERE, "_ref", "libmxref.so")

F32, F16 = 0, 2  # mshadow type flags


def build(verbose=False):
    """Compile liboracle.so (and oracle/_ref when /root/reference is present)."""
    r = subprocess.run(["make", "-C", HERE, "all"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stdout)


def have_ref():
    return os.path.exists(REF_LIB)


def _pp(arrs):
    """list of numpy arrays (or None) -> void** """
    a = (C.c_void_p * len(arrs))()
    for i, x in enumerate(arrs):
        a[i] = None if x is None else x.ctypes.data
    return a


def _i64(v):
    return np.ascontiguousarray(np.asarray(v, dtype=np.int64))


def _f32(v):
    return np.ascontiguousarray(np.asarray(v, dtype=np.float32))


def _flag(a):
    if a.dtype == np.float32:
        return F32
    if a.dtype == np.float16:
        return F16
    raise TypeError(a.dtype)


class Oracle:
    def __init__(self, backend="port", nthread=None):
        path = PORT_LIB if backend == "port" else REF_LIB
        if not os.path.exists(path):
            if backend == "port":
                build()
            else:
                raise FileNotFoundError(path)
        self.backend = backend
        # lazy binding: the reference .cc files pulled into oracle/_ref name engine / storage functions
        # that are never called on this path (and are not linked in)
        self.lib = C.CDLL(path, mode=os.RTLD_LAZY)
        self.nthread = nthread or max(1, (os.cpu_count() or 2) // 2)  # src/engine/openmp.cc:47-53
        L = self.lib
        fp = C.c_float
        L.mxref_reduce_sum.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_int64, C.c_int64, C.c_int]
        L.mxref_multi_sgd.argtypes = [C.c_int] * 4 + [C.c_void_p] * 8 + [fp, fp, fp, C.c_int]
        L.mxref_sgd_mom.argtypes = [C.c_int64] + [C.c_void_p] * 4 + [fp] * 5
        L.mxref_mp_sgd_mom.argtypes = [C.c_int64] + [C.c_void_p] * 5 + [fp] * 5
        L.mxref_adam.argtypes = [C.c_int64] + [C.c_void_p] * 5 + [fp] * 7 + [C.c_int]
        L.mxref_multi_lamb.argtypes = ([C.c_int] * 3 + [C.c_void_p] * 10 + [fp] * 7 + [C.c_int] * 2
                                       + [C.c_void_p] * 2)
        L.mxref_multi_adamw.argtypes = ([C.c_int] * 3 + [C.c_void_p] * 10 + [fp] * 5 + [C.c_int])
        L.mxref_multi_lars.argtypes = [C.c_int] + [C.c_void_p] * 5 + [fp] * 3
        L.mxref_multi_sum_sq.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.mxref_sgd_mom_rsp.argtypes = [C.c_int64, C.c_int64] + [C.c_void_p] * 5 + [fp] * 5 + [C.c_int]
        L.mxref_sgd_mom_std_rsp.argtypes = [C.c_int64] * 3 + [C.c_void_p] * 5 + [fp] * 5 + [C.c_int]
        L.mxref_adam_std_rsp.argtypes = [C.c_int64] * 3 + [C.c_void_p] * 6 + [fp] * 7 + [C.c_int]
        L.mxref_sgd_update.argtypes = [C.c_int64] + [C.c_void_p] * 4 + [fp] * 4
        L.mxref_lamb_phase1.argtypes = [C.c_int64] + [C.c_void_p] * 5 + [fp] * 6 + [C.c_int, C.c_int]
        L.mxref_lamb_phase2.argtypes = [C.c_int64] + [C.c_void_p] * 3 + [fp] * 5
        L.mxref_sgd_rsp.argtypes = [C.c_int64, C.c_int64] + [C.c_void_p] * 4 + [fp] * 4 + [C.c_int]
        L.mxref_adam_rsp.argtypes = [C.c_int64, C.c_int64] + [C.c_void_p] * 6 + [fp] * 7 + [C.c_int]
        L.mxref_adagrad_rsp.argtypes = [C.c_int64, C.c_int64] + [C.c_void_p] * 5 + [fp] * 4 + [C.c_int]
        L.mxref_sparse_retain.argtypes = [C.c_int64] + [C.c_void_p] * 5 + [C.c_int64, C.c_int64]
        L.mxref_f32_to_f16.argtypes = [C.c_int64, C.c_void_p, C.c_void_p]
        L.mxref_f32_to_f16.restype = None
        if backend == "port":
            L.oracle_unique_i64.argtypes = [C.c_int64, C.c_void_p]
            L.oracle_unique_i64.restype = C.c_int64
            L.oracle_rsp_sum.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                         C.c_void_p, C.c_void_p, C.c_int]
            L.oracle_rsp_sum.restype = C.c_int64
            L.oracle_clip_global_norm.argtypes = [C.c_int, C.c_void_p, C.c_void_p, fp, C.c_int]
            L.oracle_clip_global_norm.restype = fp

    # ---- dense reduce (CommCPU) -------------------------------------------------------------
    def reduce_sum(self, arrays, bigarray_bound=1000 * 1000, nthread=4):
        """In-place: arrays[0] += the others, CommCPU order.  Returns arrays[0].
        Defaults: MXNET_KVSTORE_BIGARRAY_BOUND / MXNET_KVSTORE_REDUCTION_NTHREADS (comm.h:105-107)."""
        if len(arrays) == 1:      # comm.h:127-129: single source is returned untouched
            return arrays[0]
        n = arrays[0].size
        assert all(a.size == n and a.dtype == arrays[0].dtype and a.flags.c_contiguous for a in arrays)
        rc = self.lib.mxref_reduce_sum(_flag(arrays[0]), len(arrays), _pp(arrays), n, bigarray_bound,
                                       nthread)
        assert rc == 0
        return arrays[0]

    # ---- optimizer ops ----------------------------------------------------------------------
    def multi_sgd(self, weights, grads, moms, w32s, lrs, wds, momentum=0.0, rescale=1.0, clip=-1.0,
                  outs=None):
        """multi_(mp_)sgd(_mom)_update, in place on weights (or into outs), moms, w32s."""
        outs = weights if outs is None else outs
        has_mom = moms is not None
        has_mp = w32s is not None
        sizes = _i64([w.size for w in weights])
        lrs, wds = _f32(lrs), _f32(wds)   # keep the arrays alive across the call
        rc = self.lib.mxref_multi_sgd(_flag(weights[0]), int(has_mom), int(has_mp), len(weights),
                                      _pp(weights), _pp(grads), _pp(moms) if has_mom else None,
                                      _pp(w32s) if has_mp else None, _pp(outs), sizes.ctypes.data,
                                      lrs.ctypes.data, wds.ctypes.data, momentum, rescale,
                                      clip, self.nthread)
        assert rc == 0

    def sgd_mom(self, w, g, mom, lr, wd, momentum, rescale=1.0, clip=-1.0):
        self.lib.mxref_sgd_mom(w.size, w.ctypes.data, mom.ctypes.data, w.ctypes.data, g.ctypes.data,
                               clip, momentum, lr, wd, rescale)

    def mp_sgd_mom(self, w16, g16, mom, w32, lr, wd, momentum, rescale=1.0, clip=-1.0):
        self.lib.mxref_mp_sgd_mom(w16.size, w16.ctypes.data, mom.ctypes.data, w16.ctypes.data,
                                  g16.ctypes.data, w32.ctypes.data, clip, momentum, lr, wd, rescale)

    def adam(self, w, g, mean, var, lr, wd, beta1=0.9, beta2=0.999, eps=1e-8, rescale=1.0, clip=-1.0):
        """adam_update, fp32, in place.  `lr` already carries the bias correction."""
        self.lib.mxref_adam(w.size, w.ctypes.data, mean.ctypes.data, var.ctypes.data, w.ctypes.data,
                            g.ctypes.data, clip, rescale, beta1, beta2, lr, wd, eps, self.nthread)

    def mp_adam(self, w16, g16, w32, mean, var, lr, wd, **kw):
        """Optimizer.update_multi_precision generic wrapper, optimizer.py:298-304:
        grad32 = grad.astype(fp32); update(master); cast(master -> weight)."""
        g32 = g16.astype(np.float32)
        self.adam(w32, g32, mean, var, lr, wd, **kw)
        self.f32_to_f16(w32, w16)

    def multi_lamb(self, weights, grads, means, vars_, w32s, lrs, wds, steps, beta1=0.9, beta2=0.999,
                   eps=1e-6, rescale=1.0, clip=-1.0, lower=-1.0, upper=-1.0, bias_correction=True):
        """_multi_(mp_)lamb_update in place; returns (sum_sq_weights, sum_sq_update) per tensor."""
        has_mp = w32s is not None
        n = len(weights)
        sizes = _i64([w.size for w in weights])
        r1 = np.zeros(n, np.float32)
        r2 = np.zeros(n, np.float32)
        st = np.ascontiguousarray(np.asarray(steps, dtype=np.int32))
        lrs, wds = _f32(lrs), _f32(wds)   # keep the arrays alive across the call
        rc = self.lib.mxref_multi_lamb(_flag(weights[0]), int(has_mp), n, _pp(weights), _pp(grads),
                                       _pp(means), _pp(vars_), _pp(w32s) if has_mp else None,
                                       _pp(weights), sizes.ctypes.data, lrs.ctypes.data,
                                       wds.ctypes.data, st.ctypes.data, beta1, beta2, eps,
                                       rescale, clip, lower, upper, int(bias_correction),
                                       self.nthread, r1.ctypes.data, r2.ctypes.data)
        assert rc == 0
        return r1, r2

    def multi_adamw(self, weights, grads, means, vars_, w32s, lrs, wds, etas, beta1=0.9, beta2=0.999,
                    eps=1e-8, rescale=1.0, clip=-1.0):
        """_multi_(mp_)adamw_update in place (contrib/adamw-inl.h:326-460)."""
        has_mp = w32s is not None
        n = len(weights)
        sizes = _i64([w.size for w in weights])
        lrs, wds, etas = _f32(lrs), _f32(wds), _f32(etas)
        rc = self.lib.mxref_multi_adamw(_flag(weights[0]), int(has_mp), n, _pp(weights), _pp(grads),
                                        _pp(means), _pp(vars_), _pp(w32s) if has_mp else None,
                                        _pp(weights), sizes.ctypes.data, lrs.ctypes.data, wds.ctypes.data,
                                        etas.ctypes.data, beta1, beta2, eps, rescale, clip, self.nthread)
        assert rc == 0

    def multi_lars(self, lrs, w_sum_sq, g_sum_sq, wds, eta, eps, rescale=1.0):
        """multi_lars (contrib/multi_lars-inl.h:60-72): layer-wise adapted learning rates."""
        lrs, w2, g2, wds = _f32(lrs), _f32(w_sum_sq), _f32(g_sum_sq), _f32(wds)
        out = np.zeros(len(lrs), np.float32)
        self.lib.mxref_multi_lars(len(lrs), out.ctypes.data, lrs.ctypes.data, w2.ctypes.data, g2.ctypes.data,
                                  wds.ctypes.data, eta, eps, rescale)
        return out

    def multi_sum_sq(self, arrays):
        out = np.zeros(len(arrays), np.float32)
        sizes = _i64([a.size for a in arrays])
        rc = self.lib.mxref_multi_sum_sq(_flag(arrays[0]), len(arrays), _pp(arrays), sizes.ctypes.data,
                                         out.ctypes.data, self.nthread)
        assert rc == 0
        return out

    def clip_global_norm(self, arrays, max_norm):
        """gluon.utils.clip_global_norm (python/mxnet/gluon/utils.py:122-161); port back-end only."""
        sizes = _i64([a.size for a in arrays])
        return float(self.lib.oracle_clip_global_norm(len(arrays), _pp(arrays), sizes.ctypes.data,
                                                      max_norm, self.nthread))

    # ---- row sparse -------------------------------------------------------------------------
    def sgd_mom_rsp(self, w, mom, idx, g, lr, wd, momentum, rescale=1.0, clip=-1.0):
        row_len = w.shape[1]
        self.lib.mxref_sgd_mom_rsp(len(idx), row_len, w.ctypes.data, mom.ctypes.data, w.ctypes.data,
                                   idx.ctypes.data, g.ctypes.data, clip, momentum, lr, wd, rescale,
                                   self.nthread)

    def sgd_mom_std_rsp(self, w, mom, idx, g, lr, wd, momentum, rescale=1.0, clip=-1.0):
        """sgd_mom_update, row-sparse grad, lazy_update=False: every row of w (optimizer_op.cc:108-140)."""
        self.lib.mxref_sgd_mom_std_rsp(w.shape[0], len(idx), w.shape[1], w.ctypes.data, mom.ctypes.data, w.ctypes.data,
                                       idx.ctypes.data, g.ctypes.data, clip, momentum, lr, wd, rescale, self.nthread)

    def adam_std_rsp(self, w, mean, var, idx, g, lr, wd, beta1=0.9, beta2=0.999, eps=1e-8, rescale=1.0, clip=-1.0):
        """adam_update, row-sparse grad, lazy_update=False (optimizer_op.cc:196-229)."""
        self.lib.mxref_adam_std_rsp(w.shape[0], len(idx), w.shape[1], w.ctypes.data, mean.ctypes.data, var.ctypes.data,
                                    w.ctypes.data, idx.ctypes.data, g.ctypes.data, clip, beta1, beta2, lr, wd, eps,
                                    rescale, self.nthread)

    def sgd_update(self, w, g, lr, wd, rescale=1.0, clip=-1.0, w32=None):
        """sgd_update / mp_sgd_update, in place (optimizer_op-inl.h:382-399,656-679)."""
        self.lib.mxref_sgd_update(w.size, w.ctypes.data, w.ctypes.data, g.ctypes.data,
                                  None if w32 is None else w32.ctypes.data, clip, lr, wd, rescale)

    def lamb_phase1(self, w, g, mean, var, t, wd=0.0, beta1=0.9, beta2=0.999, eps=1e-6, bias_correction=True,
                    rescale=1.0, clip=-1.0):
        """lamb_update_phase1 (fp32): updates mean / var in place, returns the update direction."""
        out = np.empty_like(w)
        self.lib.mxref_lamb_phase1(w.size, out.ctypes.data, mean.ctypes.data, var.ctypes.data, w.ctypes.data,
                                   g.ctypes.data, clip, rescale, beta1, beta2, wd, eps, int(t), int(bias_correction))
        return out

    def lamb_phase2(self, w, g, r1, r2, lr, lower=-1.0, upper=-1.0):
        """lamb_update_phase2 (fp32), in place on w."""
        self.lib.mxref_lamb_phase2(w.size, w.ctypes.data, w.ctypes.data, g.ctypes.data, r1, r2, lr, lower, upper)

    def sgd_rsp(self, w, idx, g, lr, wd, rescale=1.0, clip=-1.0):
        """sgd_update with a row-sparse gradient, lazy rows, no momentum (optimizer_op-inl.h:452-475)."""
        row_len = w.shape[1]
        self.lib.mxref_sgd_rsp(len(idx), row_len, w.ctypes.data, w.ctypes.data, idx.ctypes.data, g.ctypes.data,
                               clip, lr, wd, rescale, self.nthread)

    def adam_rsp(self, w, mean, var, idx, g, lr, wd, beta1=0.9, beta2=0.999, eps=1e-8, rescale=1.0,
                 clip=-1.0):
        row_len = w.shape[1]
        self.lib.mxref_adam_rsp(len(idx), row_len, w.ctypes.data, mean.ctypes.data, var.ctypes.data,
                                w.ctypes.data, idx.ctypes.data, g.ctypes.data, clip, beta1, beta2, lr,
                                wd, eps, rescale, self.nthread)

    def adagrad_rsp(self, w, hist, idx, g, lr, eps=1e-7, rescale=1.0, clip=-1.0):
        """_sparse_adagrad_update on the rows `idx` (optimizer_op-inl.h:2514-2539)."""
        row_len = w.shape[1]
        self.lib.mxref_adagrad_rsp(len(idx), row_len, w.ctypes.data, hist.ctypes.data, w.ctypes.data,
                                   idx.ctypes.data, g.ctypes.data, clip, eps, lr, rescale, self.nthread)

    @staticmethod
    def adagrad_dense(w, hist, g, lr, wd, eps=1e-7, rescale=1.0, clip=None):
        """AdaGrad.update dense branch, a chain of fp32 NDArray ops (optimizer.py:1681-1687):
        grad = grad*rescale; clip; history += square(grad); div = grad/sqrt(history+eps);
        weight += (div + weight*wd) * -lr."""
        f = np.float32
        gr = g * f(rescale)
        if clip is not None:
            gr = np.clip(gr, f(-clip), f(clip))
        hist += gr * gr
        div = gr / np.sqrt(hist + f(eps))
        w += (div + w * f(wd)) * f(-lr)

    def sparse_retain(self, src_idx, src_vals, ids):
        """Rows `ids` (int64, sorted unique) of the row-sparse (src_idx, src_vals); absent rows = 0."""
        ids = _i64(ids)
        row_len = src_vals.shape[1]
        out = np.zeros((len(ids), row_len), np.float32)
        out_idx = np.zeros(len(ids), np.int64)
        self.lib.mxref_sparse_retain(len(ids), out.ctypes.data, out_idx.ctypes.data,
                                     src_vals.ctypes.data, src_idx.ctypes.data, ids.ctypes.data,
                                     len(src_idx), row_len)
        return out_idx, out

    def unique(self, ids):
        """KVStoreLocal::Unique / UniqueImpl<cpu> (kvstore_utils.cc:32-44).  Float-typed id arrays
        are converted by integer value first (Appendix A.9)."""
        a = _i64(np.asarray(ids).reshape(-1)).copy()
        if self.backend == "port":
            m = self.lib.oracle_unique_i64(a.size, a.ctypes.data)
            return a[:m].copy()
        return np.unique(a)

    def rsp_sum(self, inputs, nthread=None):
        """ElementwiseSumRsp (ndarray_function.cc:155-176) over [(idx int64 sorted-unique, vals [nnr,L])]."""
        assert self.backend == "port"
        k = len(inputs)
        row_len = inputs[0][1].shape[1]
        nnr = _i64([len(i) for i, _ in inputs])
        total = int(nnr.sum())
        out_idx = np.zeros(max(total, 1), np.int64)
        out_vals = np.zeros((max(total, 1), row_len), np.float32)
        idxs = [_i64(i) for i, _ in inputs]
        vals = [_f32(v) for _, v in inputs]
        nu = self.lib.oracle_rsp_sum(k, _pp(idxs), _pp(vals), nnr.ctypes.data, row_len,
                                     out_idx.ctypes.data, out_vals.ctypes.data,
                                     nthread or self.nthread)
        return out_idx[:nu].copy(), out_vals[:nu].copy()

    def f32_to_f16(self, src, dst=None):
        dst = np.empty(src.shape, np.float16) if dst is None else dst
        self.lib.mxref_f32_to_f16(src.size, src.ctypes.data, dst.ctypes.data)
        return dst


# ---------------------------------------------------------------------------------------------
# Python-side scalar plumbing of the reference optimizers (host logic, float64 like CPython).
# ---------------------------------------------------------------------------------------------
def adam_lr(lr, beta1, beta2, t):
    """Adam.update: lr *= sqrt(1-beta2^t)/(1-beta1^t) in Python doubles (optimizer.py:1616-1619)."""
    coef1 = 1.0 - beta1 ** t
    coef2 = 1.0 - beta2 ** t
    return lr * (math.sqrt(coef2) / coef1)     # `lr *= math.sqrt(coef2)/coef1`


# ---------------------------------------------------------------------------------------------
# Adjudication helpers for tensors too large for the CPU reference's sequential fp32 sums.
# ---------------------------------------------------------------------------------------------
def clip_global_norm_f64(arrays, max_norm):
    """gluon.utils.clip_global_norm (python/mxnet/gluon/utils.py:122-161) with the total norm
    accumulated in float64 and rounded to fp32 once: total = sqrt(sum_i dot(x_i, x_i));
    scale = min(max_norm / (total + 1e-8), 1); x_i *= scale (always applied, like the reference).
    Returns the norm."""
    f = np.float32
    total = f(math.sqrt(sum(float(np.dot(a.reshape(-1).astype(np.float64), a.reshape(-1).astype(np.float64)))
                            for a in arrays)))
    scale = f(max_norm) / (total + f(1e-8))
    scale = min(scale, f(1.0))
    for a in arrays:
        a *= a.dtype.type(scale)
    return float(total)


def lamb_step_f64norms(w, g, m, v, lr, wd, t, beta1=0.9, beta2=0.999, eps=1e-6, rescale=1.0,
                       bias_correction=True):
    """One MultiLAMB update of ONE fp32 tensor, in place on w, m, v: the elementwise fp32 arithmetic
    of MultiLAMBKernelStep1/2 (src/operator/contrib/multi_lamb.cc:33-118) in the same order, but with
    the two sums of squares accumulated in float64.  The CPU reference accumulates them sequentially
    in fp32 (multi_sum_sq.cc:68-84), which is off by ~1e-4 relative on tensors of 10^7 elements —
    far outside the north-star tolerance — so big tensors are judged against this instead
    (SURVEY.md Appendix A.4)."""
    f = np.float32
    b1, b2 = f(beta1), f(beta2)
    sg = g * f(rescale)
    m *= b1
    m += (f(1) - b1) * sg
    v *= b2
    v += ((f(1) - b2) * sg) * sg
    if bias_correction:
        c1 = f(1) - f(np.power(b1, f(t)))
        c2 = f(1) - f(np.power(b2, f(t)))
        u = (m / c1) / (np.sqrt(v / c2) + f(eps)) + f(wd) * w
    else:
        u = m / (np.sqrt(v) + f(eps)) + f(wd) * w
    r1 = f(math.sqrt(f(np.dot(w.astype(np.float64), w.astype(np.float64)))))
    r2 = f(math.sqrt(f(np.dot(u.astype(np.float64), u.astype(np.float64)))))
    ratio = f(1.0) if (r1 == 0 or r2 == 0) else r1 / r2
    w -= (f(lr) * ratio) * u


class LocalStoreModel:
    """KVStoreLocal + CommCPU + Updater replay for dense keys on numpy arrays
    (kvstore_local.h:226-343; optimizer.py:2078-2127).  `opt` is a dict:
      kind: 'sgd' | 'adam' | 'lamb' | None;  lr, wd, momentum, beta1, beta2, eps, rescale, clip,
      multi_precision, bias_correction, lower, upper.
    With opt None, push assigns the merged value (kvstore_local.h:268-274)."""

    def __init__(self, oracle, opt=None):
        self.o = oracle
        self.opt = opt
        self.store = {}
        self.state = {}
        self.count = {}

    def init(self, key, value):
        if key in self.store:
            raise KeyError("duplicate init of key %s" % key)   # kvstore_local.h:229-232
        self.store[key] = value.copy()

    def push(self, key, values):
        merged = self.o.reduce_sum([v.copy() for v in values]) if len(values) > 1 else values[0]
        if self.opt is None or self.opt.get("kind") is None:
            self.store[key] = merged.copy()
            return
        self._update(key, merged)

    def pull(self, key):
        return self.store[key].copy()

    def states_of(self, key):
        """{which: flat fp32 array} in the store's numbering (b200kv_get_state): 0 momentum | mean |
        AdaGrad history, 1 var, 2 fp32 master weight."""
        st, kind = self.state.get(key), (self.opt or {}).get("kind")
        if st is None:
            return {}
        if kind == "sgd":
            d = {0: st[0], 2: st[1]}
        elif kind == "adagrad":
            d = {0: st[1], 2: st[0]}
        else:
            d = {0: st[1], 1: st[2], 2: st[0]}
        return {k: v for k, v in d.items() if v is not None}

    def pushpull(self, key, values):
        self.push(key, values)
        return self.pull(key)

    def _update(self, key, grad):
        op, o = self.opt, self.o
        w = self.store[key]
        t = self.count[key] = self.count.get(key, 0) + 1
        mp = op.get("multi_precision", False) and w.dtype == np.float16
        lr = op["lr"] * op.get("lr_mult", {}).get(key, 1.0)
        wd = op.get("wd", 0.0) * op.get("wd_mult", {}).get(key, 1.0)
        rescale, clip = op.get("rescale", 1.0), op.get("clip", -1.0)
        kind = op["kind"]
        flat_w, flat_g = w.reshape(-1), np.ascontiguousarray(grad.reshape(-1))
        if kind == "sgd":
            momentum = op.get("momentum", 0.0)
            if key not in self.state:
                mom = np.zeros(w.size, np.float32) if momentum != 0.0 else None
                self.state[key] = (mom, flat_w.astype(np.float32) if mp else None)
            mom, w32 = self.state[key]
            o.multi_sgd([flat_w], [flat_g], None if mom is None else [mom],
                        None if w32 is None else [w32], [lr], [wd], momentum, rescale, clip)
        elif kind == "adam":
            b1, b2, eps = op.get("beta1", 0.9), op.get("beta2", 0.999), op.get("eps", 1e-8)
            if key not in self.state:
                self.state[key] = (flat_w.astype(np.float32) if mp else None,
                                   np.zeros(w.size, np.float32), np.zeros(w.size, np.float32))
            w32, mean, var = self.state[key]
            lr_t = adam_lr(lr, b1, b2, t)
            if mp:
                o.mp_adam(flat_w, flat_g, w32, mean, var, lr_t, wd, beta1=b1, beta2=b2, eps=eps,
                          rescale=rescale, clip=clip)
            else:
                o.adam(flat_w, flat_g, mean, var, lr_t, wd, b1, b2, eps, rescale, clip)
        elif kind == "lamb":
            b1, b2, eps = op.get("beta1", 0.9), op.get("beta2", 0.999), op.get("eps", 1e-6)
            if key not in self.state:
                self.state[key] = (flat_w.astype(np.float32) if mp else None,
                                   np.zeros(w.size, np.float32), np.zeros(w.size, np.float32))
            w32, mean, var = self.state[key]
            o.multi_lamb([flat_w], [flat_g], [mean], [var], None if w32 is None else [w32], [lr], [wd],
                         [t], b1, b2, eps, rescale, clip, op.get("lower", -1.0), op.get("upper", -1.0),
                         op.get("bias_correction", True))
        elif kind == "adagrad":
            eps = op.get("eps", 1e-7)
            if key not in self.state:
                self.state[key] = (flat_w.astype(np.float32) if mp else None, np.zeros(w.size, np.float32))
            w32, hist = self.state[key]
            c = clip if clip >= 0 else None
            if mp:   # Optimizer.update_multi_precision wrapper (optimizer.py:298-304)
                Oracle.adagrad_dense(w32, hist, flat_g.astype(np.float32), lr, wd, eps, rescale, c)
                o.f32_to_f16(w32, flat_w)
            else:
                Oracle.adagrad_dense(flat_w, hist, flat_g, lr, wd, eps, rescale, c)
        else:
            raise ValueError(kind)
