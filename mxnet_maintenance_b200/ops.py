"""Store-less fused multi-tensor optimizer ops — the B200 forms of the NNVM ops a local
``Updater`` calls when ``update_on_kvstore=False``:

  multi_sgd_update / multi_sgd_mom_update / multi_mp_sgd_update / multi_mp_sgd_mom_update
      (src/operator/optimizer_op-inl.h:207-380; python wrappers python/mxnet/ndarray/contrib.py)
  adam_update over a tensor list (optimizer_op-inl.h:1252-1340), fused mixed precision
  multi_lamb_update / multi_mp_lamb_update (src/operator/contrib/multi_lamb-inl.h:51-340)
  multi_sum_sq (src/operator/contrib/multi_sum_sq-inl.h:88-103)

Arguments are lists of CUDA torch tensors updated in place (MXNet NDArrays can be passed through
``torch.from_dlpack``-style zero-copy views by the adapter).  All arithmetic happens in
libb200kv.so on the current CUDA stream.
"""
import ctypes as C

import torch

from . import _lib
from ._lib import check_call


def _pp(tensors):
    n = len(tensors)
    a = (C.c_void_p * n)()
    for i, t in enumerate(tensors):
        a[i] = t.data_ptr()
    return a


def _dtype_flag(t):
    if t.dtype == torch.float32:
        return 0
    if t.dtype == torch.float16:
        return 2
    raise TypeError("only float32 / float16 tensors are supported, got %s" % t.dtype)


def _common(weights, grads):
    assert len(weights) == len(grads) and len(weights) > 0
    for w, g in zip(weights, grads):
        assert w.is_cuda and g.is_cuda and w.is_contiguous() and g.is_contiguous()
        assert w.dtype == g.dtype and w.numel() == g.numel()
    n = len(weights)
    sizes = (C.c_int64 * n)(*[w.numel() for w in weights])
    stream = C.c_void_p(torch.cuda.current_stream(weights[0].device).cuda_stream)
    return n, sizes, stream


def _f(vals, n):
    if not isinstance(vals, (list, tuple)):
        vals = [vals] * n
    return (C.c_float * n)(*vals)


def multi_sgd_update(weights, grads, lrs, wds, moms=None, weights32=None, momentum=0.0,
                     rescale_grad=1.0, clip_gradient=-1.0):
    """moms=None: multi_(mp_)sgd_update; else multi_(mp_)sgd_mom_update.  weights32 selects mp."""
    lib = _lib.load()
    n, sizes, stream = _common(weights, grads)
    with torch.cuda.device(weights[0].device):
        check_call(lib.b200kv_multi_sgd_update(
            n, _dtype_flag(weights[0]), _pp(weights), _pp(grads),
            _pp(moms) if moms is not None else None,
            _pp(weights32) if weights32 is not None else None, sizes, _f(lrs, n), _f(wds, n),
            momentum, rescale_grad, clip_gradient if clip_gradient is not None else -1.0, stream))


def multi_adam_update(weights, grads, means, vars_, lrs, wds, weights32=None, beta1=0.9, beta2=0.999,
                      epsilon=1e-8, rescale_grad=1.0, clip_gradient=-1.0):
    lib = _lib.load()
    n, sizes, stream = _common(weights, grads)
    with torch.cuda.device(weights[0].device):
        check_call(lib.b200kv_multi_adam_update(
            n, _dtype_flag(weights[0]), _pp(weights), _pp(grads), _pp(means), _pp(vars_),
            _pp(weights32) if weights32 is not None else None, sizes, _f(lrs, n), _f(wds, n), beta1,
            beta2, epsilon, rescale_grad, clip_gradient if clip_gradient is not None else -1.0, stream))


def multi_lamb_update(weights, grads, means, vars_, lrs, wds, step_count, weights32=None, beta1=0.9,
                      beta2=0.999, epsilon=1e-6, rescale_grad=1.0, clip_gradient=-1.0,
                      lower_bound=-1.0, upper_bound=-1.0, bias_correction=True):
    lib = _lib.load()
    n, sizes, stream = _common(weights, grads)
    if not isinstance(step_count, (list, tuple)):
        step_count = [step_count] * n
    with torch.cuda.device(weights[0].device):
        check_call(lib.b200kv_multi_lamb_update(
            n, _dtype_flag(weights[0]), _pp(weights), _pp(grads), _pp(means), _pp(vars_),
            _pp(weights32) if weights32 is not None else None, sizes, _f(lrs, n), _f(wds, n),
            (C.c_int * n)(*step_count), beta1, beta2, epsilon, rescale_grad,
            clip_gradient if clip_gradient is not None else -1.0,
            lower_bound if lower_bound is not None else -1.0,
            upper_bound if upper_bound is not None else -1.0, int(bool(bias_correction)), stream))


def multi_sum_sq(arrays):
    """Returns a float32 CUDA tensor with sum(x*x) per array."""
    lib = _lib.load()
    n = len(arrays)
    for a in arrays:
        assert a.is_cuda and a.is_contiguous()
    out = torch.empty(n, dtype=torch.float32, device=arrays[0].device)
    sizes = (C.c_int64 * n)(*[a.numel() for a in arrays])
    stream = C.c_void_p(torch.cuda.current_stream(arrays[0].device).cuda_stream)
    with torch.cuda.device(arrays[0].device):
        check_call(lib.b200kv_multi_sum_sq(n, _dtype_flag(arrays[0]), _pp(arrays), sizes,
                                           C.c_void_p(out.data_ptr()), stream))
    return out


def multi_lars(lrs, weights_sum_sq, grads_sum_sq, wds, eta, eps, rescale_grad=1.0):
    """multi_lars: float32 CUDA tensors of one value per layer in, the adapted learning rates out."""
    lib = _lib.load()
    out = torch.empty_like(lrs)
    stream = C.c_void_p(torch.cuda.current_stream(lrs.device).cuda_stream)
    with torch.cuda.device(lrs.device):
        check_call(lib.b200kv_multi_lars(lrs.numel(), C.c_void_p(lrs.data_ptr()),
                                         C.c_void_p(weights_sum_sq.data_ptr()), C.c_void_p(grads_sum_sq.data_ptr()),
                                         C.c_void_p(wds.data_ptr()), eta, eps, rescale_grad,
                                         C.c_void_p(out.data_ptr()), stream))
    return out


def preloaded_multi_sgd_update(weights, grads, lrs, wds, moms=None, weights32=None, momentum=0.0,
                               rescale_grad=1.0, clip_gradient=-1.0):
    """preloaded_multi_(mp_)sgd(_mom)_update: `lrs` / `wds` are float32 CUDA tensors (one per array)."""
    lib = _lib.load()
    n, sizes, stream = _common(weights, grads)
    assert lrs.is_cuda and wds.is_cuda and lrs.numel() == n and wds.numel() == n
    with torch.cuda.device(weights[0].device):
        check_call(lib.b200kv_preloaded_multi_sgd_update(
            n, _dtype_flag(weights[0]), _pp(weights), _pp(grads), _pp(moms) if moms is not None else None,
            _pp(weights32) if weights32 is not None else None, sizes, C.c_void_p(lrs.data_ptr()),
            C.c_void_p(wds.data_ptr()), momentum, rescale_grad,
            clip_gradient if clip_gradient is not None else -1.0, stream))


def multi_adamw_update(weights, grads, means, vars_, rescale_grad, lrs, wds, etas, weights32=None,
                       beta1=0.9, beta2=0.999, epsilon=1e-8, clip_gradient=-1.0):
    """_multi_(mp_)adamw_update (python/mxnet/ndarray/contrib.py multi_adamw_update;
    src/operator/contrib/adamw-inl.h:326-460).  `rescale_grad` is a 1-element float32 CUDA tensor;
    a non-finite or zero value skips the update without any host synchronisation."""
    lib = _lib.load()
    n, sizes, stream = _common(weights, grads)
    assert rescale_grad.is_cuda and rescale_grad.dtype == torch.float32 and rescale_grad.numel() == 1
    with torch.cuda.device(weights[0].device):
        check_call(lib.b200kv_multi_adamw_update(
            n, _dtype_flag(weights[0]), _pp(weights), _pp(grads), _pp(means), _pp(vars_),
            _pp(weights32) if weights32 is not None else None, sizes, _f(lrs, n), _f(wds, n), _f(etas, n),
            beta1, beta2, epsilon, C.c_void_p(rescale_grad.data_ptr()),
            clip_gradient if clip_gradient is not None else -1.0, stream))


def multi_all_finite(arrays, out=None, init_output=True):
    """multi_all_finite: 1-element float32 CUDA tensor, 1.0 iff every value of every array is finite."""
    lib = _lib.load()
    n = len(arrays)
    if out is None:
        out = torch.empty(1, dtype=torch.float32, device=arrays[0].device)
    sizes = (C.c_int64 * n)(*[a.numel() for a in arrays])
    stream = C.c_void_p(torch.cuda.current_stream(arrays[0].device).cuda_stream)
    with torch.cuda.device(arrays[0].device):
        check_call(lib.b200kv_multi_all_finite(n, _dtype_flag(arrays[0]), _pp(arrays), sizes,
                                               int(bool(init_output)), C.c_void_p(out.data_ptr()), stream))
    return out


def reset_arrays(arrays):
    """reset_arrays: zero every array with one launch (fused zero-grad)."""
    lib = _lib.load()
    n = len(arrays)
    sizes = (C.c_int64 * n)(*[a.numel() for a in arrays])
    stream = C.c_void_p(torch.cuda.current_stream(arrays[0].device).cuda_stream)
    with torch.cuda.device(arrays[0].device):
        check_call(lib.b200kv_reset_arrays(n, _dtype_flag(arrays[0]), _pp(arrays), sizes, stream))


def clip_global_norm(arrays, max_norm):
    """gluon.utils.clip_global_norm (python/mxnet/gluon/utils.py:122-161): rescales `arrays` in
    place so that their global 2-norm is at most `max_norm`.  Returns a 2-element CUDA tensor
    {scale, total_norm}; nothing is synchronised (read it with ``.cpu()`` when needed)."""
    lib = _lib.load()
    n = len(arrays)
    for a in arrays:
        assert a.is_cuda and a.is_contiguous()
    out = torch.empty(2, dtype=torch.float32, device=arrays[0].device)
    sizes = (C.c_int64 * n)(*[a.numel() for a in arrays])
    stream = C.c_void_p(torch.cuda.current_stream(arrays[0].device).cuda_stream)
    with torch.cuda.device(arrays[0].device):
        check_call(lib.b200kv_clip_global_norm(n, _dtype_flag(arrays[0]), _pp(arrays), sizes, float(max_norm),
                                               C.c_void_p(out.data_ptr()), stream))
    return out


STATIC_MAX_ARRAYS, STATIC_MAX_CHUNKS = 256, 16384


def multi_sum_sq_static(arrays, out=None):
    """multi_sum_sq for callers that must not allocate GPU memory while they run (a framework pool
    owns the device memory, or a cudaMalloc-induced device synchronisation is unacceptable): the
    library works inside a fixed ~335 KiB array that is part of its image.  At most 256 arrays and
    16384 chunks of 4096 elements per call; pass `out` (float32, len(arrays)) to avoid even the
    result allocation.  Calls share that one array, so keep them ordered on one stream."""
    lib = _lib.load()
    n = len(arrays)
    for a in arrays:
        assert a.is_cuda and a.is_contiguous()
    if out is None:
        out = torch.empty(n, dtype=torch.float32, device=arrays[0].device)
    sizes = (C.c_int64 * n)(*[a.numel() for a in arrays])
    stream = C.c_void_p(torch.cuda.current_stream(arrays[0].device).cuda_stream)
    with torch.cuda.device(arrays[0].device):
        check_call(lib.b200kv_multi_sum_sq_static(n, _dtype_flag(arrays[0]), _pp(arrays), sizes,
                                                  C.c_void_p(out.data_ptr()), stream))
    return out


def sgd_update(weight, grad, lr, wd=0.0, rescale_grad=1.0, clip_gradient=-1.0, weight32=None):
    """sgd_update / mp_sgd_update (optimizer_op-inl.h:382-419,656-699): single tensor, no momentum,
    the (1 - lr*wd)*w form.  weight32 (fp32 master) selects the multi-precision op."""
    lib = _lib.load()
    assert weight.is_cuda and weight.is_contiguous() and grad.is_contiguous() and weight.dtype == grad.dtype
    assert (weight.dtype == torch.float16) == (weight32 is not None), "fp16 weights need weight32, fp32 must not pass it"
    stream = C.c_void_p(torch.cuda.current_stream(weight.device).cuda_stream)
    with torch.cuda.device(weight.device):
        check_call(lib.b200kv_sgd_update(_dtype_flag(weight), C.c_void_p(weight.data_ptr()), C.c_void_p(grad.data_ptr()),
                                         C.c_void_p(weight32.data_ptr()) if weight32 is not None else None,
                                         weight.numel(), lr, wd, rescale_grad, clip_gradient, stream))


def lamb_update_phase1(weight, grad, mean, var, t, wd=0.0, beta1=0.9, beta2=0.999, epsilon=1e-6,
                       bias_correction=True, rescale_grad=1.0, clip_gradient=-1.0, out=None):
    """lamb_update_phase1 (optimizer_op-inl.h:1567-1704), fp32: updates mean / var in place, returns g."""
    lib = _lib.load()
    out = torch.empty_like(weight) if out is None else out
    for x in (weight, grad, mean, var, out):
        assert x.is_cuda and x.is_contiguous() and x.dtype == torch.float32 and x.numel() == weight.numel()
    stream = C.c_void_p(torch.cuda.current_stream(weight.device).cuda_stream)
    with torch.cuda.device(weight.device):
        check_call(lib.b200kv_lamb_update_phase1(
            C.c_void_p(out.data_ptr()), C.c_void_p(weight.data_ptr()), C.c_void_p(grad.data_ptr()),
            C.c_void_p(mean.data_ptr()), C.c_void_p(var.data_ptr()), weight.numel(), beta1, beta2, epsilon, int(t),
            int(bool(bias_correction)), wd, rescale_grad, clip_gradient, stream))
    return out


def lamb_update_phase2(weight, g, r1, r2, lr, lower_bound=-1.0, upper_bound=-1.0):
    """lamb_update_phase2 (optimizer_op-inl.h:1706-1763), fp32, in place on weight; r1 / r2 are
    1-element CUDA tensors (||weight||, ||g||)."""
    lib = _lib.load()
    for x in (weight, g):
        assert x.is_cuda and x.is_contiguous() and x.dtype == torch.float32 and x.numel() == weight.numel()
    assert r1.is_cuda and r2.is_cuda and r1.dtype == torch.float32 and r2.dtype == torch.float32
    stream = C.c_void_p(torch.cuda.current_stream(weight.device).cuda_stream)
    with torch.cuda.device(weight.device):
        check_call(lib.b200kv_lamb_update_phase2(C.c_void_p(weight.data_ptr()), C.c_void_p(g.data_ptr()),
                                                 C.c_void_p(r1.data_ptr()), C.c_void_p(r2.data_ptr()), weight.numel(),
                                                 lr, lower_bound, upper_bound, stream))
