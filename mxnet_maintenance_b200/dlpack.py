"""Zero-copy tensor hand-over: Python array object -> ``DLManagedTensor*`` for the C ABI.

Works for anything that can produce a classic ``"dltensor"`` capsule:
* ``mxnet.nd.NDArray`` — ``to_dlpack_for_write()`` / ``to_dlpack_for_read()``
  (python/mxnet/ndarray/ndarray.py:4913-4987), which also wait for pending engine work;
* ``torch.Tensor`` — ``torch.utils.dlpack.to_dlpack`` (the harness used by tests and bench.py);
* any object with ``__dlpack__``.

The capsule stays owned by Python: the library only borrows the struct for the duration of a call
and never invokes its deleter, so an un-consumed capsule frees itself normally.
"""
import ctypes as C

_PyCapsule_GetPointer = C.pythonapi.PyCapsule_GetPointer
_PyCapsule_GetPointer.restype = C.c_void_p
_PyCapsule_GetPointer.argtypes = [C.py_object, C.c_char_p]

try:
    import torch as _torch
    from torch.utils.dlpack import to_dlpack as _torch_to_dlpack
except Exception:  # pragma: no cover - torch is optional for a real MXNet host
    _torch = None


def capsule_of(arr, for_write=False):
    if _torch is not None and isinstance(arr, _torch.Tensor):
        return _torch_to_dlpack(arr)
    if for_write and hasattr(arr, "to_dlpack_for_write"):
        return arr.to_dlpack_for_write()
    if hasattr(arr, "to_dlpack_for_read"):
        return arr.to_dlpack_for_read()
    if hasattr(arr, "__dlpack__"):
        return arr.__dlpack__()
    raise TypeError("cannot export %r through DLPack" % type(arr))


def pointer_of(capsule):
    return _PyCapsule_GetPointer(capsule, b"dltensor")


class HandleCache:
    """Exports arrays for one call.  Nothing is retained between calls: a DLPack capsule owns a
    reference to the tensor (PyTorch even keeps the Python object alive through it), so caching
    capsules would pin every temporary ever passed in.  Steady-state training does not pay the
    per-call export: ``B200.bind_pushpull`` exports once and keeps the capsules for as long as the
    step stays bound."""

    def get(self, arr, for_write=False):
        cap = capsule_of(arr, for_write)
        return pointer_of(cap), cap
