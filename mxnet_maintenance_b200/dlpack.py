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
import weakref

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
    """Caches one DLPack export per live torch tensor (keyed by id, validated by a weak reference
    and the current data pointer) so a steady-state step does not re-export hundreds of tensors.
    MXNet NDArrays are never cached: their getters double as engine synchronisation points."""

    def __init__(self):
        self._c = {}

    def get(self, arr, for_write=False):
        if _torch is None or not isinstance(arr, _torch.Tensor):
            cap = capsule_of(arr, for_write)
            return pointer_of(cap), cap
        key = id(arr)
        hit = self._c.get(key)
        ptr = arr.data_ptr()
        if hit is not None and hit[0]() is arr and hit[1] == ptr and hit[4] == arr.shape:
            return hit[3], hit[2]
        cap = _torch_to_dlpack(arr)
        p = pointer_of(cap)
        if len(self._c) > 65536:
            self._c.clear()
        self._c[key] = (weakref.ref(arr), ptr, cap, p, arr.shape)
        return p, cap
