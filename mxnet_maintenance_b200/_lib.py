"""ctypes binding of libb200kv.so (C ABI declared in include/b200kv.h).

There is no CPU or pure-Python fallback: if the shared library is missing the import fails with
instructions to build it, and every failing call raises :class:`B200KVError` carrying the
library's last-error text — the counterpart of ``check_call`` / ``MXNetError`` in
python/mxnet/base.py:234-246."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200KV_LIB") or os.path.join(HERE, "libb200kv.so")   # B200KV_LIB: A/B builds


class B200KVError(RuntimeError):
    """Raised when a libb200kv entry point returns -1 (mirrors mxnet.base.MXNetError)."""


class OptimizerDesc(C.Structure):
    """struct b200kv_optimizer (include/b200kv.h)."""
    _fields_ = [("kind", C.c_int32), ("multi_precision", C.c_int32), ("bias_correction", C.c_int32),
                ("momentum", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float),
                ("epsilon", C.c_float), ("rescale_grad", C.c_float), ("clip_gradient", C.c_float),
                ("lower_bound", C.c_float), ("upper_bound", C.c_float),
                ("clip_global_norm", C.c_float), ("std_update", C.c_int32)]


OPT_NONE, OPT_SGD, OPT_ADAM, OPT_LAMB, OPT_ADAGRAD = 0, 1, 2, 3, 4

_VP, _I, _F, _I64 = C.c_void_p, C.c_int, C.c_float, C.c_int64
_PI = C.POINTER(C.c_int)
_PF = C.POINTER(C.c_float)
_PVP = C.POINTER(C.c_void_p)

# name -> argtypes; every function returns int except b200kv_last_error
SIGNATURES = {
    "b200kv_version": [_PI],
    "b200kv_create": [_I, _I, _I, _PVP],
    "b200kv_create_group": [_I, _PI, _PVP],
    "b200kv_group_size": [_VP, _PI],
    "b200kv_group_set_streams": [_VP, _I, _PVP],
    "b200kv_group_view": [_VP, _I, _I, _I, _PVP],
    "b200kv_free": [_VP],
    "b200kv_set_stream": [_VP, _VP],
    "b200kv_wait": [_VP],
    "b200kv_set_symmetric_allocator": [_VP, _VP, _VP],
    "b200kv_comm_unique_id": [C.c_char_p],
    "b200kv_comm_init": [_VP, C.c_char_p],
    "b200kv_lookup_keys": [_VP, _I, C.POINTER(C.c_char_p), _I, _PI],
    "b200kv_init": [_VP, _I, _PI, _PVP],
    "b200kv_push": [_VP, _I, _PI, _PVP, _I],
    "b200kv_pull": [_VP, _I, _PI, _PVP, _I],
    "b200kv_pushpull": [_VP, _I, _PI, _PVP, _I, _PI, _PVP, _I],
    "b200kv_broadcast": [_VP, _I, _PI, _PVP, _I, _PI, _PVP, _I],
    "b200kv_bind_pushpull": [_VP, _I, _PI, _PVP, _I, _PI, _PVP, _PI],
    "b200kv_run_bound": [_VP, _I],
    "b200kv_unbind": [_VP, _I],
    "b200kv_set_profiling": [_VP, _I],
    "b200kv_profile_read": [_VP, C.POINTER(C.c_double), C.POINTER(_I64)],
    "b200kv_profile_read_phases": [_VP, C.POINTER(C.c_double), C.POINTER(_I64)],
    "b200kv_set_optimizer": [_VP, C.POINTER(OptimizerDesc)],
    "b200kv_set_hyper": [_VP, _I, _PI, _PF, _PF, _PI],
    "b200kv_last_global_norm": [_VP, _PF],
    "b200kv_view": [_VP, _I, _I, _PVP],
    "b200kv_get_state": [_VP, _I, _I, _VP, _PI],
    "b200kv_set_state": [_VP, _I, _I, _VP],
    "b200kv_init_rsp": [_VP, _I, _VP],
    "b200kv_push_rsp": [_VP, _I, _I, _PVP, _PVP, _I],
    "b200kv_rsp_merged_size": [_VP, _I, C.POINTER(_I64)],
    "b200kv_rsp_merged_copy": [_VP, _I, _VP, _VP],
    "b200kv_row_sparse_pull": [_VP, _I, _VP, _VP, _VP, C.POINTER(_I64), _I],
    "b200kv_row_sparse_pull_async": [_VP, _I, _VP, _VP, _VP, _VP, _I],
    "b200kv_get_type": [_VP, C.POINTER(C.c_char_p)],
    "b200kv_get_rank": [_VP, _PI],
    "b200kv_get_group_size": [_VP, _PI],
    "b200kv_barrier": [_VP],
    "b200kv_last_exchange": [_VP, C.POINTER(C.c_char_p)],
    "b200kv_launch_count": [_VP, C.POINTER(_I64)],
    "b200kv_multi_sgd_update": [_I, _I, _PVP, _PVP, _PVP, _PVP, C.POINTER(_I64), _PF, _PF, _F, _F,
                                _F, _VP],
    "b200kv_sgd_update": [_I, _VP, _VP, _VP, _I64, _F, _F, _F, _F, _VP],
    "b200kv_lamb_update_phase1": [_VP, _VP, _VP, _VP, _VP, _I64, _F, _F, _F, _I, _I, _F, _F, _F, _VP],
    "b200kv_lamb_update_phase2": [_VP, _VP, _VP, _VP, _I64, _F, _F, _F, _VP],
    "b200kv_multi_adam_update": [_I, _I, _PVP, _PVP, _PVP, _PVP, _PVP, C.POINTER(_I64), _PF, _PF,
                                 _F, _F, _F, _F, _F, _VP],
    "b200kv_multi_lamb_update": [_I, _I, _PVP, _PVP, _PVP, _PVP, _PVP, C.POINTER(_I64), _PF, _PF,
                                 _PI, _F, _F, _F, _F, _F, _F, _F, _I, _VP],
    "b200kv_multi_sum_sq": [_I, _I, _PVP, C.POINTER(_I64), _VP, _VP],
    "b200kv_multi_lars": [_I, _VP, _VP, _VP, _VP, _F, _F, _F, _VP, _VP],
    "b200kv_preloaded_multi_sgd_update": [_I, _I, _PVP, _PVP, _PVP, _PVP, C.POINTER(_I64), _VP, _VP, _F, _F,
                                          _F, _VP],
    "b200kv_multi_adamw_update": [_I, _I, _PVP, _PVP, _PVP, _PVP, _PVP, C.POINTER(_I64), _PF, _PF, _PF,
                                  _F, _F, _F, _VP, _F, _VP],
    "b200kv_multi_all_finite": [_I, _I, _PVP, C.POINTER(_I64), _I, _VP, _VP],
    "b200kv_reset_arrays": [_I, _I, _PVP, C.POINTER(_I64), _VP],
    "b200kv_clip_global_norm": [_I, _I, _PVP, C.POINTER(_I64), _F, _VP, _VP],
    "b200kv_multi_sum_sq_static": [_I, _I, _PVP, C.POINTER(_I64), _VP, _VP],
}

_lib = None


def load():
    """Load libb200kv.so once; fail loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "libb200kv.so not found at %s — build it with `python -c 'import __graft_entry__ as g; "
            "g.build()'` or `make -C mxnet_maintenance_b200/csrc`. There is no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError here = header / library mismatch
        fn.argtypes = argtypes
        fn.restype = C.c_int
    lib.b200kv_last_error.argtypes = []
    lib.b200kv_last_error.restype = C.c_char_p
    _lib = lib
    return lib


def check_call(ret):
    """Raise B200KVError with the library's message when `ret` is non-zero (base.py:234-246)."""
    if ret != 0:
        raise B200KVError(load().b200kv_last_error().decode("utf-8", "replace"))
