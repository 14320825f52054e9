"""The C-ABI library loads on a machine without a GPU and exports every entry point that
include/b200kv.h declares (no compute calls here); error convention = 0 / -1 + last-error string."""
import ctypes as C
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "b200kv.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200kv_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported_and_bound():
    from mxnet_maintenance_b200 import _lib
    lib = _lib.load()
    names = declared_symbols()
    assert len(names) >= 35
    for n in names:
        assert hasattr(lib, n), "libb200kv.so does not export %s" % n
    # the ctypes table binds exactly the declared functions (plus last_error)
    assert set(names) == set(_lib.SIGNATURES) | {"b200kv_last_error"}


def test_every_entry_point_cites_the_reference():
    src = open(os.path.join(ROOT, "include", "b200kv.h")).read()
    for needle in ("c_api.h:2749", "c_api.h:2776", "c_api.h:2805", "c_api.h:2947", "c_api.h:2868",
                   "kvstore_local.h", "optimizer_op-inl.h:207-380", "multi_lamb-inl.h", "multi_sum_sq-inl.h",
                   "no counterpart in the reference"):
        assert needle in src, needle


def test_error_convention_without_gpu():
    from mxnet_maintenance_b200 import _lib
    lib = _lib.load()
    v = C.c_int()
    assert lib.b200kv_version(C.byref(v)) == 0 and v.value >= 100
    # null handle -> -1 and a message, never a crash / exception across the boundary
    assert lib.b200kv_wait(None) == -1
    assert b"null store handle" in lib.b200kv_last_error()
    try:
        _lib.check_call(lib.b200kv_barrier(None))
        raise AssertionError("expected B200KVError")
    except _lib.B200KVError as e:
        assert "null store handle" in str(e)


def test_optimizer_struct_layout():
    from mxnet_maintenance_b200._lib import OptimizerDesc
    assert C.sizeof(OptimizerDesc) == 13 * 4     # 12 fields of round 1 + std_update
    assert [f[0] for f in OptimizerDesc._fields_][:4] == ["kind", "multi_precision", "bias_correction", "momentum"]
