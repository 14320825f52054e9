"""b200kv_multi_sum_sq_static — the sum-of-squares entry point for callers that must not allocate
GPU memory while they run (a framework pool already owns the device memory, or the device-wide
synchronisation of cudaMalloc is unacceptable).  It works inside one fixed array that is part of
the library image, so here it is driven straight through the C ABI and compared with the regular,
allocating b200kv_multi_sum_sq on the same inputs.  Both run the same deterministic kernels, so
the comparison is exact (tolerance 0), on two calls in a row: the second call uses fewer and
smaller arrays, so it runs on top of whatever the first call left in the shared array.  One array
is larger than a whole persistent grid's worth of chunks (5 M elements), where the order of the
fp32 additions depends on the grid size — both entry points must size their grids alike.

This file sorts last on purpose: a `-x` run has already executed every other test before it."""
import ctypes as C

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _call(lib, name, arrays, out):
    n = len(arrays)
    ptrs = (C.c_void_p * n)(*[a.data_ptr() for a in arrays])
    sizes = (C.c_int64 * n)(*[a.numel() for a in arrays])
    dtype = 0 if arrays[0].dtype == torch.float32 else 2
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    return getattr(lib, name)(n, dtype, ptrs, sizes, C.c_void_p(out.data_ptr()), stream)


def _pair(lib, arrays):
    ref = torch.full((len(arrays),), -1.0, device="cuda")
    got = torch.full((len(arrays),), -2.0, device="cuda")
    assert _call(lib, "b200kv_multi_sum_sq", arrays, ref) == 0
    assert _call(lib, "b200kv_multi_sum_sq_static", arrays, got) == 0, lib.b200kv_last_error()
    torch.cuda.synchronize()
    return ref, got


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
def test_static_sum_sq_matches_allocating_entry_point_twice(dtype):
    from mxnet_maintenance_b200 import _lib
    lib = _lib.load()
    gen = torch.Generator(device="cuda").manual_seed(5)
    first = [torch.randn(s, device="cuda", generator=gen).to(dtype)
             for s in (1, 3, 4095, 4096, 4097, 100003, 0, 1 << 20, 777, 65536, 5_000_011)]
    second = [torch.randn(s, device="cuda", generator=gen).to(dtype) for s in (5000, 2, 8193)]
    for arrays in (first, second):
        ref, got = _pair(lib, arrays)
        assert torch.equal(ref, got), (ref, got)
        truth = torch.stack([(a.double() ** 2).sum() for a in arrays]).float()
        assert torch.allclose(got, truth, rtol=1e-3 if dtype == torch.float16 else 1e-5, atol=1e-5)


def test_static_sum_sq_rejects_sizes_beyond_its_fixed_array():
    from mxnet_maintenance_b200 import _lib
    lib = _lib.load()
    out = torch.zeros(300, device="cuda")
    many = [torch.ones(4, device="cuda") for _ in range(257)]          # > 256 arrays
    assert _call(lib, "b200kv_multi_sum_sq_static", many, out) == -1
    assert b"at most 256 arrays" in lib.b200kv_last_error()
    big = [torch.empty(4096 * 16384 + 1, device="cuda")]                # 16385 chunks
    assert _call(lib, "b200kv_multi_sum_sq_static", big, out) == -1
    assert b"at most 16384 chunks" in lib.b200kv_last_error()
    # and the array is still usable afterwards
    ok = [torch.full((10,), 2.0, device="cuda")]
    assert _call(lib, "b200kv_multi_sum_sq_static", ok, out) == 0
    torch.cuda.synchronize()
    assert out[0].item() == 40.0
