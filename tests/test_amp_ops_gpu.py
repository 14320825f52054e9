"""The mixed-precision-training helper ops next to the hot path: multi_adamw /
multi_mp_adamw with a device-side loss-scale scalar, multi_all_finite, reset_arrays.  Reference
tests: tests/python/unittest/test_contrib_optimizer.py, tests/python/gpu/test_operator_gpu.py:307-456."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

SIZES = [1, 3, 4095, 4097, 100003, 64]


def _cuda(arrs):
    return [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in arrs]


@pytest.mark.parametrize("mp", [False, True])
@pytest.mark.parametrize("clip", [-1.0, 0.004])
def test_multi_adamw_bit_exact_and_skip(oracle, mp, clip):
    from mxnet_maintenance_b200 import ops
    rng = np.random.default_rng(41)
    dt = np.float16 if mp else np.float32
    w = [rng.uniform(-1, 1, s).astype(np.float32).astype(dt) for s in SIZES]
    me = [np.zeros(s, np.float32) for s in SIZES]
    va = [np.zeros(s, np.float32) for s in SIZES]
    w32 = [x.astype(np.float32) for x in w] if mp else None
    dw, dme, dva = _cuda(w), _cuda(me), _cuda(va)
    dw32 = _cuda(w32) if mp else None
    lrs, wds, etas = [1e-3, 2e-3, 5e-4, 1e-3, 1e-3, 3e-3], [0.01, 0.0, 0.1, 0.01, 0.02, 0.0], [1.0, 0.5, 2.0, 1.0, 1.0, 0.1]
    for step, rs in enumerate([0.5, 0.25, float("inf"), 0.0, 0.5]):
        g = [(rng.uniform(-1, 1, s) * 1e-2).astype(np.float32).astype(dt) for s in SIZES]
        oracle.multi_adamw(w, g, me, va, w32, lrs, wds, etas, rescale=rs, clip=clip)
        ops.multi_adamw_update(dw, _cuda(g), dme, dva, torch.tensor([rs], device="cuda"), lrs, wds, etas,
                               weights32=dw32, clip_gradient=clip)
        for a, b in zip(dw + dme + dva, w + me + va):
            assert a.cpu().numpy().tobytes() == b.tobytes(), "step %d (rescale %s)" % (step, rs)


@pytest.mark.parametrize("dt", [torch.float32, torch.float16])
def test_multi_all_finite_and_reset_arrays(dt):
    from mxnet_maintenance_b200 import ops
    gen = torch.Generator(device="cuda").manual_seed(3)
    arrs = [torch.randn(s, device="cuda", generator=gen).to(dt) for s in SIZES]
    assert ops.multi_all_finite(arrs).item() == 1.0
    for bad in (float("inf"), float("-inf"), float("nan")):
        arrs[4][77777] = bad
        assert ops.multi_all_finite(arrs).item() == 0.0
        arrs[4][77777] = 0.0
    arrs[0][0] = float("nan")
    out = torch.ones(1, device="cuda")
    ops.multi_all_finite(arrs[1:], out=out, init_output=False)     # accumulates over several calls
    assert out.item() == 1.0
    ops.multi_all_finite(arrs[:1], out=out, init_output=False)
    assert out.item() == 0.0
    ops.multi_all_finite(arrs[1:], out=out, init_output=False)     # stays 0 without init_output
    assert out.item() == 0.0
    ops.reset_arrays(arrs)
    assert all((a == 0).all() for a in arrs)


def test_multi_lars_and_preloaded_sgd(oracle):
    """LARS recipe (test_operator_gpu.py:307-456): multi_sum_sq -> multi_lars -> preloaded_multi_sgd_mom.
    multi_lars and the preloaded update are bit-exact given the same inputs; the sums of squares
    are tree-reduced on the GPU, so the end-to-end chain is compared at rtol 1e-5 / atol 1e-6."""
    from mxnet_maintenance_b200 import ops
    rng = np.random.default_rng(51)
    n = len(SIZES)
    w = [rng.uniform(-1, 1, s).astype(np.float32) for s in SIZES]
    g = [(rng.uniform(-1, 1, s) * 1e-2).astype(np.float32) for s in SIZES]
    g[1][:] = 0                                                    # a zero-gradient layer: lr unchanged
    lrs = rng.uniform(0.01, 0.2, n).astype(np.float32)
    wds = rng.uniform(0, 1e-3, n).astype(np.float32)
    w2, g2 = oracle.multi_sum_sq(w), oracle.multi_sum_sq(g)
    ref_lrs = oracle.multi_lars(lrs, w2, g2, wds, 0.001, 1e-8, 0.5)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    got_lrs = ops.multi_lars(t(lrs), t(w2), t(g2), t(wds), 0.001, 1e-8, 0.5)
    assert got_lrs.cpu().numpy().tobytes() == ref_lrs.tobytes()
    assert ref_lrs[1] == lrs[1]
    # preloaded update, bit-exact with multi_sgd_mom_update fed the same scalars
    mom = [np.zeros(s, np.float32) for s in SIZES]
    dw, dm = _cuda(w), _cuda(mom)
    ops.preloaded_multi_sgd_update(dw, _cuda(g), got_lrs, t(wds), moms=dm, momentum=0.9, rescale_grad=0.5)
    wr = [x.copy() for x in w]
    oracle.multi_sgd(wr, g, mom, None, ref_lrs, wds, 0.9, 0.5, -1.0)
    for a, b in zip(dw + dm, wr + mom):
        assert a.cpu().numpy().tobytes() == b.tobytes()
    # whole chain on the GPU
    dw2, dm2 = _cuda(w), _cuda([np.zeros(s, np.float32) for s in SIZES])
    dg = _cuda(g)
    l2 = ops.multi_lars(t(lrs), ops.multi_sum_sq(dw2), ops.multi_sum_sq(dg), t(wds), 0.001, 1e-8, 0.5)
    ops.preloaded_multi_sgd_update(dw2, dg, l2, t(wds), moms=dm2, momentum=0.9, rescale_grad=0.5)
    for a, b in zip(dw2, wr):
        np.testing.assert_allclose(a.cpu().numpy(), b, rtol=1e-5, atol=1e-6)
