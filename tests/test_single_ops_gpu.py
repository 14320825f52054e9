"""Single-tensor operator forms next to the multi-tensor ones:
sgd_update / mp_sgd_update (SGDKernel / MP_SGDKernel, src/operator/optimizer_op-inl.h:382-419,656-699) and
lamb_update_phase1 / lamb_update_phase2 (:1567-1763), bit-exact against the reference functors."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n", [1, 3, 4097, 1_000_003])
@pytest.mark.parametrize("clip", [-1.0, 0.004])
def test_sgd_update_bit_exact(oracle, n, clip):
    from mxnet_maintenance_b200 import ops
    rng = np.random.default_rng(n)
    w = rng.uniform(-1, 1, n).astype(np.float32)
    g = (rng.uniform(-1, 1, n) * 1e-2).astype(np.float32)
    dw, dg = torch.from_numpy(w).cuda(), torch.from_numpy(g).cuda()
    for _ in range(2):
        ops.sgd_update(dw, dg, 0.1, 1e-3, 0.5, clip)
        oracle.sgd_update(w, g, 0.1, 1e-3, 0.5, clip)
    assert dw.cpu().numpy().tobytes() == w.tobytes()
    # multi-precision: fp16 weight / grad, fp32 master
    w16, g16 = w.astype(np.float16), g.astype(np.float16)
    w32 = w16.astype(np.float32)
    d16, dg16, d32 = torch.from_numpy(w16).cuda(), torch.from_numpy(g16).cuda(), torch.from_numpy(w32).cuda()
    for _ in range(2):
        ops.sgd_update(d16, dg16, 0.1, 1e-3, 0.5, clip, weight32=d32)
        oracle.sgd_update(w16, g16, 0.1, 1e-3, 0.5, clip, w32)
    assert d16.cpu().numpy().tobytes() == w16.tobytes() and d32.cpu().numpy().tobytes() == w32.tobytes()


@pytest.mark.parametrize("bias_correction", [True, False])
def test_lamb_phases_bit_exact(oracle, bias_correction):
    """The way LAMB.update drives them (python/mxnet/optimizer/optimizer.py:1296-1315): phase 1, the two
    norms, phase 2.  Given the same norms the elementwise results are bit-identical; the norms
    themselves come from multi_sum_sq (tree-reduced) and are checked against float64."""
    from mxnet_maintenance_b200 import ops
    rng = np.random.default_rng(2)
    n = 300_007
    w = rng.uniform(-1, 1, n).astype(np.float32)
    m, v = np.zeros(n, np.float32), np.zeros(n, np.float32)
    dw, dm, dv = (torch.from_numpy(x.copy()).cuda() for x in (w, m, v))
    for t in (1, 2, 3):
        g = (rng.uniform(-1, 1, n) * 1e-2).astype(np.float32)
        dg = torch.from_numpy(g).cuda()
        clip = 0.004 if t == 2 else -1.0
        du = ops.lamb_update_phase1(dw, dg, dm, dv, t, wd=0.01, bias_correction=bias_correction, rescale_grad=0.5,
                                    clip_gradient=clip)
        u = oracle.lamb_phase1(w, g, m, v, t, wd=0.01, bias_correction=bias_correction, rescale=0.5, clip=clip)
        assert du.cpu().numpy().tobytes() == u.tobytes()
        assert dm.cpu().numpy().tobytes() == m.tobytes() and dv.cpu().numpy().tobytes() == v.tobytes()
        ss = ops.multi_sum_sq([dw, du])
        norms = torch.sqrt(ss)
        truth = np.sqrt([(w.astype(np.float64) ** 2).sum(), (u.astype(np.float64) ** 2).sum()])
        np.testing.assert_allclose(norms.cpu().numpy(), truth, rtol=1e-6)
        r1, r2 = norms[0:1].contiguous(), norms[1:2].contiguous()
        ops.lamb_update_phase2(dw, du, r1, r2, 1e-3, lower_bound=0.5 if t == 3 else -1.0,
                               upper_bound=10.0 if t == 3 else -1.0)
        oracle.lamb_phase2(w, u, float(r1.item()), float(r2.item()), 1e-3, lower=0.5 if t == 3 else -1.0,
                           upper=10.0 if t == 3 else -1.0)
        assert dw.cpu().numpy().tobytes() == w.tobytes()
