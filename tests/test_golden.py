"""Golden vectors produced by the reference's own functors (tests/golden/make_golden.py, run where
/root/reference exists).  CPU: the plain-C oracle must reproduce them bit for bit.  GPU: so must
the CUDA operator entry points (fp32 elementwise ops), LAMB within the north-star tolerance."""
import os

import numpy as np
import pytest

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_vectors.npz"))
SIZES = [int(x) for x in G["sizes"]]
LRS = [0.1, 0.05, 0.2, 0.1, 0.01, 0.1]
WDS = [1e-4, 0.0, 1e-3, 1e-4, 1e-2, 1e-4]


def split(flat, dt=np.float32):
    out, o = [], 0
    for s in SIZES:
        out.append(np.ascontiguousarray(flat[o:o + s]).astype(dt))
        o += s
    return out


def adam_lr(lr, b1, b2, t):
    import math
    return lr * (math.sqrt(1.0 - b2 ** t) / (1.0 - b1 ** t))


# ---------------------------------------------------------------- CPU: oracle vs golden
@pytest.mark.parametrize("tag,dt", [("f32", np.float32), ("f16", np.float16)])
@pytest.mark.parametrize("n", [2, 3, 4, 5, 8, 9])
def test_oracle_reduce(oracle, tag, dt, n):
    src = G["reduce_%s_%d_in" % (tag, n)]
    got = oracle.reduce_sum([np.ascontiguousarray(x) for x in src.copy()])
    assert got.tobytes() == G["reduce_%s_%d_out" % (tag, n)].tobytes()


@pytest.mark.parametrize("mp", [False, True])
@pytest.mark.parametrize("clip", [-1.0, 0.002])
def test_oracle_sgd(oracle, mp, clip):
    tag = "sgd_%s_%s" % ("mp" if mp else "f32", "clip" if clip > 0 else "noclip")
    dt = np.float16 if mp else np.float32
    w = split(G[tag + "_w0"], dt)
    w32 = [x.astype(np.float32) for x in w] if mp else None
    mom = [np.zeros(s, np.float32) for s in SIZES]
    for step in range(3):
        oracle.multi_sgd(w, split(G[tag + "_g%d" % step], dt), mom, w32, LRS, WDS, 0.9, 0.25, clip)
    assert np.concatenate([x.astype(np.float32) for x in w]).tobytes() == G[tag + "_w3"].tobytes()
    assert np.concatenate(mom).tobytes() == G[tag + "_mom3"].tobytes()


@pytest.mark.parametrize("mp", [False, True])
def test_oracle_adam(oracle, mp):
    tag = "adam_%s" % ("mp" if mp else "f32")
    dt = np.float16 if mp else np.float32
    w = split(G[tag + "_w0"], dt)
    w32 = [x.astype(np.float32) for x in w]
    mean = [np.zeros(s, np.float32) for s in SIZES]
    var = [np.zeros(s, np.float32) for s in SIZES]
    for step in range(3):
        g = split(G[tag + "_g%d" % step], dt)
        lr = adam_lr(1e-3, 0.9, 0.999, step + 1)
        for k in range(len(SIZES)):
            if mp:
                oracle.mp_adam(w[k], g[k], w32[k], mean[k], var[k], lr, 0.01, rescale=0.5)
            else:
                oracle.adam(w[k], g[k], mean[k], var[k], lr, 0.01, rescale=0.5)
    assert np.concatenate([x.astype(np.float32) for x in w]).tobytes() == G[tag + "_w3"].tobytes()
    assert np.concatenate(var).tobytes() == G[tag + "_var3"].tobytes()


def test_oracle_lamb_sumsq_rsp(oracle):
    w = split(G["lamb_w0"])
    mean = [np.zeros(s, np.float32) for s in SIZES]
    var = [np.zeros(s, np.float32) for s in SIZES]
    for step in range(3):
        r1, r2 = oracle.multi_lamb(w, split(G["lamb_g%d" % step]), mean, var, None, [1e-3] * 6, [0.01] * 6,
                                   [step + 1] * 6, rescale=0.5)
    assert np.concatenate(w).tobytes() == G["lamb_w3"].tobytes()
    assert r1.tobytes() == G["lamb_r1_3"].tobytes() and r2.tobytes() == G["lamb_r2_3"].tobytes()
    assert oracle.multi_sum_sq(split(G["sumsq_in"])).tobytes() == G["sumsq_out"].tobytes()
    W, mom = G["rsp_w0"].copy(), np.zeros_like(G["rsp_w0"])
    for _ in range(2):
        oracle.sgd_mom_rsp(W, mom, G["rsp_idx"], G["rsp_g"], 0.1, 1e-3, 0.9, 0.5)
    assert W.tobytes() == G["rsp_w2"].tobytes() and mom.tobytes() == G["rsp_mom2"].tobytes()
    ri, rv = oracle.sparse_retain(G["rsp_idx"], G["rsp_g"], [0, 1, 4, 40, 41, 63])
    assert ri.tobytes() == G["retain_idx"].tobytes() and rv.tobytes() == G["retain_val"].tobytes()


# ---------------------------------------------------------------- GPU: CUDA ops vs golden
def _cuda(arrs):
    import torch
    return [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in arrs]


def _flat(ts):
    return np.concatenate([t.float().cpu().numpy() for t in ts])


@pytest.mark.gpu
@pytest.mark.parametrize("mp", [False, True])
@pytest.mark.parametrize("clip", [-1.0, 0.002])
def test_cuda_sgd_golden(mp, clip):
    from mxnet_maintenance_b200 import ops
    tag = "sgd_%s_%s" % ("mp" if mp else "f32", "clip" if clip > 0 else "noclip")
    dt = np.float16 if mp else np.float32
    w = _cuda(split(G[tag + "_w0"], dt))
    w32 = [x.float() for x in w] if mp else None
    mom = _cuda([np.zeros(s, np.float32) for s in SIZES])
    for step in range(3):
        ops.multi_sgd_update(w, _cuda(split(G[tag + "_g%d" % step], dt)), LRS, WDS, moms=mom, weights32=w32,
                             momentum=0.9, rescale_grad=0.25, clip_gradient=clip)
    assert _flat(w).tobytes() == G[tag + "_w3"].tobytes()
    assert _flat(mom).tobytes() == G[tag + "_mom3"].tobytes()


@pytest.mark.gpu
@pytest.mark.parametrize("mp", [False, True])
def test_cuda_adam_golden(mp):
    from mxnet_maintenance_b200 import ops
    tag = "adam_%s" % ("mp" if mp else "f32")
    dt = np.float16 if mp else np.float32
    w = _cuda(split(G[tag + "_w0"], dt))
    w32 = [x.float() for x in w] if mp else None
    mean = _cuda([np.zeros(s, np.float32) for s in SIZES])
    var = _cuda([np.zeros(s, np.float32) for s in SIZES])
    for step in range(3):
        lr = adam_lr(1e-3, 0.9, 0.999, step + 1)
        ops.multi_adam_update(w, _cuda(split(G[tag + "_g%d" % step], dt)), mean, var, lr, 0.01, weights32=w32,
                              rescale_grad=0.5)
    assert _flat(w).tobytes() == G[tag + "_w3"].tobytes()
    assert _flat(var).tobytes() == G[tag + "_var3"].tobytes()


@pytest.mark.gpu
def test_cuda_lamb_and_sumsq_golden():
    from mxnet_maintenance_b200 import ops
    w = _cuda(split(G["lamb_w0"]))
    mean = _cuda([np.zeros(s, np.float32) for s in SIZES])
    var = _cuda([np.zeros(s, np.float32) for s in SIZES])
    for step in range(3):
        ops.multi_lamb_update(w, _cuda(split(G["lamb_g%d" % step])), mean, var, 1e-3, 0.01, step + 1,
                              rescale_grad=0.5)
    # norms are tree-reduced on the GPU, sequentially on the CPU: north-star fp32 tolerance
    np.testing.assert_allclose(_flat(w), G["lamb_w3"], rtol=1e-5, atol=1e-6)
    ss = ops.multi_sum_sq(_cuda(split(G["sumsq_in"]))).cpu().numpy()
    np.testing.assert_allclose(ss, G["sumsq_out"], rtol=1e-5, atol=1e-5)   # test_operator_gpu.py:277-305
    # run-to-run determinism, bitwise (test_operator_gpu.py:281-283)
    ss2 = ops.multi_sum_sq(_cuda(split(G["sumsq_in"]))).cpu().numpy()
    assert ss.tobytes() == ss2.tobytes()


@pytest.mark.gpu
@pytest.mark.parametrize("max_norm", [0.05, 1e6])
def test_cuda_clip_global_norm(oracle, max_norm):
    """ops.clip_global_norm vs the oracle's restatement of gluon.utils.clip_global_norm."""
    from mxnet_maintenance_b200 import ops
    rng = np.random.default_rng(9)
    arrs = [(rng.uniform(-1, 1, s) * 1e-2).astype(np.float32) for s in (1, 77, 4096, 100003, 4097)]
    dev = _cuda(arrs)
    out = ops.clip_global_norm(dev, max_norm).cpu().numpy()
    ref = [a.copy() for a in arrs]
    total = oracle.clip_global_norm(ref, max_norm)
    truth = float(np.sqrt(sum((a.astype(np.float64) ** 2).sum() for a in arrs)))
    assert abs(out[1] - truth) <= 1e-6 * truth and abs(total - truth) <= 1e-5 * truth
    for d, r in zip(dev, ref):
        np.testing.assert_allclose(d.cpu().numpy(), r, rtol=1e-5, atol=1e-6)
    if max_norm > 1:
        for d, a in zip(dev, arrs):
            assert d.cpu().numpy().tobytes() == a.tobytes()       # scale == 1: untouched bits


ETAS = [1.0, 0.5, 2.0, 1.0, 1.0, 0.1]


def test_oracle_adamw_lars_adagrad(oracle):
    w = split(G["adamw_w0"])
    mean = [np.zeros(s, np.float32) for s in SIZES]
    var = [np.zeros(s, np.float32) for s in SIZES]
    for step, rs in enumerate([0.5, float("inf"), 0.25]):
        oracle.multi_adamw(w, split(G["adamw_g%d" % step]), mean, var, None, LRS, WDS, ETAS, rescale=rs, clip=0.004)
    assert np.concatenate(w).tobytes() == G["adamw_w3"].tobytes()
    assert np.concatenate(var).tobytes() == G["adamw_var3"].tobytes()
    li = G["lars_in"]
    assert oracle.multi_lars(li[0], li[1], li[2], li[3], 0.001, 1e-8, 0.5).tobytes() == G["lars_out"].tobytes()
    W, h = G["rsp_w0"].copy(), np.zeros_like(G["rsp_w0"])
    oracle.adagrad_rsp(W, h, G["rsp_idx"], G["rsp_g"], 0.1, 1e-7, 0.5)
    oracle.adagrad_rsp(W, h, G["rsp_idx"], G["rsp_g"], 0.1, 1e-7, 0.5, 0.003)
    assert W.tobytes() == G["adagrad_rsp_w2"].tobytes() and h.tobytes() == G["adagrad_rsp_h2"].tobytes()


@pytest.mark.gpu
def test_cuda_adamw_lars_golden():
    import torch
    from mxnet_maintenance_b200 import ops
    w = _cuda(split(G["adamw_w0"]))
    mean = _cuda([np.zeros(s, np.float32) for s in SIZES])
    var = _cuda([np.zeros(s, np.float32) for s in SIZES])
    for step, rs in enumerate([0.5, float("inf"), 0.25]):
        ops.multi_adamw_update(w, _cuda(split(G["adamw_g%d" % step])), mean, var,
                               torch.tensor([rs], device="cuda"), LRS, WDS, ETAS, clip_gradient=0.004)
    assert _flat(w).tobytes() == G["adamw_w3"].tobytes()
    assert _flat(var).tobytes() == G["adamw_var3"].tobytes()
    li = [torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in G["lars_in"]]
    got = ops.multi_lars(li[0], li[1], li[2], li[3], 0.001, 1e-8, 0.5).cpu().numpy()
    assert got.tobytes() == G["lars_out"].tobytes()
