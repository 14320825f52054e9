"""Generates tests/golden/ref_vectors.npz from the REFERENCE's own functors (oracle/_ref, compiled
in place from /root/reference).  Run in the build container only:

    python tests/golden/make_golden.py

The fixtures let the GPU box (which has no /root/reference) and later rounds check both the plain-C
oracle and the CUDA path against outputs of the reference code itself."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.oracle import Oracle, adam_lr  # noqa: E402


def cases(o):
    rng = np.random.default_rng(20260611)
    out = {}
    sizes = [1, 3, 7, 130, 1000, 4097]

    def arrs(scale=1.0, dt=np.float32):
        return [(rng.uniform(-1, 1, s) * scale).astype(np.float32).astype(dt) for s in sizes]

    # dense reduce, CommCPU grouping, fp32 + fp16, 2..9 copies
    for dt, tag in ((np.float32, "f32"), (np.float16, "f16")):
        for n in (2, 3, 4, 5, 8, 9):
            src = [(rng.uniform(-1, 1, 1027) * 1e-2).astype(np.float32).astype(dt) for _ in range(n)]
            out["reduce_%s_%d_in" % (tag, n)] = np.stack(src)
            out["reduce_%s_%d_out" % (tag, n)] = o.reduce_sum([x.copy() for x in src])
    # multi_sgd_mom_update (fp32) / multi_mp_sgd_mom_update, 3 steps
    lrs = [0.1, 0.05, 0.2, 0.1, 0.01, 0.1]
    wds = [1e-4, 0.0, 1e-3, 1e-4, 1e-2, 1e-4]
    for mp in (False, True):
        for clip in (-1.0, 0.002):
            tag = "sgd_%s_%s" % ("mp" if mp else "f32", "clip" if clip > 0 else "noclip")
            dt = np.float16 if mp else np.float32
            w = arrs(dt=dt)
            w32 = [x.astype(np.float32) for x in w] if mp else None
            mom = [np.zeros(s, np.float32) for s in sizes]
            out[tag + "_w0"] = np.concatenate([x.astype(np.float32) for x in w])
            for step in range(3):
                g = arrs(1e-2, dt)
                out[tag + "_g%d" % step] = np.concatenate([x.astype(np.float32) for x in g])
                o.multi_sgd(w, g, mom, w32, lrs, wds, 0.9, 0.25, clip)
                if step == 2:
                    out[tag + "_w3"] = np.concatenate([x.astype(np.float32) for x in w])
                    out[tag + "_mom3"] = np.concatenate(mom)
    # adam_update fp32 + mixed-precision wrapper
    for mp in (False, True):
        tag = "adam_%s" % ("mp" if mp else "f32")
        dt = np.float16 if mp else np.float32
        w = arrs(dt=dt)
        w32 = [x.astype(np.float32) for x in w]
        mean = [np.zeros(s, np.float32) for s in sizes]
        var = [np.zeros(s, np.float32) for s in sizes]
        out[tag + "_w0"] = np.concatenate([x.astype(np.float32) for x in w])
        for step in range(3):
            g = arrs(1e-2, dt)
            out[tag + "_g%d" % step] = np.concatenate([x.astype(np.float32) for x in g])
            lr = adam_lr(1e-3, 0.9, 0.999, step + 1)
            for k in range(len(sizes)):
                if mp:
                    o.mp_adam(w[k], g[k], w32[k], mean[k], var[k], lr, 0.01, rescale=0.5)
                else:
                    o.adam(w[k], g[k], mean[k], var[k], lr, 0.01, rescale=0.5)
            if step == 2:
                out[tag + "_w3"] = np.concatenate([x.astype(np.float32) for x in w])
                out[tag + "_var3"] = np.concatenate(var)
    # multi_lamb_update fp32, 3 steps
    w = arrs()
    mean = [np.zeros(s, np.float32) for s in sizes]
    var = [np.zeros(s, np.float32) for s in sizes]
    out["lamb_w0"] = np.concatenate(w)
    for step in range(3):
        g = arrs(1e-2)
        out["lamb_g%d" % step] = np.concatenate(g)
        r1, r2 = o.multi_lamb(w, g, mean, var, None, [1e-3] * 6, [0.01] * 6, [step + 1] * 6, rescale=0.5)
        if step == 2:
            out["lamb_w3"] = np.concatenate(w)
            out["lamb_r1_3"] = r1
            out["lamb_r2_3"] = r2
    # multi_sum_sq
    a = arrs()
    out["sumsq_in"] = np.concatenate(a)
    out["sumsq_out"] = o.multi_sum_sq(a)
    # lazy row-sparse SGD-momentum + sparse_retain
    W = rng.uniform(-1, 1, (64, 16)).astype(np.float32)
    idx = np.array([0, 3, 4, 17, 40, 63], np.int64)
    gr = (rng.uniform(-1, 1, (6, 16)) * 1e-2).astype(np.float32)
    mom = np.zeros_like(W)
    out["rsp_w0"], out["rsp_idx"], out["rsp_g"] = W.copy(), idx, gr
    o.sgd_mom_rsp(W, mom, idx, gr, 0.1, 1e-3, 0.9, 0.5)
    o.sgd_mom_rsp(W, mom, idx, gr, 0.1, 1e-3, 0.9, 0.5)
    out["rsp_w2"], out["rsp_mom2"] = W.copy(), mom.copy()
    ri, rv = o.sparse_retain(idx, gr, [0, 1, 4, 40, 41, 63])
    out["retain_idx"], out["retain_val"] = ri, rv
    # _multi_adamw_update (fp32), 3 steps incl. one skipped (non-finite loss scale)
    w = arrs()
    mean = [np.zeros(s, np.float32) for s in sizes]
    var = [np.zeros(s, np.float32) for s in sizes]
    out["adamw_w0"] = np.concatenate(w)
    for step, rs in enumerate([0.5, float("inf"), 0.25]):
        g = arrs(1e-2)
        out["adamw_g%d" % step] = np.concatenate(g)
        o.multi_adamw(w, g, mean, var, None, lrs, wds, [1.0, 0.5, 2.0, 1.0, 1.0, 0.1], rescale=rs, clip=0.004)
    out["adamw_w3"], out["adamw_var3"] = np.concatenate(w), np.concatenate(var)
    # multi_lars
    l_in = [rng.uniform(0.01, 0.2, 40).astype(np.float32), (rng.uniform(0, 5, 40) * (rng.random(40) > 0.2)).astype(np.float32),
            (rng.uniform(0, 1e-2, 40) * (rng.random(40) > 0.2)).astype(np.float32), rng.uniform(0, 1e-3, 40).astype(np.float32)]
    out["lars_in"] = np.stack(l_in)
    out["lars_out"] = o.multi_lars(l_in[0], l_in[1], l_in[2], l_in[3], 0.001, 1e-8, 0.5)
    # lazy row-sparse AdaGrad
    Wa, ha = out["rsp_w0"].copy(), np.zeros_like(out["rsp_w0"])
    o.adagrad_rsp(Wa, ha, idx, gr, 0.1, 1e-7, 0.5)
    o.adagrad_rsp(Wa, ha, idx, gr, 0.1, 1e-7, 0.5, 0.003)
    out["adagrad_rsp_w2"], out["adagrad_rsp_h2"] = Wa, ha
    out["sizes"] = np.asarray(sizes, np.int64)
    return out


if __name__ == "__main__":
    o = Oracle("ref", nthread=4)
    d = cases(o)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_vectors.npz")
    np.savez_compressed(path, **d)
    print("wrote", path, os.path.getsize(path), "bytes,", len(d), "arrays")
