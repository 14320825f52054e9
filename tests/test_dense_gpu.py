"""Parity of the CUDA path with the CPU oracle for the dense gradient-sync + update path, through
the kvstore plugin API (which calls the C ABI).  fp32 paths are compared BIT-EXACTLY wherever the
reduction order is the reference's (device-local copies are summed in CommCPU order); LAMB norms
use a tree reduction instead of the CPU's sequential sum, so LAMB is compared at the north-star
tolerance rtol 1e-5 / atol 1e-6 (fp32) and rtol 2e-3 (fp16 weights)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

RTOL32, ATOL32 = 1e-5, 1e-6     # BASELINE.json north_star: fp32 weights / states per step
RTOL16 = 2e-3                   # fp16-gradient / fp32-master paths


def _mb():
    import mxnet_maintenance_b200 as mb
    return mb


def _shapes(name):
    from mxnet_maintenance_b200 import shapes
    if name == "ragged":   # edge cases: 1 element, non-multiples of 4, chunk boundaries +-1
        return [(1,), (3,), (5, 1), (4095,), (4096,), (4097,), (2, 4099), (8192,), (33, 7, 3), (100003,)]
    return shapes.CONFIGS[name]()


def _run(oracle, shape_name, opt_kind, ncopies, steps=3, mp=False, clip=None, gclip=None, batch=True,
         momentum=0.9, seed=0, wd=1e-4):
    from oracle.oracle import LocalStoreModel
    mb = _mb()
    rng = np.random.default_rng(1234 + seed)
    shp = _shapes(shape_name)
    wdt = np.float16 if mp else np.float32
    w0 = [rng.uniform(-1, 1, s).astype(np.float32).astype(wdt) for s in shp]
    keys = list(range(len(shp)))
    # LAMB runs at its C4 learning rate; the others at a deliberately large one
    lr = 1e-3 if opt_kind == "lamb" else 0.05
    common = dict(learning_rate=lr, wd=wd, rescale_grad=1.0 / 8, clip_gradient=clip, multi_precision=mp)
    odict = dict(lr=lr, wd=wd, rescale=1.0 / 8, clip=clip if clip else -1.0, multi_precision=mp)
    if opt_kind == "sgd":
        opt = mb.optimizer.SGD(momentum=momentum, **common)
        odict.update(kind="sgd", momentum=momentum)
    elif opt_kind == "adam":
        opt = mb.optimizer.Adam(**common)
        odict.update(kind="adam")
    elif opt_kind == "lamb":
        opt = mb.optimizer.LAMB(**common)
        odict.update(kind="lamb")
    elif opt_kind == "adagrad":
        opt = mb.optimizer.AdaGrad(eps=1e-7, **common)
        odict.update(kind="adagrad", eps=1e-7)
    else:
        opt = None
        odict = None
    kv = mb.create("b200")
    if opt is not None:
        kv.set_optimizer(opt)
    if gclip is not None:
        kv.clip_global_norm = gclip
    model = LocalStoreModel(oracle, odict)
    dev_w = [torch.from_numpy(w).cuda() for w in w0]
    for k, w in zip(keys, w0):
        model.init(k, w)
    kv.broadcast(keys, dev_w, [[w] for w in dev_w])
    results = []
    for step in range(steps):
        grads = [[(rng.uniform(-1, 1, s) * 1e-2).astype(np.float32).astype(wdt) for _ in range(ncopies)]
                 for s in shp]
        dev_g = [[torch.from_numpy(g).cuda() for g in gs] for gs in grads]
        if batch:
            kv.pushpull(keys, dev_g, out=[[w] for w in dev_w])
        else:
            for k in keys:
                kv.pushpull(k, dev_g[k], out=dev_w[k])
        kv.wait()
        if gclip is not None:
            merged = [oracle.reduce_sum([g.copy() for g in gs]).astype(np.float32) for gs in grads]
            oracle.clip_global_norm(merged, gclip)
            ref = [model.pushpull(k, [merged[k].astype(wdt)]) for k in keys]
        else:
            ref = [model.pushpull(k, grads[k]) for k in keys]
        got = [w.cpu().numpy() for w in dev_w]
        if step == steps - 1 and opt is not None:
            # "weights AND optimizer states" (north star): every state array the reference keeps
            # (momentum | mean, var | history, fp32 master) is compared like one more key
            for k in keys:
                for which, want in sorted(model.states_of(k).items()):
                    have = kv.get_state(k, which)
                    assert have is not None, "key %d: state %d missing in the store" % (k, which)
                    got.append(have.reshape(-1))
                    ref.append(want)
        results.append((got, ref))
    return kv, model, results


def _assert_bits(results):
    for step, (got, ref) in enumerate(results):
        for k, (a, b) in enumerate(zip(got, ref)):
            assert a.tobytes() == b.tobytes(), "key %d step %d: max |diff| %g" % (
                k, step, np.abs(a.astype(np.float64) - b.astype(np.float64)).max())


def _assert_close(results, rtol, atol):
    for step, (got, ref) in enumerate(results):
        for k, (a, b) in enumerate(zip(got, ref)):
            np.testing.assert_allclose(a.astype(np.float32), b.astype(np.float32), rtol=rtol, atol=atol,
                                       err_msg="key %d step %d" % (k, step))


@pytest.mark.parametrize("ncopies", [1, 2, 4, 5, 8])
def test_sgd_mom_ragged_bit_exact(oracle, ncopies):
    _, _, res = _run(oracle, "ragged", "sgd", ncopies)
    _assert_bits(res)


@pytest.mark.parametrize("clip", [None, 0.0005])
@pytest.mark.parametrize("momentum", [0.9, 0.0])
def test_sgd_variants_bit_exact(oracle, clip, momentum):
    _, _, res = _run(oracle, "ragged", "sgd", 3, clip=clip, momentum=momentum)
    _assert_bits(res)


def test_sgd_mom_mlp_4copies_bit_exact(oracle):
    """C1: Gluon MLP, 4 device copies, SGD momentum 0.9 (the reference's CPU-runnable config)."""
    _, _, res = _run(oracle, "mlp", "sgd", 4)
    _assert_bits(res)


def test_sgd_mom_resnet50_bit_exact(oracle):
    """C2 at one GPU: all 161 arrays in one pushpull."""
    kv, _, res = _run(oracle, "resnet50", "sgd", 1)
    _assert_bits(res)
    # broadcast unpack + 3 fused update launches, nothing per key
    assert kv.launch_count() <= 8


def test_per_key_calls_match_batched(oracle):
    _, _, res = _run(oracle, "mlp", "sgd", 2, batch=False)
    _assert_bits(res)


def test_mp_sgd_mom_bit_exact(oracle):
    """multi_mp_sgd_mom_update: fp16 grads/weights, fp32 master + momentum (C3 kernel)."""
    _, _, res = _run(oracle, "ragged", "sgd", 1, mp=True)
    _assert_bits(res)
    _, _, res = _run(oracle, "ragged", "sgd", 4, mp=True)   # fp16 CommCPU sum: rounds after each add
    _assert_bits(res)


@pytest.mark.parametrize("clip", [None, 0.0005])
def test_adam_bit_exact(oracle, clip):
    _, _, res = _run(oracle, "ragged", "adam", 2, clip=clip, wd=0.01)
    _assert_bits(res)


def test_mp_adam_bit_exact(oracle):
    """fused fp16-grad -> fp32 Adam -> fp16 weight equals the reference's cast/update/cast wrapper."""
    _, _, res = _run(oracle, "ragged", "adam", 1, mp=True, wd=0.01)
    _assert_bits(res)


def test_lamb_close(oracle):
    _, _, res = _run(oracle, "ragged", "lamb", 1, wd=0.01)
    _assert_close(res, RTOL32, ATOL32)
    _, _, res = _run(oracle, "ragged", "lamb", 4, wd=0.01, clip=0.001)
    _assert_close(res, RTOL32, ATOL32)


def test_mp_lamb_close(oracle):
    _, _, res = _run(oracle, "ragged", "lamb", 1, mp=True, wd=0.01)
    _assert_close(res, RTOL16, 1e-3)


def test_lamb_global_norm_clip_close(oracle):
    """C4 recipe: allreduce -> clip_global_norm -> multi_lamb_update, run on the store."""
    kv, _, res = _run(oracle, "ragged", "lamb", 2, wd=0.01, gclip=0.05)
    _assert_close(res, RTOL32, ATOL32)
    assert kv.last_global_norm() > 0.05


def test_no_optimizer_assigns_merged(oracle):
    """push without an updater replaces the stored value by the merged one (kvstore_local.h:268-274)."""
    _, _, res = _run(oracle, "ragged", None, 4)
    _assert_bits(res)


def test_bert_base_mp_sgd_one_step(oracle):
    """C3 shapes (109M parameters), one step, bit-exact."""
    _, _, res = _run(oracle, "bert_base", "sgd", 1, mp=True, steps=1)
    _assert_bits(res)


@pytest.mark.parametrize("kind", ["sgd", "adam", "lamb"])
def test_bound_step_equals_pushpull(kind):
    """bind_pushpull replays must be bit-identical with plain pushpull calls on the same inputs,
    including the per-step scalars (lr schedule, lr/wd multipliers, update counts)."""
    import mxnet_maintenance_b200 as mb

    class Sched:
        base_lr = 0.02

        def __call__(self, n):
            return self.base_lr / (1 + n)

    class P:
        def __init__(self, lm, wm):
            self.lr_mult, self.wd_mult = lm, wm

    shp = _shapes("ragged")
    keys = list(range(len(shp)))
    rng = np.random.default_rng(7)
    w0 = [rng.uniform(-1, 1, s).astype(np.float32) for s in shp]
    g = [[(rng.uniform(-1, 1, s) * 1e-2).astype(np.float32) for s in shp] for _ in range(4)]
    res = []
    for bound in (False, True):
        O = {"sgd": mb.optimizer.SGD, "adam": mb.optimizer.Adam, "lamb": mb.optimizer.LAMB}[kind]
        kw = dict(momentum=0.9) if kind == "sgd" else {}
        opt = O(learning_rate=0.02, lr_scheduler=Sched(), wd=1e-3,
                param_dict={k: P(1.0 + 0.25 * (k % 3), 0.5 * (k % 2)) for k in keys}, **kw)
        kv = mb.create("b200")
        kv.set_optimizer(opt)
        w = [torch.from_numpy(x).cuda() for x in w0]
        gd = [torch.empty_like(x) for x in w]
        kv.broadcast(keys, w, [[x] for x in w])
        step = kv.bind_pushpull(keys, gd, out=[[x] for x in w]) if bound else None
        for it in range(4):
            for d, h in zip(gd, g[it]):
                d.copy_(torch.from_numpy(h))
            if it == 2:
                opt.param_dict[1].lr_mult = 3.0     # multipliers may change mid-training
            if bound:
                step()
            else:
                kv.pushpull(keys, gd, out=[[x] for x in w])
        kv.wait()
        res.append([x.cpu().numpy() for x in w])
        assert opt.num_update == 4 and opt._index_update_count[0] == 4
    for a, b in zip(*res):
        assert a.tobytes() == b.tobytes()


@pytest.mark.parametrize("kind,mp", [("sgd", False), ("sgd", True), ("adam", False), ("lamb", False)])
def test_store_views_update_in_place(oracle, kind, mp):
    """Arrays that are zero-copy views of the store (b200kv_view) take the copy-free path; results
    must not change (bit-exact vs the oracle for SGD / Adam)."""
    from oracle.oracle import LocalStoreModel
    import mxnet_maintenance_b200 as mb
    shp = _shapes("ragged")
    keys = list(range(len(shp)))
    rng = np.random.default_rng(21)
    wdt = np.float16 if mp else np.float32
    w0 = [rng.uniform(-1, 1, s).astype(np.float32).astype(wdt) for s in shp]
    lr = 1e-3 if kind == "lamb" else 0.05
    O = {"sgd": mb.optimizer.SGD, "adam": mb.optimizer.Adam, "lamb": mb.optimizer.LAMB}[kind]
    kw = dict(momentum=0.9) if kind == "sgd" else {}
    kv = mb.create("b200")
    kv.set_optimizer(O(learning_rate=lr, wd=0.01, rescale_grad=0.5, multi_precision=mp, **kw))
    od = dict(kind=kind, lr=lr, wd=0.01, rescale=0.5, multi_precision=mp)
    if kind == "sgd":
        od["momentum"] = 0.9
    model = LocalStoreModel(oracle, od)
    kv.init(keys, [torch.from_numpy(w).cuda() for w in w0])
    for k, w in zip(keys, w0):
        model.init(k, w)
    wv = [kv.weight_view(k, s) for k, s in zip(keys, shp)]
    gv = [kv.grad_view(k, s) for k, s in zip(keys, shp)]
    for w, h in zip(wv, w0):
        assert w.cpu().numpy().tobytes() == h.tobytes()      # the view shows the initial value
    base = kv.launch_count()
    for step in range(3):
        g = [(rng.uniform(-1, 1, s) * 1e-2).astype(np.float32).astype(wdt) for s in shp]
        for d, h in zip(gv, g):
            d.copy_(torch.from_numpy(h))
        kv.pushpull(keys, gv, out=[[w] for w in wv])
        kv.wait()
        for k in keys:
            ref = model.pushpull(k, [g[k]])
            got = wv[k].cpu().numpy()
            if kind == "lamb":
                np.testing.assert_allclose(got, ref, rtol=RTOL32, atol=ATOL32)
            else:
                assert got.tobytes() == ref.tobytes()
    per_step = (kv.launch_count() - base) / 3
    assert per_step <= (4 if kind == "lamb" else 1) + (1 if mp else 0)
    # a plain pull still returns the stored value
    out = [torch.empty_like(w) for w in wv]
    kv.pull(keys, out=out)
    for a, b in zip(out, wv):
        assert torch.equal(a, b)


def test_mixed_dtypes_late_init_and_optimizer_switch(oracle):
    """One pushpull over fp16 and fp32 keys (BERT-style fp32 LayerNorm inside an fp16 model), keys
    initialised after the first push (a second flat segment), and an optimizer change that resets
    the state — each step compared with the oracle."""
    from oracle.oracle import LocalStoreModel
    import mxnet_maintenance_b200 as mb
    rng = np.random.default_rng(31)
    spec = [((300, 7), np.float16), ((64,), np.float32), ((4097,), np.float16), ((33,), np.float32)]
    late = [((1000,), np.float32), ((129,), np.float16)]

    def mk(shape, dt, scale=1.0):
        return (rng.uniform(-1, 1, shape) * scale).astype(np.float32).astype(dt)

    kv = mb.create("b200")
    kv.set_optimizer(mb.optimizer.SGD(learning_rate=0.05, momentum=0.9, wd=1e-3, multi_precision=True))
    model = LocalStoreModel(oracle, dict(kind="sgd", lr=0.05, momentum=0.9, wd=1e-3, multi_precision=True))
    w = {}

    def add(k, shape, dt):
        h = mk(shape, dt)
        model.init(k, h)
        w[k] = torch.from_numpy(h).cuda()
        kv.broadcast(k, w[k], out=[w[k]])

    def step(keys, specs):
        g = [mk(s, dt, 1e-2) for s, dt in specs]
        kv.pushpull(keys, [torch.from_numpy(x).cuda() for x in g], out=[[w[k]] for k in keys])
        kv.wait()
        for k, x in zip(keys, g):
            ref = model.pushpull(k, [x])
            assert w[k].cpu().numpy().tobytes() == ref.tobytes(), "key %d" % k

    for k, (s, dt) in enumerate(spec):
        add(k, s, dt)
    keys = list(range(len(spec)))
    step(keys, spec)
    step(keys, spec)
    for j, (s, dt) in enumerate(late):           # new keys after the first push: a new segment
        add(len(spec) + j, s, dt)
    allkeys, allspec = keys + [4, 5], spec + late
    step(allkeys, allspec)
    step([5, 0, 3], [late[1], spec[0], spec[3]])   # a subset, out of order
    # switching the optimizer resets the state (new Updater in the reference, kvstore.py:553-600)
    kv.set_optimizer(mb.optimizer.Adam(learning_rate=1e-3, wd=0.01, multi_precision=True))
    model.opt = dict(kind="adam", lr=1e-3, wd=0.01, multi_precision=True)
    model.state, model.count = {}, {}
    step(allkeys, allspec)
    step(allkeys, allspec)
    # and a pull still returns the stored (fp16 / fp32) values
    outs = [torch.empty_like(w[k]) for k in allkeys]
    kv.pull(allkeys, out=outs)
    for k, o in zip(allkeys, outs):
        assert torch.equal(o, w[k])


def test_uneven_copy_counts_in_one_call(oracle):
    """Keys with different numbers of device copies in the same pushpull (separate launch classes)."""
    _, _, _ = None, None, None
    from oracle.oracle import LocalStoreModel
    import mxnet_maintenance_b200 as mb
    rng = np.random.default_rng(32)
    shp = [(100,), (5000,), (77, 3)]
    ncop = [1, 3, 2]
    kv = mb.create("b200")
    kv.set_optimizer(mb.optimizer.SGD(learning_rate=0.1, momentum=0.9))
    model = LocalStoreModel(oracle, dict(kind="sgd", lr=0.1, momentum=0.9))
    w0 = [rng.uniform(-1, 1, s).astype(np.float32) for s in shp]
    dev = [torch.from_numpy(x).cuda() for x in w0]
    for k, x in enumerate(w0):
        model.init(k, x)
    kv.broadcast([0, 1, 2], dev, out=[[d] for d in dev])
    for _ in range(2):
        g = [[(rng.uniform(-1, 1, s) * 1e-2).astype(np.float32) for _ in range(n)] for s, n in zip(shp, ncop)]
        kv.pushpull([0, 1, 2], [[torch.from_numpy(x).cuda() for x in gs] for gs in g], out=[[d] for d in dev])
        kv.wait()
        for k in range(3):
            assert dev[k].cpu().numpy().tobytes() == model.pushpull(k, g[k]).tobytes()


def test_bert_large_lamb_clip_full_size(oracle):
    """C4 at full size (391 arrays, 335,141,888 parameters): one LAMB step with clip_global_norm on
    the store vs the oracle (clip_global_norm restatement + the reference's MultiLAMB functors).
    Weights within the north-star fp32 tolerance; the global norm and the per-tensor sums of
    squares are also checked against a float64 truth (the CPU reference accumulates 31M-element
    tensors sequentially in fp32, the GPU reduces pairwise — SURVEY.md Appendix A.4)."""
    import mxnet_maintenance_b200 as mb
    from mxnet_maintenance_b200 import shapes, ops
    shp = shapes.bert_large()
    keys = list(range(len(shp)))
    rng = np.random.default_rng(4)
    w0 = [rng.uniform(-1, 1, s).astype(np.float32) for s in shp]
    g0 = [(rng.uniform(-1, 1, s) * 1e-2).astype(np.float32) for s in shp]
    kv = mb.create("b200")
    kv.set_optimizer(mb.optimizer.LAMB(learning_rate=1e-3, beta1=0.9, beta2=0.999, epsilon=1e-6, wd=0.01))
    kv.clip_global_norm = 1.0
    dev_w = [torch.from_numpy(w).cuda() for w in w0]
    dev_g = [torch.from_numpy(g).cuda() for g in g0]
    kv.broadcast(keys, dev_w, [[w] for w in dev_w])
    kv.pushpull(keys, dev_g, out=[[w] for w in dev_w])
    kv.wait()
    truth_norm = float(np.sqrt(sum(float((g.astype(np.float64) ** 2).sum()) for g in g0)))
    assert abs(kv.last_global_norm() - truth_norm) <= 2e-6 * truth_norm
    # sums of squares of the largest tensor: GPU tree reduction vs float64
    big = int(np.argmax([shapes.numel(s) for s in shp]))
    ss = ops.multi_sum_sq([dev_g[big]]).item()
    t64 = float((g0[big].astype(np.float64) ** 2).sum())
    assert abs(ss - t64) <= 1e-6 * t64
    # oracle step (reference CPU functors: sums of squares accumulated sequentially in fp32)
    gc = [g.copy().reshape(-1) for g in g0]
    oracle.clip_global_norm(gc, 1.0)
    wr = [w.copy().reshape(-1) for w in w0]
    me = [np.zeros(w.size, np.float32) for w in w0]
    va = [np.zeros(w.size, np.float32) for w in w0]
    oracle.multi_lamb(wr, gc, me, va, None, [1e-3] * len(shp), [0.01] * len(shp), [1] * len(shp), eps=1e-6)
    f = np.float32
    worst_gpu = worst_cpu = 0.0
    for k in keys:
        got = dev_w[k].cpu().numpy().reshape(-1)
        if w0[k].size <= (1 << 20):
            # up to ~1M elements the CPU's sequential fp32 sums are still accurate: direct parity
            np.testing.assert_allclose(got, wr[k], rtol=RTOL32, atol=ATOL32, err_msg="key %d" % k)
            continue
        # Larger tensors: the CPU kernel's `sum += x*x` in fp32 stalls once the sum dwarfs the addends
        # (multi_sum_sq.cc:68-84), so its trust ratio is off by far more than the tolerance.  Adjudicate
        # with the same fp32 elementwise step-1 math and float64-accurate norms.
        w, g = w0[k].reshape(-1), gc[k]
        b1, b2, eps, wd, lr = f(0.9), f(0.999), f(1e-6), f(0.01), f(1e-3)
        m = b1 * f(0) + (f(1) - b1) * g
        v = b2 * f(0) + ((f(1) - b2) * g) * g
        u = (m / (f(1) - b1)) / (np.sqrt(v / (f(1) - b2)) + eps) + wd * w
        r1 = np.sqrt(f((w.astype(np.float64) ** 2).sum()))
        r2 = np.sqrt(f((u.astype(np.float64) ** 2).sum()))
        truth = w - (lr * (r1 / r2)) * u
        np.testing.assert_allclose(got, truth, rtol=RTOL32, atol=ATOL32, err_msg="key %d vs float64 norms" % k)
        worst_gpu = max(worst_gpu, float(np.abs(got - truth).max()))
        worst_cpu = max(worst_cpu, float(np.abs(wr[k] - truth).max()))
    # the GPU path is closer to the exact-norm result than the CPU reference on every big tensor
    assert worst_gpu < 1e-6 and worst_gpu < worst_cpu


@pytest.mark.parametrize("mp", [False, True])
@pytest.mark.parametrize("clip", [None, 0.0005])
def test_adagrad_bit_exact(oracle, mp, clip):
    """AdaGrad on the store (dense branch of AdaGrad.update, optimizer.py:1681-1687)."""
    _, _, res = _run(oracle, "ragged", "adagrad", 2, mp=mp, clip=clip, wd=0.01)
    _assert_bits(res)


@pytest.mark.parametrize("kind", ["sgd", "adam", "lamb", "adamw"])
@pytest.mark.parametrize("dt", [torch.float32, torch.float16])
def test_no_out_of_bounds_writes(kind, dt):
    """compute-sanitizer is unavailable on this pool, so guard every array with canary regions and
    check that no kernel wrote outside [begin, end) — for odd sizes, odd alignments (arrays that
    start 4 / 2 bytes past a 16-byte boundary take the scalar path) and chunk-boundary sizes."""
    from mxnet_maintenance_b200 import ops
    sizes = [1, 2, 3, 5, 63, 4095, 4096, 4097, 8191, 12289]
    mp = dt == torch.float16
    PAD = 64

    def guarded(n, dtype, shift):
        buf = torch.full((n + 2 * PAD + shift,), 7.0, device="cuda", dtype=dtype)
        return buf, buf[PAD + shift:PAD + shift + n]

    results = []
    for shift in (0, 1, 3):
        gen = torch.Generator(device="cuda").manual_seed(5)     # same values at every alignment
        bufs, w, g, s0, s1, w32 = [], [], [], [], [], []
        for n in sizes:
            for lst, dtype in ((w, dt), (g, dt), (s0, torch.float32), (s1, torch.float32), (w32, torch.float32)):
                b, v = guarded(n, dtype, shift)
                v.copy_(torch.rand(n, device="cuda", generator=gen) * 0.01)
                bufs.append((b, n, shift))
                lst.append(v)
        m32 = w32 if mp else None
        if kind == "sgd":
            ops.multi_sgd_update(w, g, 0.1, 1e-3, moms=s0, weights32=m32, momentum=0.9)
        elif kind == "adam":
            ops.multi_adam_update(w, g, s0, s1, 1e-3, 0.01, weights32=m32)
        elif kind == "lamb":
            ops.multi_lamb_update(w, g, s0, s1, 1e-3, 0.01, 1, weights32=m32)
        else:
            ops.multi_adamw_update(w, g, s0, s1, torch.ones(1, device="cuda"), 1e-3, 0.01, 1.0, weights32=m32)
        results.append([x.clone() for x in w + s0 + s1 + w32])
        ops.clip_global_norm(g, 1e-3)
        ops.reset_arrays(g)
        torch.cuda.synchronize()
        for b, n, sh in bufs:
            assert (b[:PAD + sh] == 7.0).all() and (b[PAD + sh + n:] == 7.0).all(), (kind, n, sh)
        assert all((x == 0).all() for x in g)
    # the scalar (unaligned) path computes the same bits as the 128-bit vector path
    # (LAMB's norms are summed in a different order on the two paths: tolerance instead of bits)
    for other in results[1:]:
        for a, b in zip(results[0], other):
            if kind == "lamb":
                assert torch.allclose(a.float(), b.float(), rtol=2e-3 if mp else 1e-5, atol=1e-6)
            else:
                assert torch.equal(a, b)


def test_layout_planner_matches_store():
    """tests/layout_planner.py (used by the gloo tests to shard a CPU update like the store does) against
    the REAL layout: element offsets of every key inside its segment, read back through the
    addresses of the store's own views (b200kv_view)."""
    import layout_planner
    mb = _mb()
    shp = _shapes("ragged") + _shapes("mlp")
    kv = mb.create("b200")
    keys = list(range(len(shp)))
    kv.init(keys, [torch.zeros(s, device="cuda") for s in shp])
    sizes = [int(np.prod(s)) for s in shp]
    offs, total, sl = layout_planner.segment_layout(sizes, 1)
    seg_w, seg_g = kv.segment_views(0)
    assert seg_w.numel() == total == sl
    for k, o in zip(keys, offs):
        assert (kv.weight_view(k).data_ptr() - seg_w.data_ptr()) // 4 == o
        assert (kv.grad_view(k).data_ptr() - seg_g.data_ptr()) // 4 == o
