"""ONE process driving several GPUs (b200kv_create_group) — the reference's own usage model for
kvstore 'device' / 'nccl': one array per device in every push / pull list (CommDevice::Reduce,
src/kvstore/comm.h:503-550; gluon.Trainer, python/mxnet/gluon/trainer.py:392-395).

The multi-rank machinery (peer-memory reduce-scatter + sharded update + all-gather kernel, in-kernel
flag barriers, partial-sum exchange, sharded optimizer state) is exercised on ONE GPU through
"virtual ranks" — a group that lists device 0 several times — so these tests run on the 1-GPU box
too; with two or more GPUs the same cases also run across real devices.  Expected values: the CPU
oracle's CommCPU reduce over the device copies in list order, then the reference optimizer functor;
ranks are summed in CommCPU order on the GPU as well, so fp32 / fp16 results are compared BIT-EXACTLY
(LAMB: tree-reduced norms, north-star tolerance)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

RTOL32, ATOL32 = 1e-5, 1e-6
RAGGED = [(1,), (3,), (5, 1), (4095,), (4096,), (4097,), (2, 4099), (8192,), (33, 7, 3), (100003,)]


def _mb():
    import mxnet_maintenance_b200 as mb
    return mb


def _device_sets():
    sets = [[0, 0], [0, 0, 0], [0] * 4, [0] * 8]
    n = torch.cuda.device_count()
    if n >= 2:
        sets.append([0, 1])
    if n >= 4:
        sets.append([0, 1, 2, 3])
    return sets


def _ids(devs):
    return "x".join(map(str, devs))


def _run(oracle, devs, kind, shp=RAGGED, steps=3, mp=False, gclip=None, batch=True, momentum=0.9, adjudicate=False):
    from oracle.oracle import LocalStoreModel
    mb = _mb()
    W = len(devs)
    rng = np.random.default_rng(99)
    wdt = np.float16 if mp else np.float32
    w0 = [rng.uniform(-1, 1, s).astype(np.float32).astype(wdt) for s in shp]
    keys = list(range(len(shp)))
    lr = 1e-3 if kind == "lamb" else 0.05
    common = dict(learning_rate=lr, wd=1e-3, rescale_grad=1.0 / 8, multi_precision=mp)
    odict = dict(lr=lr, wd=1e-3, rescale=1.0 / 8, multi_precision=mp, kind=kind)
    if kind == "sgd":
        opt = mb.optimizer.SGD(momentum=momentum, **common)
        odict["momentum"] = momentum
    elif kind == "adam":
        opt = mb.optimizer.Adam(**common)
    elif kind == "lamb":
        opt = mb.optimizer.LAMB(**common)
    else:
        opt, odict = None, None
    kv = mb.B200(devices=devs)
    assert kv.num_workers == 1 and kv.rank == 0 and kv.devices == list(devs)
    if opt is not None:
        kv.set_optimizer(opt)
    if gclip is not None:
        kv.clip_global_norm = gclip
    model = LocalStoreModel(oracle, odict)
    for k, w in zip(keys, w0):
        model.init(k, w)
    # one weight copy per device of the group; the init value lives on the first device only
    dev_w = [[torch.empty(w.shape, dtype=torch.float16 if mp else torch.float32, device="cuda:%d" % d)
              for d in devs] for w in w0]
    kv.broadcast(keys, [torch.from_numpy(w).to("cuda:%d" % devs[0]) for w in w0], dev_w)
    for k in keys:
        for c in dev_w[k]:
            assert c.cpu().numpy().tobytes() == w0[k].tobytes()
    res = []
    for step in range(steps):
        grads = [[(rng.uniform(-1, 1, s) * 1e-2).astype(np.float32).astype(wdt) for _ in devs] for s in shp]
        dev_g = [[torch.from_numpy(g).to("cuda:%d" % d) for g, d in zip(gs, devs)] for gs in grads]
        if batch:
            kv.pushpull(keys, dev_g, out=dev_w)
        else:                                    # gluon.Trainer's loop: one call per parameter
            for k in keys:
                kv.pushpull(k, dev_g[k], out=dev_w[k], priority=-k)
        kv.wait()
        torch.cuda.synchronize()
        ends = None
        if gclip is not None:
            merged = [oracle.reduce_sum([g.copy() for g in gs]).astype(np.float32) for gs in grads]
            oracle.clip_global_norm(merged, gclip)
            ref = [model.pushpull(k, [merged[k].astype(wdt)]) for k in keys]
        else:
            ends = _adam_interval(oracle, model, odict, keys, grads) if adjudicate and kind == "adam" and not mp else None
            ref = [model.pushpull(k, grads[k]) for k in keys]
        got = [[c.cpu().numpy() for c in dev_w[k]] for k in keys]
        if gclip is None and ends is not None:
            for k in keys:
                _accept_inside_interval(model, k, got[k][0], ref, ends[k])
        res.append((got, ref))
    states = []
    if opt is not None:
        for k in keys:
            for which, want in sorted(model.states_of(k).items()):
                have = kv.get_state(k, which)
                assert have is not None
                states.append((have.reshape(-1), want))
    return kv, res, states


def _adam_interval(oracle, model, odict, keys, grads):
    """Adam divides by sqrt(v)+eps: where rescale*sum(g) + wd*w is ~0, a one-ulp change of the summed
    gradient (the NVSwitch adds the copies in another order than CommCPU) is amplified by up to lr/eps.
    Returns, per key, the two reference updates from the same state with the merged gradient moved by
    -/+ delta, delta = W * eps32 * sum_r |g_r| (the reduction-order uncertainty)."""
    import copy
    from oracle.oracle import LocalStoreModel
    out = {}
    for k in keys:
        W = len(grads[k])
        merged = oracle.reduce_sum([g.copy() for g in grads[k]])
        delta = W * np.finfo(np.float32).eps * sum(np.abs(g).astype(np.float64) for g in grads[k])
        ends = []
        for sign in (-1.0, 1.0):
            m2 = LocalStoreModel(oracle, odict)
            m2.store = {k: model.store[k].copy()}
            m2.state = {k: copy.deepcopy(model.state[k])} if k in model.state else {}
            m2.count = {k: model.count.get(k, 0)}
            gk = (merged.astype(np.float64) + sign * delta).astype(np.float32)
            ends.append(m2.pushpull(k, [gk]).astype(np.float64))
        out[k] = ends
    return out


def _accept_inside_interval(model, k, got, ref, ends, rtol=RTOL32, atol=ATOL32):
    """An element outside the tolerance is accepted only inside the interval above; it then keeps that
    offset in the weight for good, so later steps are judged from the weights the GPU really holds."""
    r64, g64 = ref[k].astype(np.float64), got.astype(np.float64)
    tol = atol + rtol * np.abs(r64)
    far = np.abs(g64 - r64) > tol
    if not far.any():
        return
    lo, hi = np.minimum(*ends), np.maximum(*ends)
    inside = (g64 >= lo - tol) & (g64 <= hi + tol)
    assert bool(inside[far].all()), "key %d: %d elements outside tolerance AND outside the reduction-order interval" % (
        k, int((far & ~inside).sum()))
    ref[k] = np.where(far, got, ref[k]).astype(ref[k].dtype)
    model.store[k] = got.copy()


def _check(res, states, exact, rtol=RTOL32, atol=ATOL32):
    for step, (got, ref) in enumerate(res):
        for k, (copies, b) in enumerate(zip(got, ref)):
            for d, a in enumerate(copies):
                # every device copy holds the same bits, whatever the comparison with the oracle is
                assert a.tobytes() == copies[0].tobytes(), "key %d step %d: device copies differ" % (k, step)
                if exact:
                    assert a.tobytes() == b.tobytes(), "key %d step %d copy %d: max |diff| %g" % (
                        k, step, d, np.abs(a.astype(np.float64) - b.astype(np.float64)).max())
                else:
                    np.testing.assert_allclose(a.astype(np.float32), b.astype(np.float32), rtol=rtol, atol=atol)
    for have, want in states:
        if exact:
            assert have.tobytes() == want.tobytes()
        else:
            np.testing.assert_allclose(have, want, rtol=rtol, atol=atol)


@pytest.mark.parametrize("devs", _device_sets(), ids=_ids)
def test_group_sgd_momentum_bit_exact(oracle, devs):
    kv, res, states = _run(oracle, devs, "sgd")
    assert kv.last_exchange() in ("peer", "nvls")
    # 4 real devices: the NVSwitch sums the copies in its own order (north-star tolerance); everything
    # else adds the ranks in CommCPU order, bit for bit
    _check(res, states, exact=kv.last_exchange() == "peer")


@pytest.mark.parametrize("devs", [[0, 0], [0] * 4], ids=_ids)
@pytest.mark.parametrize("kind,mp", [("sgd", True), ("adam", False), ("adam", True)])
def test_group_mp_and_adam_bit_exact(oracle, devs, kind, mp):
    _, res, states = _run(oracle, devs, kind, mp=mp)
    _check(res, states, exact=True)


@pytest.mark.parametrize("devs", [[0, 0], [0] * 4], ids=_ids)
@pytest.mark.parametrize("gclip", [None, 0.05])
def test_group_lamb(oracle, devs, gclip):
    """LAMB (+ global-norm clip): norms are tree-reduced per shard and exchanged through peer memory."""
    _, res, states = _run(oracle, devs, "lamb", gclip=gclip)
    _check(res, states, exact=False)


@pytest.mark.parametrize("devs", [[0, 0], [0] * 3], ids=_ids)
def test_group_no_optimizer_assigns_merged(oracle, devs):
    """kvstore_local.h:268-274: without an updater the merged value replaces the stored one."""
    _, res, states = _run(oracle, devs, None)
    _check(res, states, exact=True)


@pytest.mark.parametrize("devs", [[0, 0], [0] * 4], ids=_ids)
def test_group_per_key_trainer_loop(oracle, devs):
    """gluon.Trainer calls pushpull once PER PARAMETER (trainer.py:392-395): every call is its own
    fused exchange over whichever shards the key touches (ranks whose shard of the key is empty
    still take part in the barriers)."""
    from mxnet_maintenance_b200 import shapes
    _, res, states = _run(oracle, devs, "sgd", shp=shapes.mlp() + [(7,), (4097,)], batch=False)
    _check(res, states, exact=True)


@pytest.mark.parametrize("devs", [[0, 0]] + ([[0, 1]] if torch.cuda.device_count() >= 2 else []), ids=_ids)
def test_group_trainer_known_answer(devs):
    """The reference's test_gluon_trainer.py:84-103 closed form, driven the way Trainer drives a
    kvstore over 2 contexts (trainer.py:141-162,371-395): ones gradients on both devices, lr 1,
    momentum .5 -> -2, then lr_mult .5 -> -4."""
    m = _mb()

    class Param:
        lr_mult, wd_mult = 1.0, 1.0
    p = Param()
    opt = m.optimizer.SGD(learning_rate=1.0, momentum=0.5, param_dict={0: p})
    kv = m.B200(devices=devs)
    kv.set_optimizer(opt)
    copies = [torch.zeros(10, device="cuda:%d" % d) for d in devs]
    kv.broadcast(0, copies[0], out=copies)
    for expect in (-2.0, -4.0):
        opt.rescale_grad = 1.0
        grads = [torch.ones(10, device="cuda:%d" % d) for d in devs]
        kv.pushpull(0, grads, out=copies, priority=0)
        kv.wait()
        assert all((c == expect).all() for c in copies)
        p.lr_mult = 0.5


def test_group_rejects_a_missing_device():
    """KVStoreNCCL requires the same device set on every call (kvstore_nccl.h:244-250); so does a group."""
    m = _mb()
    kv = m.B200(devices=[0, 0])
    w = [torch.zeros(8, device="cuda") for _ in range(2)]
    kv.broadcast(0, w[0], out=w)
    kv.pushpull(0, [torch.ones(8, device="cuda"), torch.ones(8, device="cuda")], out=w)
    kv.wait()
    assert all((c == 2).all() for c in w)
    kv3 = m.B200(devices=[0, 0, 0])
    w3 = [torch.zeros(8, device="cuda") for _ in range(3)]
    kv3.broadcast(0, w3[0], out=w3)
    with pytest.raises(m.B200KVError, match="one value per device"):
        kv3.pushpull([0, 0], [torch.ones(8, device="cuda")] * 2, out=w3[:2])     # 2 copies for 3 members


def test_group_views_and_bound_step(oracle):
    """Zero-copy views per member + a replayable bound step: nothing but the fused kernel runs."""
    from oracle.oracle import LocalStoreModel
    m = _mb()
    devs = [0, 0]
    shp = [(4097,), (33, 7), (100003,)]
    rng = np.random.default_rng(5)
    w0 = [rng.uniform(-1, 1, s).astype(np.float32) for s in shp]
    keys = list(range(len(shp)))
    kv = m.B200(devices=devs)
    kv.set_optimizer(m.optimizer.SGD(learning_rate=0.05, momentum=0.9, wd=1e-3))
    model = LocalStoreModel(oracle, dict(kind="sgd", lr=0.05, momentum=0.9, wd=1e-3))
    kv.init(keys, [torch.from_numpy(w).cuda() for w in w0])
    for k, w in zip(keys, w0):
        model.init(k, w)
    wv = [[kv.weight_view(k, s, member=r) for r in range(2)] for k, s in zip(keys, shp)]
    gv = [[kv.grad_view(k, s, member=r) for r in range(2)] for k, s in zip(keys, shp)]
    step = kv.bind_pushpull(keys, gv, out=wv)
    base = kv.launch_count()
    for it in range(3):
        grads = [[(rng.uniform(-1, 1, s) * 1e-2).astype(np.float32) for _ in devs] for s in shp]
        for k in keys:
            for r in range(2):
                gv[k][r].copy_(torch.from_numpy(grads[k][r]))
        step()
        kv.wait()
        for k in keys:
            ref = model.pushpull(k, grads[k])
            for r in range(2):
                assert wv[k][r].cpu().numpy().tobytes() == ref.tobytes()
    assert kv.launch_count() - base == 3 * 2          # one fused kernel per member per step


@pytest.mark.parametrize("devs", [[0, 0], [0] * 3, [0] * 4] + ([[0, 1]] if torch.cuda.device_count() >= 2 else []),
                         ids=_ids)
def test_group_row_sparse(oracle, devs):
    """Row-sparse key in a group: one row-sparse gradient per device, union-sum in device order
    (ndarray_function.cc:81-113), lazy SGD-momentum rows, then row_sparse_pull.  With an optimizer the
    table and its momentum are dealt cyclically (row id % members) over the members (3000 rows over 3 members also
    covers a ragged last shard)."""
    m = _mb()
    num_rows, row_len = 3001, 64
    table = np.random.default_rng(3).uniform(-1, 1, (num_rows, row_len)).astype(np.float32)
    for with_opt in (False, True):
        kv = m.B200(devices=devs)
        if with_opt:
            kv.set_optimizer(m.optimizer.SGD(learning_rate=0.1, momentum=0.9, wd=1e-3, rescale_grad=0.5))
        kv.init_row_sparse(0, torch.from_numpy(table).cuda())
        ref, mom = table.copy(), np.zeros_like(table)
        for step in range(2):
            rg = np.random.default_rng(900 + step)
            copies = []
            for _ in devs:
                idx = np.unique(rg.integers(0, num_rows, 400)).astype(np.int64)
                copies.append((idx, (rg.uniform(-1, 1, (len(idx), row_len)) * 1e-2).astype(np.float32)))
            kv.push(0, [m.RowSparse(torch.from_numpy(v).to("cuda:%d" % d), torch.from_numpy(i).to("cuda:%d" % d),
                                    (num_rows, row_len)) for (i, v), d in zip(copies, devs)])
            midx, mval = oracle.rsp_sum(copies)
            if with_opt:
                oracle.sgd_mom_rsp(ref, mom, midx, mval, 0.1, 1e-3, 0.9, 0.5)
            else:
                mg = kv.rsp_merged(0)
                assert mg.indices.cpu().numpy().tobytes() == midx.tobytes()
                assert mg.data.cpu().numpy().tobytes() == mval.tobytes()
        if with_opt:
            for d in sorted(set(devs)):         # any device may pull: rows come from every member's shard
                out = m.RowSparse(None, None, (0, 0))
                kv.row_sparse_pull(0, out=out, row_ids=torch.arange(num_rows, device="cuda:%d" % d))
                assert out.data.cpu().numpy().tobytes() == ref.tobytes()


def test_group_save_load_states(tmp_path, oracle):
    """Sharded optimizer state of a group gathers into / scatters from the reference's checkpoint layout."""
    m = _mb()
    shp = [(10, 10), (4097,)]

    def make():
        kv = m.B200(devices=[0, 0])
        kv.set_optimizer(m.optimizer.Adam(learning_rate=0.01, multi_precision=True))
        w = [[torch.linspace(-1, 1, int(np.prod(s)), device="cuda").half().view(s) for _ in range(2)] for s in shp]
        kv.broadcast([0, 1], [c[0] for c in w], w)
        return kv, w
    g = [[torch.full(s, 0.01 * (r + 1), device="cuda").half() for r in range(2)] for s in shp]
    kv1, w1 = make()
    for _ in range(2):
        kv1.pushpull([0, 1], g, out=w1)
    f = str(tmp_path / "states")
    kv1.save_optimizer_states(f, dump_optimizer=True)
    kv2, w2 = make()
    kv2.load_optimizer_states(f)
    kv2.pull([0, 1], out=w2)
    kv2.wait()
    for a, b in zip(w1, w2):            # the loaded fp32 masters define the fp16 value on every member
        for x, y in zip(a, b):
            assert torch.equal(x, y)
    kv1.pushpull([0, 1], g, out=w1)
    kv2.pushpull([0, 1], g, out=w2)
    kv1.wait(); kv2.wait()
    for a, b in zip(w1, w2):
        for x, y in zip(a, b):
            assert torch.equal(x, y)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_lazy_device_discovery_makes_a_group(oracle):
    """mx.kv.create('b200') takes no arguments: init values are held back until a call names the
    per-device arrays, then a group over those devices comes up (trainer.py:155-160)."""
    m = _mb()
    kv = m.create("b200")
    kv.set_optimizer(m.optimizer.SGD(learning_rate=1.0))
    kv.init(0, torch.zeros(16, device="cuda:0"))
    outs = [torch.empty(16, device="cuda:0"), torch.empty(16, device="cuda:1")]
    kv.pushpull(0, [torch.ones(16, device="cuda:0"), torch.ones(16, device="cuda:1")], out=outs)
    kv.wait()
    assert kv.devices == [0, 1]
    assert all((o == -2).all() for o in outs)
