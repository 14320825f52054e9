"""kvstore API semantics through the plugin class, following the reference's own unit tests
(tests/python/unittest/test_kvstore.py, test_kvstore_custom.py, test_gluon_trainer.py) with torch
CUDA tensors standing in for NDArrays."""
import os
import pickle

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

SHAPE = (4, 4)
KEYS = [5, 7, 11]
STR_KEYS = ["b", "c", "d"]


def mb():
    import mxnet_maintenance_b200 as m
    return m


def zeros():
    return torch.zeros(SHAPE, device="cuda")


def ones(v=1.0):
    return torch.full(SHAPE, float(v), device="cuda")


def init_kv(str_keys=False):
    kv = mb().create("b200")
    if str_keys:
        kv.init("a", zeros())
        kv.init(STR_KEYS, [zeros() for _ in STR_KEYS])
    else:
        kv.init(3, zeros())
        kv.init(KEYS, [zeros() for _ in KEYS])
    return kv


@pytest.mark.parametrize("s", [False, True])
def test_single_kv_pair(s):
    """test_kvstore.py:55-67"""
    kv = init_kv(s)
    key = "a" if s else 3
    kv.push(key, ones())
    val = torch.empty(SHAPE, device="cuda")
    kv.pull(key, out=val)
    assert (val == 1).all()


@pytest.mark.parametrize("s", [False, True])
def test_list_kv_pair_and_aggregator(s):
    """test_kvstore.py:98-175: lists of keys; 4 device copies sum to 4, ones*2 to 8."""
    kv = init_kv(s)
    key, keys = ("a", STR_KEYS) if s else (3, KEYS)
    kv.push(keys, [ones(4)] * len(keys))
    val = [torch.empty(SHAPE, device="cuda") for _ in keys]
    kv.pull(keys, out=val)
    assert all((v == 4).all() for v in val)
    ndev = 4
    vals = [ones() for _ in range(ndev)]
    outs = [torch.empty(SHAPE, device="cuda") for _ in range(ndev)]
    kv.push(key, vals)
    kv.pull(key, out=outs)
    assert all((o == ndev).all() for o in outs)
    vals = [[ones(2) for _ in range(ndev)] for _ in keys]
    outs = [[torch.empty(SHAPE, device="cuda") for _ in range(ndev)] for _ in keys]
    kv.push(keys, vals)
    kv.pull(keys, out=outs)
    assert all((o == ndev * 2).all() for out in outs for o in out)


def test_pushpull_in_place_and_broadcast():
    """test_kvstore_custom.py:38-140: broadcast, pushpull with / without out (in-place semantics)."""
    kv = mb().create("b200")
    outs = [torch.empty(SHAPE, device="cuda") for _ in range(3)]
    kv.broadcast("w", ones(3), out=outs)
    assert all((o == 3).all() for o in outs)
    vals = [ones(1) for _ in range(4)]
    kv.pushpull("w", vals)                       # out=None: results land in the values
    assert all((v == 4).all() for v in vals)
    out = torch.empty(SHAPE, device="cuda")
    kv.pushpull("w", [ones(2), ones(5)], out=out)
    assert (out == 7).all()
    assert kv.type == "b200" and kv.rank == 0 and kv.num_workers == 1
    assert kv.is_capable("optimizer") is True
    with pytest.raises(ValueError):
        kv.is_capable("no-such-capability")


def test_errors_follow_the_reference():
    """kvstore_local.h:229-232,289,322-325,453-454; test_kvstore.py:288-346."""
    m = mb()
    kv = init_kv()
    with pytest.raises(m.B200KVError, match="duplicate init"):
        kv.init(3, zeros())
    with pytest.raises(m.B200KVError, match="has not been initialized"):
        kv.pull(99, out=zeros())
    with pytest.raises(m.B200KVError, match="has not been initialized"):
        kv.push(99, ones())
    with pytest.raises(m.B200KVError, match="mixing int and string"):
        kv.init("x", zeros())
    with pytest.raises(m.B200KVError, match="elements"):
        kv.push(3, torch.ones(5, device="cuda"))
    with pytest.raises(m.B200KVError):
        kv.push(3, torch.ones(SHAPE))             # CPU tensor: no CPU path
    with pytest.raises(m.B200KVError, match="compression"):
        kv.set_gradient_compression({"type": "2bit"})
    kvs = init_kv(True)
    with pytest.raises(m.B200KVError, match="has not been initialized"):
        kvs.pull("zzz", out=zeros())


def test_default_sgd_docstring_example():
    """kvstore.py:569-581: kv.set_optimizer(SGD(lr=.01)); push(ones) from weight 0 -> -0.01... wait for
    rescale 1: weight = 0 - 0.01*1."""
    m = mb()
    kv = m.create("b200")
    w = zeros()
    kv.init("3", w)
    kv.set_optimizer(m.optimizer.SGD(learning_rate=0.01))
    kv.push("3", ones())
    kv.pull("3", out=w)
    assert torch.equal(w, torch.full(SHAPE, np.float32(0) - np.float32(0.01) * np.float32(1.0), device="cuda"))


def test_trainer_flow_known_answer():
    """test_gluon_trainer.py:84-103 replayed the way Trainer drives a kvstore (trainer.py:141-162,
    371-395): broadcast to all device copies, per-key pushpull(i, grads, out=weights, priority=-i),
    rescale_grad = 1/batch_size set every step; 2 device copies, lr 1, momentum .5 -> -2, then
    lr_mult .5 -> -4."""
    m = mb()

    class Param:
        lr_mult, wd_mult = 1.0, 1.0
    p = Param()
    opt = m.optimizer.SGD(learning_rate=1.0, momentum=0.5, param_dict={0: p})
    kv = m.create("b200")
    assert kv.is_capable("optimizer")
    kv.set_optimizer(opt)                                   # before any init/broadcast (trainer.py:254-256)
    copies = [torch.zeros(10, device="cuda") for _ in range(2)]
    kv.broadcast(0, copies[0], out=copies)
    for expect in (-2.0, -4.0):
        opt.rescale_grad = 1.0 / 1                           # trainer.py:338-339
        grads = [torch.ones(10, device="cuda") for _ in range(2)]
        kv.pushpull(0, grads, out=copies, priority=0)
        assert all((c == expect).all() for c in copies)
        p.lr_mult = 0.5


@pytest.mark.parametrize("kind,mp", [("sgd", False), ("sgd", True), ("adam", False), ("lamb", True)])
def test_save_load_optimizer_states(tmp_path, kind, mp):
    """Trainer.save_states / load_states contract (trainer.py:468-527; kvstore.py:641-690): the pickle
    holds {index: state} in the reference's tuple layout (+ the optimizer when dump_optimizer)."""
    m = mb()
    dt = torch.float16 if mp else torch.float32
    O = {"sgd": m.optimizer.SGD, "adam": m.optimizer.Adam, "lamb": m.optimizer.LAMB}[kind]
    kw = dict(momentum=0.9) if kind == "sgd" else {}

    def make():
        kv = m.create("b200")
        kv.set_optimizer(O(learning_rate=0.01, multi_precision=mp, **kw))
        w = [torch.linspace(-1, 1, 100, device="cuda").to(dt).view(10, 10), torch.linspace(0, 1, 7, device="cuda").to(dt)]
        kv.broadcast([0, 1], w, out=[[x] for x in w])
        return kv, w
    g = [torch.full((10, 10), 0.01, device="cuda").to(dt), torch.full((7,), -0.02, device="cuda").to(dt)]
    kv1, w1 = make()
    for _ in range(2):
        kv1.pushpull([0, 1], g, out=[[x] for x in w1])
    f = str(tmp_path / "states")
    kv1.save_optimizer_states(f, dump_optimizer=True)
    states, opt = pickle.loads(open(f, "rb").read())
    st = states[0]
    # every state array has the WEIGHT's shape (zeros(weight.shape), optimizer.py:583-600,1602-1607),
    # so a native store / Trainer can load the file
    if kind == "sgd":
        assert (isinstance(st, tuple) and len(st) == 2 and st[1].dtype == np.float32 and
                st[0].shape == (10, 10) and st[1].shape == (10, 10)) if mp else st.shape == (10, 10)
    else:
        assert (isinstance(st[1], tuple) and st[0].shape == (10, 10) and st[1][0].shape == (10, 10)) if mp \
            else (len(st) == 2 and st[0].shape == (10, 10) and st[1].shape == (10, 10))
    assert states[1].shape == (7,) if (kind == "sgd" and not mp) else True
    assert opt.num_update == 2
    # continue in a fresh store from the checkpoint == continue in the original
    kv2, w2 = make()
    for a, b in zip(w2, w1):
        a.copy_(b)
    kv2.init([10, 11], w2) if False else None
    kv2b = m.create("b200")
    kv2b.set_optimizer(O(learning_rate=0.01, multi_precision=mp, **kw))
    kv2b.broadcast([0, 1], w1, out=[[x] for x in w2])
    kv2b.load_optimizer_states(f)
    assert kv2b._updater.optimizer.num_update == 2          # trainer.py:517-518 reads _updater.optimizer
    kv1.pushpull([0, 1], g, out=[[x] for x in w1])
    kv2b.pushpull([0, 1], g, out=[[x] for x in w2])
    kv1.wait(); kv2b.wait()
    for a, b in zip(w1, w2):
        assert torch.equal(a, b)


def test_pull_ignores_row_sparse_outs():
    """kvstore_local.h:374-385: pull with ignore_sparse=True silently skips row_sparse outputs."""
    m = mb()
    kv = init_kv()
    rsp = m.RowSparse(None, None, SHAPE)
    kv.pull(3, out=rsp)                                      # no error, nothing written
    assert rsp.data is None
    with pytest.raises(m.B200KVError):
        kv.pull(3, out=rsp, ignore_sparse=False)


def test_auto_batch_per_key_trainer_loop(oracle=None):
    """gluon.Trainer's per-parameter loop (trainer.py:392-395: pushpull(i, grads, out=weights,
    priority=-i)) under auto_batch: the calls are fused into one launch per step and give the same
    bits as immediate per-key calls."""
    m = mb()
    from mxnet_maintenance_b200 import shapes
    shp = shapes.mlp() + [(33,), (4097,)]
    rng = np.random.default_rng(3)
    w0 = [rng.uniform(-1, 1, s).astype(np.float32) for s in shp]
    gs = [[(rng.uniform(-1, 1, s) * 1e-2).astype(np.float32) for s in shp] for _ in range(4)]
    res, launches = [], []
    for auto in (False, True):
        kv = m.create("b200")
        kv.auto_batch = auto
        kv.set_optimizer(m.optimizer.SGD(learning_rate=0.1, momentum=0.9, wd=1e-4))
        w = [torch.from_numpy(x).cuda() for x in w0]
        g = [torch.empty_like(x) for x in w]                     # persistent gradient buffers
        for i, x in enumerate(w):
            kv.broadcast(i, x, out=[x])
        kv.wait()
        base = kv.launch_count()
        for it in range(4):
            for d, h in zip(g, gs[it]):
                d.copy_(torch.from_numpy(h))
            for i in range(len(w)):
                kv.pushpull(i, [g[i]], out=[w[i]], priority=-i)
        kv.wait()
        launches.append(kv.launch_count() - base)
        res.append([x.cpu().numpy() for x in w])
    for a, b in zip(*res):
        assert a.tobytes() == b.tobytes()
    assert launches[0] == 4 * len(shp) and launches[1] == 4     # one fused launch per step
    # a partial set of keys is executed at the next flush point (wait / pull / ...)
    kv.pushpull(0, [g[0]], out=[w[0]])
    before = w[0].clone()
    kv.wait()
    assert not torch.equal(before, w[0])


def test_handle_cache_does_not_pin_temporaries():
    """Exported DLPack capsules hold a reference to the tensor storage; the per-store handle cache
    must let go of them when the tensor dies (row_sparse_pull allocates fresh outputs every call)."""
    import gc
    m = mb()
    kv = m.create("b200")
    w = torch.zeros(1 << 20, device="cuda")
    kv.init(0, w)
    torch.cuda.synchronize()
    gc.collect()
    base = torch.cuda.memory_allocated()
    for _ in range(20):
        t = torch.empty(8 << 20, device="cuda")          # 32 MB temporary pulled into
        kv.pull(0, out=torch.empty(1 << 20, device="cuda"))
        kv._handles([t], True)
        del t
    kv.wait()
    gc.collect()
    assert torch.cuda.memory_allocated() - base < (40 << 20)


@pytest.mark.parametrize("str_keys", [False, True])
def test_custom_updater(str_keys):
    """test_kvstore.py:226-280 (test_updater): ``kv._set_updater(lambda key, recv, local: local += recv)``;
    every push of ones from 4 device copies adds 4 to the stored value (kvstore_local.h:259-267)."""
    kv = init_kv(str_keys)
    key, keys = ("a", STR_KEYS) if str_keys else (3, KEYS)
    seen = []

    def updater(k, recv, local):
        seen.append(k)
        local += recv
    kv._set_updater(updater)
    ndev = 4
    vals = [ones() for _ in range(ndev)]
    outs = [torch.empty(SHAPE, device="cuda") for _ in range(ndev)]
    kv.push(key, vals)
    kv.pull(key, out=outs)
    assert all((o == ndev).all() for o in outs) and seen == [key]
    for i in range(1, 4):                                   # accumulates over pushes
        kv.push(keys, [[ones() for _ in range(ndev)] for _ in keys])
        got = [[torch.empty(SHAPE, device="cuda") for _ in range(ndev)] for _ in keys]
        kv.pull(keys, out=got)
        assert all((o == ndev * i).all() for row in got for o in row)
    out = torch.empty(SHAPE, device="cuda")
    kv.pushpull(key, [ones(2), ones(3)], out=out)           # pushpull runs the updater too
    assert (out == ndev + 5).all()


def test_custom_updater_on_a_group():
    """The same updater over a store that drives several (virtual) devices: the reduce is the fused
    peer-memory exchange, the callback runs on every device copy of the stored value."""
    m = mb()
    kv = m.B200(devices=[0, 0])
    w = [torch.zeros(SHAPE, device="cuda") for _ in range(2)]
    kv.broadcast(7, w[0], out=w)

    def nightly_test_optimizer(k, recv, local):             # tests/nightly/test_kvstore.py:80-99: w += rescale * g
        local += 0.5 * recv
    kv._set_updater(nightly_test_optimizer)
    for i in range(1, 4):
        kv.pushpull(7, [ones(), ones(3)], out=w)
        kv.wait()
        assert all((c == 2.0 * i).all() for c in w)


def test_priority_buckets_launch_early():
    """trainer.py:392-395 pushes parameter i with priority -i.  With auto_batch and auto_buckets = 3
    a bucket is launched the moment its last key has been handed over — before the remaining keys
    arrive — and the results equal the unbucketed run bit for bit."""
    m = mb()
    from mxnet_maintenance_b200 import shapes
    shp = shapes.mlp() + [(33,), (4097,), (70000,)]
    rng = np.random.default_rng(4)
    w0 = [rng.uniform(-1, 1, s).astype(np.float32) for s in shp]
    gs = [[(rng.uniform(-1, 1, s) * 1e-2).astype(np.float32) for s in shp] for _ in range(3)]
    res = []
    for buckets in (1, 3):
        kv = m.create("b200")
        kv.auto_batch, kv.auto_buckets = True, buckets
        kv.set_optimizer(m.optimizer.SGD(learning_rate=0.1, momentum=0.9, wd=1e-4))
        w = [torch.from_numpy(x).cuda() for x in w0]
        g = [torch.empty_like(x) for x in w]
        for i, x in enumerate(w):
            kv.broadcast(i, x, out=[x])
        kv.wait()
        early = []
        for it in range(3):
            for d, h in zip(g, gs[it]):
                d.copy_(torch.from_numpy(h))
            base = kv.launch_count()
            for i in range(len(w)):
                kv.pushpull(i, [g[i]], out=[w[i]], priority=-i)
                if i == len(w) - 2:
                    early.append(kv.launch_count() - base)      # launches before the LAST key was pushed
            kv.wait()
        res.append([x.cpu().numpy() for x in w])
        if buckets == 1:
            assert early == [0, 0, 0]
        else:
            assert early[0] == 0 and early[1] >= 1 and early[2] >= 1   # buckets are known from step 2 on
    for a, b in zip(*res):
        assert a.tobytes() == b.tobytes()
