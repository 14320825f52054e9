"""Host-side logic that needs no GPU: plugin registry, key/value flattening, optimizer scalar
plumbing (python/mxnet/optimizer/optimizer.py:411-508,1609-1619), shape lists, the oracle's
KVStoreLocal model against the reference's closed-form known answers."""
import math

import numpy as np
import pytest

import mxnet_maintenance_b200 as mb
from mxnet_maintenance_b200 import optimizer as opt, shapes
from mxnet_maintenance_b200.kvstore import _flatten


def test_registry_and_create():
    # KVStoreBase.register lower-cases the class name (base.py:219-243); create() looks it up
    assert mb.KVStoreBase.kv_registry["b200"] is mb.B200
    assert issubclass(mb.B200, mb.KVStoreBase)
    with pytest.raises(ValueError):
        mb.create("no-such-store")
    with pytest.warns(UserWarning):       # overriding an existing name warns, like the reference
        @mb.KVStoreBase.register
        class B200(mb.B200):              # noqa: F811
            pass
    mb.KVStoreBase.kv_registry["b200"] = mb.B200


def test_plugin_interface_methods_exist():
    for name in ("broadcast", "pushpull", "set_optimizer", "is_capable", "save_optimizer_states",
                 "load_optimizer_states", "init", "push", "pull", "row_sparse_pull",
                 "set_gradient_compression", "_barrier"):
        assert callable(getattr(mb.B200, name))
    for prop in ("type", "rank", "num_workers"):
        assert isinstance(getattr(mb.B200, prop), property)
    assert mb.KVStoreBase.OPTIMIZER == "optimizer"


class _Arr:  # minimal array stand-in
    def __init__(self):
        self.__dlpack__ = lambda: None


def test_flatten_follows_ctype_key_value():
    a, b, c = _Arr(), _Arr(), _Arr()
    assert _flatten(3, a) == ([3], [a], False)
    assert _flatten(3, [a, b]) == ([3, 3], [a, b], False)            # device copies of one key
    assert _flatten([3, 5], [a, b]) == ([3, 5], [a, b], False)
    assert _flatten(["x", "y"], [[a, b], [c]]) == (["x", "x", "y"], [a, b, c], True)
    with pytest.raises(AssertionError):
        _flatten([3, "x"], [a, b])                                   # mixed key types


class _P:
    def __init__(self, lr_mult=1.0, wd_mult=1.0):
        self.lr_mult, self.wd_mult = lr_mult, wd_mult


def test_resolve_step_counts_scheduler_multipliers():
    class Sched:
        base_lr = 0.1

        def __call__(self, n):
            return 0.1 * (0.5 ** n)
    o = opt.SGD(learning_rate=0.1, lr_scheduler=Sched(), wd=0.01, momentum=0.9,
                param_dict={0: _P(2.0, 0.0), 1: _P()})
    lrs, wds, steps = opt.resolve_step(o, [0, 1])
    assert steps == [1, 1] and o.num_update == 1
    assert lrs == [0.1 * 0.5 * 2.0, 0.1 * 0.5] and wds == [0.0, 0.01]
    lrs, wds, steps = opt.resolve_step(o, [1])
    assert steps == [2] and o.num_update == 2 and lrs == [0.1 * 0.25]
    # idx2name-based wd_mult: names not ending in _weight/_gamma get 0 (optimizer.py:375-397)
    o2 = opt.SGD(param_idx2name={0: "fc_weight", 1: "fc_bias"}, wd=0.1)
    o2.set_wd_mult({})
    assert o2._get_wds([0, 1]) == [0.1, 0.0]


def test_adam_bias_correction_folded_into_lr():
    o = opt.Adam(learning_rate=1e-3)
    for t in (1, 2, 3):
        lrs, _, steps = opt.resolve_step(o, [7])
        assert steps == [t]
        assert lrs[0] == 1e-3 * (math.sqrt(1.0 - 0.999 ** t) / (1.0 - 0.9 ** t))   # `lr *= sqrt(c2)/c1`


def test_kind_of_and_unsupported():
    assert opt.kind_of(opt.LAMB()) == "lamb"

    assert opt.kind_of(opt.AdaGrad()) == "adagrad"

    class RMSProp(opt.Optimizer):
        pass
    with pytest.raises(ValueError):
        opt.kind_of(RMSProp())


def test_shape_lists_match_baseline_configs():
    assert len(shapes.mlp()) == 6 and sum(map(shapes.numel, shapes.mlp())) == 669706
    r = shapes.resnet50()
    assert len(r) == 161 and sum(map(shapes.numel, r)) == 25557032
    assert sum(1 for s in r if shapes.numel(s) <= 2048) == 107 and max(map(shapes.numel, r)) == 2359296
    assert len(shapes.bert_base()) == 199 and sum(map(shapes.numel, shapes.bert_base())) == 109482240
    assert len(shapes.bert_large()) == 391 and sum(map(shapes.numel, shapes.bert_large())) == 335141888


# ---- the reference's closed-form known answers, replayed on the CPU oracle -----------------------
def test_known_answer_trainer_two_devices(oracle):
    """tests/python/unittest/test_gluon_trainer.py:84-103: 2 contexts, w=0, grad=1 on each, lr 1,
    momentum .5 -> -2; then lr_mult .5 -> -4."""
    from oracle.oracle import LocalStoreModel
    m = LocalStoreModel(oracle, dict(kind="sgd", lr=1.0, momentum=0.5, rescale=1.0, lr_mult={}))
    m.init(0, np.zeros(10, np.float32))
    ones = [np.ones(10, np.float32), np.ones(10, np.float32)]
    assert (m.pushpull(0, ones) == -2).all()
    m.opt["lr_mult"] = {0: 0.5}
    assert (m.pushpull(0, ones) == -4).all()


def test_known_answer_aggregator_and_default_sgd(oracle):
    """test_kvstore.py:143-175 (4 devices of ones -> 4, ones*2 -> 8); kvstore.py docstring
    (:569-581): default SGD lr .01 on a gradient of ones from weight 0 -> -0.01... in fp32."""
    from oracle.oracle import LocalStoreModel
    m = LocalStoreModel(oracle, None)
    m.init(3, np.zeros((4, 4), np.float32))
    assert (m.pushpull(3, [np.ones((4, 4), np.float32)] * 4) == 4).all()
    assert (m.pushpull(3, [np.ones((4, 4), np.float32) * 2] * 4) == 8).all()
    m2 = LocalStoreModel(oracle, dict(kind="sgd", lr=0.01))
    m2.init(0, np.zeros(3, np.float32))
    out = m2.pushpull(0, [np.ones(3, np.float32)])
    assert (out == np.float32(0) + (np.float32(0) - np.float32(0.01) * np.float32(0) * np.float32(0)
                                    - np.float32(0.01) * np.float32(1.0) * np.float32(1.0))).all()
    with pytest.raises(KeyError):
        m2.init(0, np.zeros(3, np.float32))          # duplicate init (kvstore_local.h:229-232)


def test_known_answer_row_sparse(oracle):
    """test_kvstore.py:178-224: row-sparse aggregate == numpy sum; :69-96 pull retains requested rows."""
    rng = np.random.default_rng(0)
    dense = [np.where(rng.random((20, 1)) < 0.4, rng.uniform(-1, 1, (20, 6)), 0).astype(np.float32) for _ in range(4)]
    inputs = []
    for d in dense:
        idx = np.nonzero(d.any(axis=1))[0].astype(np.int64)
        inputs.append((idx, d[idx]))
    idx, val = oracle.rsp_sum(inputs)
    full = np.zeros((20, 6), np.float32)
    full[idx] = val
    expect = ((dense[0] + dense[1]) + dense[2]) + dense[3]
    np.testing.assert_allclose(full, expect, rtol=1e-6, atol=1e-7)
    ri, rv = oracle.sparse_retain(idx, val, oracle.unique([19, 0, 0, 5, 19]))
    assert ri.tolist() == [0, 5, 19]
    assert (rv == full[[0, 5, 19]]).all()


def test_bound_step_refresh_equals_resolve_step():
    """BoundStep._refresh (the vectorised per-step scalar path of bind_pushpull) must ship exactly the
    floats that the reference-order resolve_step produces — checked on the CPU with a fake library
    that records the b200kv_set_hyper arguments."""
    import ctypes as C
    from mxnet_maintenance_b200.kvstore import BoundStep

    class Sched:
        base_lr = 0.02

        def __call__(self, n):
            return self.base_lr * (0.9 ** n)

    class FakeLib:
        def __init__(self):
            self.calls = []

        def b200kv_set_hyper(self, h, n, keys, lrs, wds, steps):
            f = C.cast(lrs, C.POINTER(C.c_float))
            g = C.cast(wds, C.POINTER(C.c_float))
            self.calls.append(([f[i] for i in range(n)], [g[i] for i in range(n)], [steps[i] for i in range(n)]))
            return 0

        def b200kv_run_bound(self, h, bid):
            return 0

    class FakeKV:
        def __init__(self, optimizer):
            self._lib, self._h = FakeLib(), None
            self._updater = opt.Updater(optimizer)

        def _keys(self, keys, use_str, create=False):
            return (C.c_int * len(keys))(*keys)

        def _send_optimizer(self):
            pass

    keys = list(range(7))
    for kind in ("sgd", "adam", "lamb"):
        def make():
            O = {"sgd": opt.SGD, "adam": opt.Adam, "lamb": opt.LAMB}[kind]
            return O(learning_rate=0.02, lr_scheduler=Sched(), wd=1e-3,
                     param_dict={k: _P(1.0 + 0.25 * (k % 3), 0.5 * (k % 2)) for k in keys})
        o_ref, o_fast = make(), make()
        kv = FakeKV(o_fast)
        opt.resolve_step(o_fast, keys)            # what bind_pushpull does for the first replay
        step = BoundStep(kv, 1, keys, None)
        expect = [opt.resolve_step(o_ref, keys)]
        step()                                    # first replay: scalars already resolved by bind
        for it in range(3):
            if it == 1:
                o_ref.param_dict[2].lr_mult = 3.0
                o_fast.param_dict[2].lr_mult = 3.0
            lrs, wds, steps = opt.resolve_step(o_ref, keys)
            step()
            got = kv._lib.calls[-1]
            f32 = lambda xs: [C.c_float(x).value for x in xs]
            assert got[0] == f32(lrs) and got[1] == f32(wds) and got[2] == steps, (kind, it)
        assert o_fast.num_update == o_ref.num_update == 4
        assert o_fast._index_update_count == o_ref._index_update_count
