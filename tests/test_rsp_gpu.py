"""Row-sparse path (BASELINE config C5): ReduceRowSparse + lazy row update on push, unique + row
gather on row_sparse_pull, through the kvstore plugin API.  Index arrays must be BIT-EXACT
(int64, sorted, unique); values are bit-exact too because rows are accumulated in the reference's
device order (src/ndarray/ndarray_function.cc:81-113)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _kv(table, opt=None):
    import mxnet_maintenance_b200 as mb
    kv = mb.create("b200")
    if opt is not None:
        kv.set_optimizer(opt)
    kv.init_row_sparse(0, torch.from_numpy(table).cuda())
    return kv, mb


def _grads(rng, num_rows, row_len, ncopies, nnr, zipf=False):
    out = []
    for _ in range(ncopies):
        if zipf:
            raw = np.minimum(rng.zipf(1.05, nnr) - 1, num_rows - 1)
        else:
            raw = rng.integers(0, num_rows, nnr)
        idx = np.unique(raw).astype(np.int64)           # row-sparse invariant: sorted unique
        val = (rng.uniform(-1, 1, (len(idx), row_len)) * 1e-2).astype(np.float32)
        out.append((idx, val))
    return out


def _push(kv, mb, grads):
    vals = [mb.RowSparse(torch.from_numpy(v).cuda(), torch.from_numpy(i).cuda(), (0, v.shape[1]))
            for i, v in grads]
    kv.push(0, vals)


def _pull(kv, mb, ids):
    out = mb.RowSparse(None, None, (0, 0))
    kv.row_sparse_pull(0, out=out, row_ids=ids)
    return out.indices.cpu().numpy(), out.data.cpu().numpy()


@pytest.mark.parametrize("ncopies", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("row_len", [512, 3])
def test_reduce_row_sparse_no_updater(oracle, ncopies, row_len):
    rng = np.random.default_rng(ncopies * 10 + row_len)
    num_rows = 5000
    table = rng.uniform(-1, 1, (num_rows, row_len)).astype(np.float32)
    kv, mb = _kv(table)
    grads = _grads(rng, num_rows, row_len, ncopies, 700)
    _push(kv, mb, grads)
    merged = kv.rsp_merged(0)
    ref_idx, ref_val = oracle.rsp_sum(grads)
    got_idx = merged.indices.cpu().numpy()
    assert got_idx.dtype == np.int64 and got_idx.tobytes() == ref_idx.tobytes()
    assert merged.data.cpu().numpy().tobytes() == ref_val.tobytes()
    # the merged value replaced the stored one: pulling absent rows gives zeros, all ids are kept
    ids = rng.integers(0, num_rows, (37, 11))
    pi, pv = _pull(kv, mb, torch.from_numpy(ids).cuda())
    uniq = oracle.unique(ids)
    ri, rv = oracle.sparse_retain(ref_idx, ref_val, uniq)
    assert pi.tobytes() == uniq.tobytes() == ri.tobytes()
    assert pv.tobytes() == rv.tobytes()


@pytest.mark.parametrize("dtype", [torch.int64, torch.int32, torch.float32])
def test_row_sparse_pull_unique_ids(oracle, dtype):
    """test_kvstore.py:69-96: random duplicate ids of any shape / dtype; all rows of the table."""
    rng = np.random.default_rng(3)
    num_rows, row_len = 4099, 64
    table = rng.uniform(-1, 1, (num_rows, row_len)).astype(np.float32)
    kv, mb = _kv(table)
    for shape in [(1,), (5, 7), (3000,), (4096,), (4097,), (20000,)]:
        ids = rng.integers(0, num_rows, shape)
        pi, pv = _pull(kv, mb, torch.from_numpy(ids).cuda().to(dtype))
        uniq = oracle.unique(ids)
        assert pi.dtype == np.int64 and pi.tobytes() == uniq.tobytes()
        assert pv.tobytes() == table[uniq].tobytes()


def test_large_row_id(oracle):
    """tests/python/gpu/test_kvstore_gpu.py:126-134: row id 793470."""
    num_rows, row_len = 793471 + 5, 8
    table = np.zeros((num_rows, row_len), np.float32)
    table[793470] = 7.0
    kv, mb = _kv(table)
    pi, pv = _pull(kv, mb, torch.tensor([793470, 3, 793470], device="cuda"))
    assert pi.tolist() == [3, 793470] and (pv[1] == 7.0).all() and (pv[0] == 0).all()


@pytest.mark.parametrize("kind", ["sgd", "sgd_nomom", "adam", "sgd_clip", "adam_clip", "adagrad", "adagrad_clip"])
def test_lazy_row_update(oracle, kind):
    import mxnet_maintenance_b200 as mb0
    rng = np.random.default_rng(11)
    num_rows, row_len = 3000, 128
    table = rng.uniform(-1, 1, (num_rows, row_len)).astype(np.float32)
    clip = 0.004 if "clip" in kind else None
    if kind.startswith("sgd"):
        mom = 0.0 if "nomom" in kind else 0.9
        opt = mb0.optimizer.SGD(learning_rate=0.1, momentum=mom, wd=1e-3, rescale_grad=0.5, clip_gradient=clip)
    elif kind.startswith("adagrad"):
        opt = mb0.optimizer.AdaGrad(learning_rate=0.05, eps=1e-7, rescale_grad=0.5, clip_gradient=clip)
    else:
        opt = mb0.optimizer.Adam(learning_rate=1e-3, wd=0.01, rescale_grad=0.5, clip_gradient=clip)
    kv, mb = _kv(table, opt)
    ref = table.copy()
    s0, s1 = np.zeros_like(ref), np.zeros_like(ref)
    from oracle.oracle import adam_lr
    for step in range(3):
        grads = _grads(rng, num_rows, row_len, 4, 500)
        _push(kv, mb, grads)
        midx, mval = oracle.rsp_sum(grads)
        if kind.startswith("sgd") and mom == 0.0:
            # momentum-free sparse SGD is SGDDnsRspKernel: (1 - lr*wd)*w - lr*g (optimizer_op-inl.h:452-475)
            oracle.sgd_rsp(ref, midx, mval, 0.1, 1e-3, 0.5, clip if clip else -1.0)
        elif kind.startswith("sgd"):
            oracle.sgd_mom_rsp(ref, s0, midx, mval, 0.1, 1e-3, mom, 0.5, clip if clip else -1.0)
        elif kind.startswith("adagrad"):
            oracle.adagrad_rsp(ref, s0, midx, mval, 0.05, 1e-7, 0.5, clip if clip else -1.0)
        else:
            oracle.adam_rsp(ref, s0, s1, midx, mval, adam_lr(1e-3, 0.9, 0.999, step + 1), 0.01, rescale=0.5,
                            clip=clip if clip else -1.0)
    pi, pv = _pull(kv, mb, torch.arange(num_rows, device="cuda"))
    assert pi.tolist() == list(range(num_rows))
    assert pv.tobytes() == ref.tobytes()


def test_empty_and_uninitialised_inputs(oracle):
    rng = np.random.default_rng(5)
    num_rows, row_len = 100, 16
    table = rng.uniform(-1, 1, (num_rows, row_len)).astype(np.float32)
    kv, mb = _kv(table)
    e = (np.zeros(0, np.int64), np.zeros((0, row_len), np.float32))
    g = _grads(rng, num_rows, row_len, 1, 10)[0]
    _push(kv, mb, [e, g, e])          # uninitialised copies are skipped (ndarray_function.cc:81,129)
    m = kv.rsp_merged(0)
    assert m.indices.cpu().numpy().tobytes() == g[0].tobytes()
    assert m.data.cpu().numpy().tobytes() == g[1].tobytes()
    _push(kv, mb, [e, e])             # nothing initialised -> empty stored value (ndarray_function.cu:113-123)
    assert kv.rsp_merged(0).indices.numel() == 0
    pi, pv = _pull(kv, mb, torch.tensor([5, 1], device="cuda"))
    assert pi.tolist() == [1, 5] and (pv == 0).all()


def test_c5_full_size(oracle):
    """C5: 1,048,576 x 512 table, 65,536 Zipf ids per device with heavy duplicates, 4 devices,
    lazy SGD-momentum; then row_sparse_pull with the raw (unsorted, duplicated) ids."""
    import mxnet_maintenance_b200 as mb0
    rng = np.random.default_rng(1234)
    num_rows, row_len = 1048576, 512
    table = np.zeros((num_rows, row_len), np.float32)
    hot = rng.uniform(-1, 1, (4096, row_len)).astype(np.float32)
    table[:4096] = hot
    kv, mb = _kv(table, mb0.optimizer.SGD(learning_rate=0.1, momentum=0.9, wd=1e-4))
    grads = _grads(rng, num_rows, row_len, 4, 65536, zipf=True)
    _push(kv, mb, grads)
    midx, mval = oracle.rsp_sum(grads)
    raw = np.minimum(rng.zipf(1.05, 65536) - 1, num_rows - 1)
    pi, pv = _pull(kv, mb, torch.from_numpy(raw).cuda())
    uniq = np.unique(raw)
    assert pi.tobytes() == uniq.astype(np.int64).tobytes()
    # oracle update restricted to the touched rows
    rows = np.union1d(midx, uniq)
    sub = table[rows].copy()
    mom = np.zeros_like(sub)
    pos = np.searchsorted(rows, midx).astype(np.int64)
    oracle.sgd_mom_rsp(sub, mom, pos, mval, 0.1, 1e-4, 0.9)
    assert pv.tobytes() == sub[np.searchsorted(rows, uniq)].tobytes()


def test_pull_with_millions_of_ids(oracle):
    """> 1M / > 2M ids exercise the separate-scan variants of the sort and unique kernels."""
    rng = np.random.default_rng(77)
    num_rows, row_len = 300000, 4
    table = rng.uniform(-1, 1, (num_rows, row_len)).astype(np.float32)
    kv, mb = _kv(table)
    for n in (1_100_000, 2_300_000):
        ids = rng.integers(0, num_rows, n)
        pi, pv = _pull(kv, mb, torch.from_numpy(ids).cuda())
        uniq = np.unique(ids).astype(np.int64)
        assert pi.tobytes() == uniq.tobytes()
        assert pv.tobytes() == table[uniq].tobytes()


def test_row_ids_are_validated(oracle):
    """A pushed id outside the table (or an unsorted / duplicated index list) must not corrupt memory:
    the rows are skipped on the device and the fault is reported at the next synchronisation point;
    a pulled id outside the table is an error."""
    import mxnet_maintenance_b200 as mb0
    rng = np.random.default_rng(9)
    num_rows, row_len = 1000, 8
    table = rng.uniform(-1, 1, (num_rows, row_len)).astype(np.float32)
    for ncopies in (1, 2):
        kv, mb = _kv(table, mb0.optimizer.SGD(learning_rate=0.1))
        good = (np.array([1, 5, 9], np.int64), np.ones((3, row_len), np.float32))
        bad = (np.array([2, 7, num_rows + 3], np.int64), np.ones((3, row_len), np.float32))
        _push(kv, mb, [good, bad][:ncopies] if ncopies == 2 else [bad])
        with pytest.raises(mb.B200KVError, match="outside"):
            kv.wait()
        kv.wait()                                       # reported once
        unsorted = (np.array([9, 5, 5], np.int64), np.ones((3, row_len), np.float32))
        _push(kv, mb, [unsorted] + ([good] if ncopies == 2 else []))
        with pytest.raises(mb.B200KVError, match="sorted and unique"):
            kv.wait()
        with pytest.raises(mb.B200KVError, match="outside"):
            _pull(kv, mb, torch.tensor([3, num_rows], device="cuda"))
        pi, pv = _pull(kv, mb, torch.tensor([0, 999], device="cuda"))     # untouched rows are intact
        assert pi.tolist() == [0, 999] and pv.tobytes() == table[[0, 999]].tobytes()


def test_row_sparse_pull_async(oracle):
    """sync=False leaves the unique-row count on the device (no host synchronisation in the call)."""
    rng = np.random.default_rng(11)
    num_rows, row_len = 4099, 32
    table = rng.uniform(-1, 1, (num_rows, row_len)).astype(np.float32)
    kv, mb = _kv(table)
    ids = rng.integers(0, num_rows, 5000)
    out = mb.RowSparse(None, None, (0, 0))
    kv.row_sparse_pull(0, out=out, row_ids=torch.from_numpy(ids).cuda(), sync=False)
    n = int(out.nnr.item())
    uniq = np.unique(ids).astype(np.int64)
    assert n == len(uniq)
    assert out.indices[:n].cpu().numpy().tobytes() == uniq.tobytes()
    assert out.data[:n].cpu().numpy().tobytes() == table[uniq].tobytes()


@pytest.mark.parametrize("kind", ["sgd", "sgd_clip", "adam", "adam_clip"])
@pytest.mark.parametrize("ncopies", [1, 3])
def test_standard_row_update(oracle, kind, ncopies):
    """lazy_update=False: EVERY row of the table is updated, rows absent from the merged gradient with a
    zero gradient (SGDMomStdDnsRspDnsKernel / AdamStdDnsRspDnsKernel, optimizer_op.cc:108-286)."""
    import mxnet_maintenance_b200 as mb0
    from oracle.oracle import adam_lr
    rng = np.random.default_rng(21)
    num_rows, row_len = 777, 24
    table = rng.uniform(-1, 1, (num_rows, row_len)).astype(np.float32)
    clip = 0.004 if "clip" in kind else None
    if kind.startswith("sgd"):
        opt = mb0.optimizer.SGD(learning_rate=0.1, momentum=0.9, wd=1e-3, rescale_grad=0.5, clip_gradient=clip,
                                lazy_update=False)
    else:
        opt = mb0.optimizer.Adam(learning_rate=1e-3, wd=0.01, rescale_grad=0.5, clip_gradient=clip, lazy_update=False)
    kv, mb = _kv(table, opt)
    ref = table.copy()
    s0, s1 = np.zeros_like(ref), np.zeros_like(ref)
    for step in range(3):
        grads = _grads(rng, num_rows, row_len, ncopies, 60)
        _push(kv, mb, grads)
        midx, mval = oracle.rsp_sum(grads)
        if kind.startswith("sgd"):
            oracle.sgd_mom_std_rsp(ref, s0, midx, mval, 0.1, 1e-3, 0.9, 0.5, clip if clip else -1.0)
        else:
            oracle.adam_std_rsp(ref, s0, s1, midx, mval, adam_lr(1e-3, 0.9, 0.999, step + 1), 0.01, rescale=0.5,
                                clip=clip if clip else -1.0)
    pi, pv = _pull(kv, mb, torch.arange(num_rows, device="cuda"))
    assert pv.tobytes() == ref.tobytes()
