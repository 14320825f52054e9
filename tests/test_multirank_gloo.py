"""world_size-2 CPU tests (gloo) of the multi-rank HOST logic: both ranks derive the same segment
layout, their shards tile every key exactly once, and a sharded update (each rank updating only
the elements it owns, then exchanging) equals the single-process oracle update bit for bit.  The
per-element arithmetic here is the test's own use of the oracle as a checker — the product's
update runs only on the GPU."""
import os
import sys

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from mxnet_maintenance_b200 import shapes
    import layout_planner as planner
    from oracle.oracle import Oracle
    o = Oracle("port", nthread=1)
    shp = shapes.mlp() + [(1,), (63,), (65,), (4097,)]
    sizes = [shapes.numel(s) for s in shp]
    offs, total, sl = planner.segment_layout(sizes, world)
    # 1. every rank computes the same plan
    plan = torch.tensor(offs + [total, sl], dtype=torch.int64)
    ref = plan.clone()
    dist.broadcast(ref, src=0)
    assert torch.equal(plan, ref)
    # 2. shards tile each key exactly once
    mine = planner.shard_ranges(sizes, world, rank)
    cover = torch.zeros(sum(sizes), dtype=torch.int32)
    base = np.cumsum([0] + sizes[:-1])
    for (b, e), o0 in zip(mine, base):
        cover[o0 + b:o0 + e] += 1
    dist.all_reduce(cover)
    assert int(cover.min()) == 1 and int(cover.max()) == 1
    # 3. sharded SGD-momentum == whole-array oracle update (gradients summed over ranks first)
    rng = np.random.default_rng(5)
    w = [rng.uniform(-1, 1, n).astype(np.float32) for n in sizes]
    grads = [[(np.random.default_rng(100 * r + k).uniform(-1, 1, n) * 1e-2).astype(np.float32)
              for k, n in enumerate(sizes)] for r in range(world)]
    full_w = [x.copy() for x in w]
    full_m = [np.zeros(n, np.float32) for n in sizes]
    merged = [o.reduce_sum([grads[r][k].copy() for r in range(world)]) for k in range(len(sizes))]
    o.multi_sgd(full_w, merged, full_m, None, [0.1] * len(sizes), [1e-4] * len(sizes), 0.9, 0.5, -1.0)
    new_w = [torch.zeros(n) for n in sizes]
    for k, (b, e) in enumerate(mine):
        if e > b:
            t = torch.from_numpy(grads[rank][k]).clone()
            dist.all_reduce(t)                       # two ranks: a + b is order independent
        else:
            t = torch.from_numpy(grads[rank][k]).clone()
            dist.all_reduce(t)
        ws, ms, gs = w[k][b:e].copy(), np.zeros(e - b, np.float32), t.numpy()[b:e].copy()
        if e > b:
            o.multi_sgd([ws], [gs], [ms], None, [0.1], [1e-4], 0.9, 0.5, -1.0)
            new_w[k][b:e] = torch.from_numpy(ws)
    for k in range(len(sizes)):
        dist.all_reduce(new_w[k])                    # "all-gather" of disjoint shards
        assert new_w[k].numpy().tobytes() == full_w[k].tobytes(), "key %d" % k
    # 4. pieces() agrees with shard_ranges()
    for (k, r, b, n) in planner.pieces(sizes, world):
        if r == rank:
            assert mine[k] == (b, b + n)
    q.put((rank, "ok"))
    dist.destroy_process_group()


def test_two_rank_host_logic():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29000 + (os.getpid() % 500)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    got = sorted(q.get(timeout=5) for _ in range(2))
    assert got == [(0, "ok"), (1, "ok")]


def test_planner_matches_store_rules():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import layout_planner as planner
    offs, total, sl = planner.segment_layout([1, 64, 65, 1000], 8)
    assert offs == [0, 64, 128, 256] and total % (8 * 64) == 0 and sl * 8 == total
    assert planner.segment_layout([], 2)[1] == 128
    ps = planner.pieces([1000, 1000], 2)
    assert sum(p[3] for p in ps) == 2000
