"""Multi-GPU parity check, one process per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29533 tests/multi_gpu_check.py

Every rank pushes its own gradient set; the expected result is the CPU oracle's CommCPU reduce over
the ranks' gradients followed by the reference optimizer functor.  With two ranks an fp32 / fp16
sum of two numbers is order-independent, so N=2 must be BIT-EXACT; for N>2 NCCL's reduction order
differs from CommCPU's and the north-star tolerances apply (fp32 rtol 1e-5 / atol 1e-6 per step,
fp16-gradient paths rtol 2e-3)."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import mxnet_maintenance_b200 as mb
    from oracle.oracle import Oracle, LocalStoreModel
    oracle = Oracle("port")
    shp = [(1,), (3,), (4095,), (4097,), (2, 4099), (33, 7, 3), (100003,), (64,), (300, 300)]
    keys = list(range(len(shp)))
    failures = []

    def run(kind, mp=False, gclip=None, batch=True, steps=3, ncopies=1):
        wdt = np.float16 if mp else np.float32
        rng_w = np.random.default_rng(77)
        w0 = [rng_w.uniform(-1, 1, s).astype(np.float32).astype(wdt) for s in shp]
        lr = 1e-3 if kind == "lamb" else 0.05
        common = dict(learning_rate=lr, wd=0.01, rescale_grad=1.0 / 8, multi_precision=mp)
        odict = dict(lr=lr, wd=0.01, rescale=1.0 / 8, multi_precision=mp, kind=kind)
        if kind == "sgd":
            opt = mb.optimizer.SGD(momentum=0.9, **common); odict["momentum"] = 0.9
        elif kind == "adam":
            opt = mb.optimizer.Adam(**common)
        elif kind == "lamb":
            opt = mb.optimizer.LAMB(**common)
        else:
            opt, odict = None, None
        kv = mb.create("b200")
        assert kv.num_workers == world and kv.rank == rank
        if opt is not None:
            kv.set_optimizer(opt)
        if gclip is not None:
            kv.clip_global_norm = gclip
        model = LocalStoreModel(oracle, odict)
        # rank 0's value wins the broadcast: give the other ranks garbage to prove it
        dev_w = [torch.from_numpy(w if rank == 0 else w * 0 + 7).cuda() for w in w0]
        for k, w in zip(keys, w0):
            model.init(k, w)
        kv.broadcast(keys, dev_w, [[w] for w in dev_w])
        for step in range(steps):
            # gradients of every rank (all ranks can compute all of them from the seeds)
            allg = []
            for r in range(world):
                rg = np.random.default_rng(1000 * step + r)
                allg.append([[(rg.uniform(-1, 1, s) * 1e-2).astype(np.float32).astype(wdt)
                              for _ in range(ncopies)] for s in shp])
            mine = [[torch.from_numpy(g).cuda() for g in gs] for gs in allg[rank]]
            if batch:
                kv.pushpull(keys, mine, out=[[w] for w in dev_w])
            else:
                for k in keys:
                    kv.pushpull(k, mine[k], out=dev_w[k])
            kv.wait()
            # oracle: device-local copies first (CommCPU order), then across ranks (CommCPU order)
            merged = []
            for k in keys:
                per_rank = [oracle.reduce_sum([g.copy() for g in allg[r][k]]) for r in range(world)]
                merged.append(oracle.reduce_sum([g.copy() for g in per_rank]))
            if gclip is not None:
                m32 = [m.astype(np.float32) for m in merged]
                oracle.clip_global_norm(m32, gclip)
                merged = [m.astype(wdt) for m in m32]
            interval = {}
            if kind == "adam" and not mp:
                # the same reference update from the same state, merged gradient shifted by -/+ delta
                import copy
                for k in keys:
                    delta = (world * np.finfo(np.float32).eps *
                             sum(np.abs(g).astype(np.float64) for r in range(world) for g in allg[r][k]))
                    ends = []
                    for sign in (-1.0, 1.0):
                        m2 = LocalStoreModel(oracle, odict)
                        m2.store = {k: model.store[k].copy()}
                        m2.state = {k: copy.deepcopy(model.state[k])} if k in model.state else {}
                        m2.count = {k: model.count.get(k, 0)}
                        gk = (merged[k].astype(np.float64) + sign * delta).astype(np.float32)
                        ends.append(m2.pushpull(k, [gk]).astype(np.float64))
                    interval[k] = ends
            ref = [model.pushpull(k, [merged[k]]) for k in keys]
            for k in keys:
                got = dev_w[k].cpu().numpy()
                # peer-memory path (whole-model calls): ranks are summed in CommCPU order => exact
                # for any world size; NCCL path (per-key calls): exact only for two ranks
                # NVLS path (default when the fabric has multicast): the switch's summation order is
                # unspecified => tolerance beyond two ranks
                nvls = os.environ.get("B200KV_NVLS", "1" if world >= 4 else "0") == "1"
                p2p = os.environ.get("B200KV_P2P", "1") != "0" and not nvls and batch and kind is not None
                exact = (world == 2 or p2p) and kind != "lamb" and gclip is None
                if exact:
                    ok = got.tobytes() == ref[k].tobytes()
                else:
                    rtol, atol = (2e-3, 1e-3) if mp else (1e-5, 1e-6)
                    close = np.isclose(got.astype(np.float32), ref[k].astype(np.float32), rtol=rtol, atol=atol)
                    ok = bool(close.all())
                    if kind == "adam" and not ok and not mp:
                        # Adam divides by sqrt(v)+eps: where rescale*sum(g) + wd*w is ~0 a one-ulp change
                        # of the summed gradient (NCCL / NVSwitch add in another order than CommCPU) is
                        # amplified by up to lr/eps.  Such an element is accepted only if it lies inside
                        # the interval the reference update spans when the merged gradient moves by the
                        # reduction-order uncertainty delta = W * eps32 * sum_r |g_r| (plus the tolerance).
                        lo, hi = interval[k]
                        g32 = got.astype(np.float64)
                        tol = atol + rtol * np.abs(ref[k].astype(np.float64))
                        ok = bool(((g32 >= np.minimum(lo, hi) - tol) & (g32 <= np.maximum(lo, hi) + tol)).all())
                        if ok:
                            # an element accepted through its interval keeps that offset in the weight
                            # for good: later steps are judged from the weights the GPU really holds
                            model.store[k] = got.copy()
                if not ok:
                    d = np.abs(got.astype(np.float64) - ref[k].astype(np.float64)).max()
                    failures.append("%s mp=%s gclip=%s batch=%s step %d key %d maxdiff %g" % (
                        kind, mp, gclip, batch, step, k, d))
        return kv, model

    run("sgd")
    run("sgd", batch=False)            # per-key calls: grouped ncclReduce / ncclBroadcast pieces
    run("sgd", ncopies=3)
    run("sgd", mp=True)
    run("adam")
    run("adam", mp=True)
    kv, model = run("lamb")
    run("lamb", gclip=0.05)
    run(None)                          # no optimizer: all-reduce, merged value assigned
    run(None, batch=False)

    # zero-copy store views: no pack / unpack, same results
    def run_views(kind):
        rng_w = np.random.default_rng(78)
        w0 = [rng_w.uniform(-1, 1, s).astype(np.float32) for s in shp]
        lr = 1e-3 if kind == "lamb" else 0.05
        O = {"sgd": mb.optimizer.SGD, "lamb": mb.optimizer.LAMB}[kind]
        kw = dict(momentum=0.9) if kind == "sgd" else {}
        kv = mb.create("b200")
        kv.set_optimizer(O(learning_rate=lr, wd=0.01, rescale_grad=1.0 / 8, **kw))
        od = dict(kind=kind, lr=lr, wd=0.01, rescale=1.0 / 8)
        if kind == "sgd":
            od["momentum"] = 0.9
        model = LocalStoreModel(oracle, od)
        kv.init(keys, [torch.from_numpy(w).cuda() for w in w0])
        for k, w in zip(keys, w0):
            model.init(k, w)
        wv = [kv.weight_view(k, s) for k, s in zip(keys, shp)]
        gv = [kv.grad_view(k, s) for k, s in zip(keys, shp)]
        step_fn = kv.bind_pushpull(keys, gv, out=[[w] for w in wv])
        for step in range(3):
            allg = [[(np.random.default_rng(5000 * step + r).uniform(-1, 1, s) * 1e-2).astype(np.float32)
                     for s in shp] for r in range(world)]
            for d, h in zip(gv, allg[rank]):
                d.copy_(torch.from_numpy(h))
            step_fn()
            kv.wait()
            for k in keys:
                merged = oracle.reduce_sum([allg[r][k].copy() for r in range(world)])
                ref = model.pushpull(k, [merged])
                got = wv[k].cpu().numpy()
                bits = kind == "sgd" and (world == 2 or
                                          os.environ.get("B200KV_NVLS", "1" if world >= 4 else "0") != "1")
                ok = got.tobytes() == ref.tobytes() if bits else np.allclose(got, ref, rtol=1e-5, atol=1e-6)
                if not ok:
                    failures.append("views %s step %d key %d" % (kind, step, k))
    run_views("sgd")
    run_views("lamb")

    # row-sparse keys: every rank pushes 2 device copies; union-sum in (rank, device) order
    def run_rsp(with_opt):
        num_rows, row_len = 3000, 64
        table = np.random.default_rng(3).uniform(-1, 1, (num_rows, row_len)).astype(np.float32)
        kv = mb.create("b200")
        if with_opt:
            kv.set_optimizer(mb.optimizer.SGD(learning_rate=0.1, momentum=0.9, wd=1e-3, rescale_grad=0.5))
        kv.init_row_sparse(0, torch.from_numpy(table if rank == 0 else table * 0).cuda())
        ref, mom = table.copy(), np.zeros_like(table)
        for step in range(2):
            allg = []
            for r in range(world):
                rg = np.random.default_rng(900 * step + r)
                for _ in range(2):
                    idx = np.unique(rg.integers(0, num_rows, 400)).astype(np.int64)
                    allg.append((r, idx, (rg.uniform(-1, 1, (len(idx), row_len)) * 1e-2).astype(np.float32)))
            mine = [mb.RowSparse(torch.from_numpy(v).cuda(), torch.from_numpy(i).cuda(), (num_rows, row_len))
                    for (r, i, v) in allg if r == rank]
            kv.push(0, mine)
            midx, mval = oracle.rsp_sum([(i, v) for (_, i, v) in allg])
            if with_opt:
                oracle.sgd_mom_rsp(ref, mom, midx, mval, 0.1, 1e-3, 0.9, 0.5)
            else:
                m = kv.rsp_merged(0)
                if m.indices.cpu().numpy().tobytes() != midx.tobytes() or m.data.cpu().numpy().tobytes() != mval.tobytes():
                    failures.append("rsp merged step %d" % step)
        if with_opt:
            out = mb.RowSparse(None, None, (0, 0))
            kv.row_sparse_pull(0, out=out, row_ids=torch.arange(num_rows, device="cuda"))
            if out.data.cpu().numpy().tobytes() != ref.tobytes():
                failures.append("rsp lazy update")
    run_rsp(False)
    run_rsp(True)

    # sharded optimizer state gathers back to the full per-key arrays (checkpointing), at EVERY world
    # size: bit-exact on the peer-memory path (ranks summed in CommCPU order), within the fp32
    # tolerance when the NVSwitch did the summing
    kv2, model2 = run("sgd", steps=2)
    st = kv2._collect_states()
    exact_states = kv2.last_exchange() in ("peer", "peer-push") or world == 2
    for k in keys:
        mom_ref = model2.state[k][0]
        have = st[k].reshape(-1)
        good = np.array_equal(have, mom_ref) if exact_states else np.allclose(have, mom_ref, rtol=1e-5, atol=1e-6)
        if not good:
            failures.append("state gather key %d (%s)" % (k, kv2.last_exchange()))
    kv3, model3 = run("adam", mp=True, steps=2)
    for k in keys:
        for which, want in model3.states_of(k).items():
            have = kv3.get_state(k, which)
            exact3 = kv3.last_exchange() in ("peer", "peer-push") or world == 2
            good = have is not None and (np.array_equal(have.reshape(-1), want) if exact3 else
                                         np.allclose(have.reshape(-1), want, rtol=2e-3, atol=1e-5))
            if not good:
                failures.append("mp-adam state %d of key %d (%s)" % (which, k, kv3.last_exchange()))

    # ---- the BASELINE configurations at full size: bench.py's own pre-flight (weights of every rank
    # and all gathered optimizer states vs the CommCPU + reference-functor oracle)
    if os.environ.get("MULTI_GPU_CHECK_FULL", "1") == "1":
        sys.path.insert(0, ROOT)
        import bench
        args = bench.parse_args_for_check(world)
        env = bench.Env(args, init=False)
        for wl in ("resnet50", "bert_base_sgd", "bert_base_adam", "bert_large"):
            line = bench.run_dense(args, env, wl, 3, 2, False, 2)
            if rank == 0:
                par = line["parity"]
                print("FULL", wl, "exchange", par["exchange"], "bit_exact", par["bit_exact"], "max_err_over_tol",
                      par["max_err_over_tol"], "pass", par["pass"], flush=True)
                must_be_exact = par["exchange"] in ("peer", "peer-push") and wl != "bert_large"
                if not par["pass"] or (must_be_exact and not par["bit_exact"]):
                    failures.append("full-size %s: %s" % (wl, json.dumps(par)[:600]))
        line = bench.run_rsp(args, env, 3, 2, False)
        if rank == 0:
            par = line["parity"]
            print("FULL rsp indices_bit_exact", par["indices_bit_exact"], "rows bit_exact", par["bit_exact"], flush=True)
            if not (par["indices_bit_exact"] and par["bit_exact"]):
                failures.append("full-size rsp: %s" % json.dumps(par)[:600])

    t = torch.tensor([len(failures)], device="cuda")
    dist.all_reduce(t)
    if rank == 0:
        for f in failures[:20]:
            print("FAIL", f)
        print("MULTI_GPU_CHECK", "FAILED" if t.item() else "OK", "world", world)
    dist.destroy_process_group()
    sys.exit(1 if t.item() else 0)


if __name__ == "__main__":
    main()
