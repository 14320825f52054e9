#!/usr/bin/env python
"""Gradient-sync + update benchmark (BASELINE.json metric: grad-sync+update steps/s).

One *step* = every parameter array of the model pushed from every GPU, reduced, updated once, and
the new weights present in every GPU's copy — the loop of the reference's own benchmark
tools/bandwidth/measure.py:112-121 (push all, pull all, wait), with the optimizer on the store.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload resnet50] [--impl reference]

Prints ONE JSON line (rank 0).  `value` times the step with gradients already resident in HBM;
`e2e` times the same step through the kvstore plugin call with HOST buffers (pinned host gradients
copied in, updated weights copied back, both inside the timed region).  `--impl reference` times
the reference's own CPU path (CommCPU reduce + CPU optimizer functors from oracle/_ref, or the
bit-identical C restatement when that is absent) on the host cores.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (shape list, dtype, optimizer, bytes/elem of the fused update, description)
    "mlp": ("mlp", "f32", "sgd", 20, "C1 Gluon MLP 784-512-512-10 gradient set: pushpull + multi_sgd_mom_update"),
    "resnet50": ("resnet50", "f32", "sgd", 20,
                 "C2 ResNet-50 v1 gradient set (161 fp32 arrays, 25,557,032 params): pushpull + multi_sgd_mom_update"),
    "bert_base_sgd": ("bert_base", "f16", "sgd", 20,
                      "C3 BERT-base fp16 grads + fp32 masters (199 arrays, 109,482,240 params): multi_mp_sgd_mom_update"),
    "bert_base_adam": ("bert_base", "f16", "adam", 28,
                       "C3 BERT-base fp16 grads + fp32 masters (199 arrays, 109,482,240 params): fused mp adam_update"),
    "bert_large": ("bert_large", "f32", "lamb", 28,
                   "C4 BERT-large (391 fp32 arrays, 335,141,888 params): multi_lamb_update + global-norm clip"),
    "rsp": ("rsp", "f32", "sgd", 20,
            "C5 row_sparse embedding gradients (1,048,576 x 512 table, 65,536 Zipf ids per GPU with duplicates): "
            "ReduceRowSparse + lazy sgd_mom_update + row_sparse_pull"),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--workload", default="resnet50", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--e2e-steps", type=int, default=20)
    ap.add_argument("--cpu-steps", type=int, default=0, help="0 = pick a count that takes ~10-20 s")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-views", action="store_true",
                    help="use caller-owned gradient/weight arrays (adds the pack / unpack / out copies) "
                         "instead of zero-copy views into the store")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def opt_kwargs(kind, mp):
    if kind == "sgd":
        return dict(learning_rate=0.1, momentum=0.9, wd=1e-4, rescale_grad=1.0 / 256, multi_precision=mp)
    if kind == "adam":
        return dict(learning_rate=1e-4, beta1=0.9, beta2=0.999, epsilon=1e-8, wd=0.01,
                    rescale_grad=1.0 / 256, multi_precision=mp)
    return dict(learning_rate=1e-3, beta1=0.9, beta2=0.999, epsilon=1e-6, wd=0.01, bias_correction=True,
                rescale_grad=1.0 / 256, multi_precision=mp)


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's CommCPU reduce + CPU optimizer functors on the host cores
# ------------------------------------------------------------------------------------------------
def cpu_run(args, ncopies, steps, warmup, budget_s=None):
    from mxnet_maintenance_b200 import shapes
    from oracle.oracle import Oracle, have_ref, adam_lr
    shp_name, dt, kind, _, _ = WORKLOADS[args.workload]
    shp = shapes.CONFIGS[shp_name]()
    backend = "ref" if have_ref() else "port"
    cores = os.cpu_count() or 1
    o = Oracle(backend, nthread=cores)
    rng = np.random.default_rng(1234)
    wdt = np.float16 if dt == "f16" else np.float32
    w = [rng.uniform(-1, 1, s).astype(np.float32).astype(wdt).reshape(-1) for s in shp]
    g = [[(rng.uniform(-1, 1, s) * 1e-2).astype(np.float32).astype(wdt).reshape(-1) for s in shp]
         for _ in range(ncopies)]
    n = len(shp)
    mp = dt == "f16"
    w32 = [x.astype(np.float32) for x in w] if mp else None
    s0 = [np.zeros(x.size, np.float32) for x in w]
    s1 = [np.zeros(x.size, np.float32) for x in w]
    kw = opt_kwargs(kind, mp)
    merged = [np.empty_like(x) for x in w]

    def step(t):
        # CommCPU::Reduce copies src[0] into the merge buffer, then ReduceSumCPU (comm.h:121-207)
        for k in range(n):
            np.copyto(merged[k], g[0][k])
            if ncopies > 1:
                o.reduce_sum([merged[k]] + [g[c][k] for c in range(1, ncopies)], nthread=min(4, cores))
        if kind == "sgd":
            o.multi_sgd(w, merged, s0, w32, [kw["learning_rate"]] * n, [kw["wd"]] * n, kw["momentum"],
                        kw["rescale_grad"], -1.0)
        elif kind == "adam":
            lr = adam_lr(kw["learning_rate"], kw["beta1"], kw["beta2"], t)
            for k in range(n):   # Adam has no aggregated path in the reference (aggregate_num = 0)
                if mp:
                    o.mp_adam(w[k], merged[k], w32[k], s0[k], s1[k], lr, kw["wd"], beta1=kw["beta1"],
                              beta2=kw["beta2"], eps=kw["epsilon"], rescale=kw["rescale_grad"])
                else:
                    o.adam(w[k], merged[k], s0[k], s1[k], lr, kw["wd"], kw["beta1"], kw["beta2"],
                           kw["epsilon"], kw["rescale_grad"])
        else:
            if backend == "port":
                o.clip_global_norm(merged, 1.0)
            else:
                Oracle("port", nthread=cores).clip_global_norm(merged, 1.0)
            o.multi_lamb(w, merged, s0, s1, w32, [kw["learning_rate"]] * n, [kw["wd"]] * n, [t] * n,
                         kw["beta1"], kw["beta2"], kw["epsilon"], kw["rescale_grad"])
        # CommCPU::Broadcast: copy the stored value to every device copy (comm.h:209-224)
        for c in range(ncopies):
            for k in range(n):
                np.copyto(g[c][k], w[k])

    for t in range(1, warmup + 1):
        step(t)
    t0 = time.perf_counter()
    done = 0
    for t in range(warmup + 1, warmup + steps + 1):
        step(t)
        done += 1
        if budget_s is not None and time.perf_counter() - t0 > budget_s and done >= 3:
            break
    dt_s = time.perf_counter() - t0
    return done / dt_s, done, cores, ("reference" if backend == "ref" else "port")


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    shp_name, dt, kind, bpe, desc = WORKLOADS[args.workload]
    steps = args.steps if args.steps <= 50 else 50
    sps, done, cores, kind_s = cpu_run(args, max(args.gpus, 1), steps, min(args.warmup, 3), budget_s=120)
    line = {
        "impl": "reference", "metric": "grad-sync+update steps/s", "value": sps, "unit": "steps/s",
        "n_gpus": args.gpus, "steps": done, "warmup": min(args.warmup, 3), "ms_per_step": 1e3 / sps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": dt,
        "data": "synthetic",
        "config": {"workload": desc, "device_copies": max(args.gpus, 1), "optimizer": kind,
                   "where": "host cores: CommCPU reduce + CPU optimizer functors + broadcast memcpy, "
                            "gradients already in host memory"},
        "cpu_baseline": {"value": sps, "unit": "steps/s", "cores": cores, "kind": kind_s,
                         "sample": "%d full steps of the same workload" % done},
        "e2e": {"value": sps, "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# clocks sampler (NVML): SM clock + throttle reasons DURING the timed region
# ------------------------------------------------------------------------------------------------
class Clocks:
    def __init__(self, index):
        self.samples, self.reasons, self.stop = [], set(), False
        self.max_mhz = None
        self.active = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.th = threading.Thread(target=self._loop, daemon=True)
            self.th.start()
        except Exception as e:  # pragma: no cover
            self.nv = None
            self.err = str(e)

    def _loop(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self.stop:
            if self.active:
                try:
                    self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                    try:
                        r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                    except Exception:
                        r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                    for bit, name in names.items():
                        if r & bit:
                            self.reasons.add(name)
                except Exception:
                    pass
            time.sleep(0.002)

    def result(self):
        self.stop = True
        if self.nv is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "error": getattr(self, "err", "")}
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


# ------------------------------------------------------------------------------------------------
# C5: row-sparse push (+ lazy update) and row_sparse_pull
# ------------------------------------------------------------------------------------------------
def rsp_inputs(rank, step, num_rows, row_len, nids):
    rng = np.random.default_rng(1234 + 1000 * step + rank)
    raw = np.minimum(rng.zipf(1.05, nids) - 1, num_rows - 1).astype(np.int64)
    idx = np.unique(raw)
    val = (rng.uniform(-1, 1, (len(idx), row_len)) * 1e-2).astype(np.float32)
    pull_ids = np.minimum(rng.zipf(1.05, nids) - 1, num_rows - 1).astype(np.int64)
    return idx, val, pull_ids


def run_rsp_cpu(args, world, steps):
    from mxnet_maintenance_b200 import shapes
    from oracle.oracle import Oracle
    num_rows, row_len = shapes.ROW_SPARSE_TABLE
    cores = os.cpu_count() or 1
    o = Oracle("port", nthread=cores)
    table = np.zeros((num_rows, row_len), np.float32)
    mom = np.zeros_like(table)
    ins = [rsp_inputs(r, 0, num_rows, row_len, shapes.ROW_SPARSE_IDS) for r in range(world)]
    t0 = time.perf_counter()
    for _ in range(steps):
        midx, mval = o.rsp_sum([(i, v) for i, v, _ in ins])
        o.sgd_mom_rsp(table, mom, midx, mval, 0.1, 1e-4, 0.9)
        for _, _, pid in ins:
            u = o.unique(pid)
            out = table[u]          # full table: SparseRetainCopyRetainedRowsFromDns (direct row copy)
    return steps / (time.perf_counter() - t0), cores


def run_rsp(args):
    import torch
    import torch.distributed as dist
    import mxnet_maintenance_b200 as mb
    from mxnet_maintenance_b200 import shapes
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    num_rows, row_len = shapes.ROW_SPARSE_TABLE
    nids = shapes.ROW_SPARSE_IDS
    kv = mb.create("b200")
    kv.set_optimizer(mb.optimizer.SGD(learning_rate=0.1, momentum=0.9, wd=1e-4))
    kv.init_row_sparse(0, torch.zeros((num_rows, row_len), device=dev))
    nvar = 4   # rotate through a few input sets so consecutive steps differ
    sets = []
    for s in range(nvar):
        idx, val, pid = rsp_inputs(rank, s, num_rows, row_len, nids)
        sets.append((mb.RowSparse(torch.from_numpy(val).to(dev), torch.from_numpy(idx).to(dev), (num_rows, row_len)),
                     torch.from_numpy(pid).to(dev), len(idx)))
    out = mb.RowSparse(None, None, (0, 0))

    def step(i):
        g, pid, _ = sets[i % nvar]
        kv.push(0, [g])
        kv.row_sparse_pull(0, out=out, row_ids=pid)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    clocks = Clocks(local) if rank == 0 else None
    for i in range(max(args.warmup, 3)):
        step(i)
    barrier()
    l0 = kv.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if clocks:
        clocks.active = True
    steps = min(args.steps, 100)
    e0.record()
    for i in range(steps):
        step(i)
    e1.record()
    barrier()
    if clocks:
        clocks.active = False
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    launches = kv.launch_count() - l0
    # e2e: host ids + rows in, pulled rows out
    h = [(s[0].data.cpu().pin_memory(), s[0].indices.cpu().pin_memory(), s[1].cpu().pin_memory()) for s in sets]
    def e2e_step(i):
        v, ix, pid = h[i % nvar]
        g = mb.RowSparse(v.to(dev, non_blocking=True), ix.to(dev, non_blocking=True), (num_rows, row_len))
        kv.push(0, [g])
        kv.row_sparse_pull(0, out=out, row_ids=pid.to(dev, non_blocking=True))
        return out.data.cpu()
    e2e_step(0)
    barrier()
    t0 = time.perf_counter()
    for i in range(args.e2e_steps):
        res = e2e_step(i)
    barrier()
    e2e_s = time.perf_counter() - t0
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    ms_step = ms / steps
    peak, peak_src = peaks()
    nnr = float(np.mean([s[2] for s in sets]))
    uniq_pull = float(np.mean([len(np.unique(s[1].cpu().numpy())) for s in sets]))
    # algorithmic bytes: reduce reads nnr*(8+4L) per copy, lazy update 20 B per
    # merged element, pull reads + writes U'*4L (+ ids)
    alg = world * nnr * (8 + 4 * row_len) + 20.0 * (world * nnr * 0.8) * row_len + 2 * uniq_pull * 4 * row_len + 8 * nids
    line = {
        "metric": "grad-sync+update steps/s", "value": 1e3 / ms_step, "unit": "steps/s", "n_gpus": world,
        "steps": steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOADS["rsp"][4], "optimizer": "sgd momentum (lazy row update)",
                   "rows_per_push_per_gpu": nnr, "unique_rows_pulled": uniq_pull,
                   "l2": "touched rows (%.0f MB) fit the 126 MB L2 only partly; inputs rotate over %d sets"
                         % (nnr * row_len * 4 * 3 / 1e6, nvar)},
        "roofline": {"bound": "hbm", "achieved": alg / (ms_step * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                     "frac": alg / (ms_step * 1e-3) / 1e9 / peak, "traffic": None,
                     "kernel": "radix sort + rsp_reduce_kernel<sgd> + rsp_gather_kernel (whole step)",
                     "algorithmic_bytes_per_step": alg, "peak_source": peak_src,
                     "note": "whole-step time incl. the row-count read-back that row_sparse_pull needs"},
        "e2e": {"value": args.e2e_steps / e2e_s, "unit": "steps/s",
                "h2d_bytes_per_step": int(nnr * (8 + 4 * row_len) + 8 * nids),
                "d2h_bytes_per_step": int(uniq_pull * 4 * row_len), "steps": args.e2e_steps},
        "gpu_launches": launches, "clocks": clocks.result() if clocks else None,
    }
    if not args.no_cpu_baseline and world == 1:
        cs, cores = run_rsp_cpu(args, 1, 5)
        line["cpu_baseline"] = {"value": cs, "unit": "steps/s", "cores": cores, "kind": "port",
                                "sample": "5 steps: ElementwiseSumRsp + SGDMomDnsRspDnsKernel + Unique + row copy"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    import mxnet_maintenance_b200 as mb
    from mxnet_maintenance_b200 import shapes

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == max(args.gpus, 1), "launch with torchrun --nproc-per-node %d" % args.gpus
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    shp_name, dt, kind, bpe, desc = WORKLOADS[args.workload]
    shp = shapes.CONFIGS[shp_name]()
    nelem = sum(shapes.numel(s) for s in shp)
    mp = dt == "f16"
    tdt = torch.float16 if mp else torch.float32
    esz = 2 if mp else 4
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234 + rank)
    wgen = torch.Generator(device=dev)
    wgen.manual_seed(99)
    # flat device buffers; the per-array tensors are views (128-byte aligned like MXNet's pooled
    # allocations), so the e2e leg can move a whole gradient set with one copy per direction
    offs, tot = [], 0
    for s in shp:
        offs.append(tot)
        tot += (shapes.numel(s) + 63) // 64 * 64
    flat_w = ((torch.rand(tot, device=dev, generator=wgen) * 2 - 1)).to(tdt)
    flat_g = ((torch.rand(tot, device=dev, generator=gen) * 2 - 1) * 1e-2).to(tdt)
    weights = [flat_w[o:o + shapes.numel(s)].view(s) for o, s in zip(offs, shp)]
    grads = [flat_g[o:o + shapes.numel(s)].view(s) for o, s in zip(offs, shp)]
    keys = list(range(len(shp)))

    kv = mb.create("b200")
    O = getattr(mb.optimizer, {"sgd": "SGD", "adam": "Adam", "lamb": "LAMB"}[kind])
    kv.set_optimizer(O(**opt_kwargs(kind, mp)))
    if kind == "lamb":
        kv.clip_global_norm = 1.0
    outs = [[w] for w in weights]
    kv.broadcast(keys, weights, outs)
    kv.wait()
    if not args.no_views:
        # the framework's parameter / gradient arrays ARE the store's buffers (b200kv_view): no
        # pack, unpack or second weight copy exists on the path
        src_g = flat_g
        weights = [kv.weight_view(k, s) for k, s in zip(keys, shp)]
        grads = [kv.grad_view(k, s) for k, s in zip(keys, shp)]
        for g, o, s in zip(grads, offs, shp):
            g.copy_(src_g[o:o + shapes.numel(s)].view(s))
        flat_w, flat_g = kv.segment_views(keys[0])
        tot = flat_w.numel()
        outs = [[w] for w in weights]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    clocks = Clocks(local) if rank == 0 else None

    # the step: one replayable pushpull of the whole gradient set (same arrays every step)
    step = kv.bind_pushpull(keys, grads, out=outs)

    # ---- value: gradients resident in HBM -------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    kv.set_profiling(True)
    kv.profile_read()
    l0 = kv.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if clocks:
        clocks.active = True
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    phases = kv.profile_read_phases()
    kernel_ms, kernel_n = phases["update"]
    kv.set_profiling(False)
    launches = kv.launch_count() - l0
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_step = ms / args.steps
    sps = 1e3 / ms_step

    # host enqueue cost per step, measured without ring back-pressure (6 steps after a sync)
    barrier()
    t_h0 = time.perf_counter()
    for _ in range(6):
        step()
    t_enq = (time.perf_counter() - t_h0) / 6
    barrier()

    # ---- e2e: host gradients in, host weights out, through the plugin call ------------------------
    # Pinned host buffers; H2D of step k+1 and D2H of step k ride separate streams (PCIe is full
    # duplex) while the update runs on the compute stream, ordered by events.
    h_g = torch.empty(tot, dtype=tdt).pin_memory()
    h_g.copy_(flat_g)
    h_w = torch.empty(tot, dtype=tdt).pin_memory()
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    main = torch.cuda.current_stream()
    ev_in = [torch.cuda.Event() for _ in range(2)]
    ev_upd = torch.cuda.Event()
    ev_out = torch.cuda.Event()

    def e2e_loop(n):
        ev_upd.record(main)
        ev_out.record(main)
        for _ in range(n):
            with torch.cuda.stream(s_in):
                s_in.wait_event(ev_upd)          # the previous update has consumed the gradients
                flat_g.copy_(h_g, non_blocking=True)
                ev_in[0].record(s_in)
            main.wait_event(ev_in[0])
            main.wait_event(ev_out)              # the previous weights have left the device
            step()
            ev_upd.record(main)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_upd)
                h_w.copy_(flat_w, non_blocking=True)
                ev_out.record(s_out)
        main.wait_event(ev_out)

    e2e_loop(3)
    barrier()
    t0 = time.perf_counter()
    e2e_loop(args.e2e_steps)
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    if clocks:
        clocks.active = False
    e2e_sps = args.e2e_steps / e2e_s

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline --------------------------------------------------------------------------------
    peak, peak_src = peaks()
    if world == 1:
        alg = float(bpe) * nelem
        k_ms = kernel_ms / max(kernel_n, 1)
        ach = alg / (k_ms * 1e-3) / 1e9
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic_r1.json")
        if os.path.exists(tpath) and not args.no_views:
            with open(tpath) as f:
                traffic = json.load(f).get(args.workload, {}).get("traffic")
        roof = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": traffic, "traffic_source": "profiles/traffic_r1.json (ncu --set full, one launch)",
                "kernel": "mta_kernel (fused multi-tensor %s update)" % kind,
                "algorithmic_bytes_per_step": alg, "bytes_per_element": bpe,
                "kernel_ms_per_step": k_ms, "kernel_share_of_step": k_ms / ms_step,
                "launches_per_step": launches / args.steps, "peak_source": peak_src,
                "note": "kernel time = CUDA events recorded by the library on its stream around the "
                        "update kernel(s) of every timed step (b200kv_profile_read)"}
    else:
        # Bytes that must cross NVLink per GPU per direction in one step: the reduce-scatter half
        # ((N-1)/N of the gradient bytes in) and the all-gather half ((N-1)/N of the weight bytes
        # in) share each direction, so they add (SURVEY.md §8(d) "un-overlapped" bound).  Peak =
        # the measured peer-copy bandwidth this pool's B200s reach, 770 GB/s per direction
        # (B200_PROFILING.md; 900 GB/s nominal is reported beside it).
        sbytes = float(nelem) * esz
        link_bytes = 2.0 * sbytes * (world - 1) / world
        k_ms = kernel_ms / max(kernel_n, 1)
        t_upd = float(bpe) * nelem / world / (peak * 1e9)
        ach_kernel = link_bytes / (k_ms * 1e-3) / 1e9
        ach_step = link_bytes / (ms_step * 1e-3) / 1e9
        # single fused kernel (SGD / Adam on the peer-memory path): time that kernel; otherwise
        # (NCCL path, or LAMB whose peer traffic is split over a reduce pass and the last update
        # pass) the link bytes are spread over the step, so the step time is the denominator
        p2p_on = os.environ.get("B200KV_P2P", "1") != "0" and kind != "lamb"
        nv_env = os.environ.get("B200KV_NVLS")
        nvls_on = (os.environ.get("B200KV_P2P", "1") != "0" and dt == "f32" and
                   (nv_env == "1" or (nv_env is None and world >= 4 and kind != "lamb")))
        roof = {"bound": "nvlink", "achieved": ach_kernel if p2p_on else ach_step, "peak": 770.0, "unit": "GB/s",
                "frac": (ach_kernel if p2p_on else ach_step) / 770.0, "traffic": None,
                "kernel": ("mta_kernel over peer memory: fused reduce-scatter + %s update + all-gather" % kind)
                          if p2p_on else "ncclReduceScatter + sharded update + ncclAllGather",
                "exchange": "nvls multicast (multimem.ld_reduce / multimem.st)" if nvls_on else
                            ("peer memory loads/stores" if os.environ.get("B200KV_P2P", "1") != "0" else "nccl"),
                "nvls_actual_link_bytes_per_direction": sbytes * (1.0 + 1.0 / world) if nvls_on else None,
                "link_bytes_per_direction_per_step": link_bytes, "kernel_ms_per_step": k_ms,
                "step_achieved_GBps": ach_step, "step_frac_of_measured_peak": ach_step / 770.0,
                "step_frac_of_nominal_900": ach_step / 900.0,
                "step_lower_bound_ms_nominal": (link_bytes / 900e9 + t_upd) * 1e3,
                "phase_ms_per_step": {k: v[0] / max(v[1], 1) for k, v in phases.items()},
                "peak_source": "measured peer copy 770 GB/s per direction (B200_PROFILING.md); nominal 900",
                "note": "achieved = algorithmic collective bytes per direction (SURVEY 8d: reduce-scatter + "
                        "all-gather, 2*S*(N-1)/N) / fused-kernel time (CUDA events recorded by the library "
                        "around the kernel on rank 0); with NVLS the switch sums and replicates, so fewer "
                        "bytes actually cross the links and the fraction can approach or exceed 1"}

    line = {
        "metric": "grad-sync+update steps/s", "value": sps, "unit": "steps/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": dt, "data": "synthetic",
        "config": {"workload": desc, "optimizer": kind, "update_on_kvstore": True,
                   "arrays": "caller-owned (pack/unpack copies)" if args.no_views else
                             "zero-copy views into the store (b200kv_view)",
                   "l2": "working set %.0f MB per step exceeds the 126 MB L2; no flush needed"
                         % ((esz * 2 + (16 if kind != "sgd" else 8) + (4 if mp else 0)) * nelem / 1e6),
                   "parallelism": "dp%d reduce-scatter / sharded update / all-gather" % world if world > 1
                                  else "single GPU, fused update in place"},
        "roofline": roof,
        "e2e": {"value": e2e_sps, "unit": "steps/s", "h2d_bytes_per_step": nelem * esz,
                "d2h_bytes_per_step": nelem * esz, "steps": args.e2e_steps,
                "what": "pinned host gradients -> device, kvstore pushpull (bound), updated weights -> pinned "
                        "host; H2D / update / D2H pipelined on three streams"},
        "gpu_launches": launches,
        "host_enqueue_ms_per_step": 1e3 * t_enq,
        "clocks": clocks.result() if clocks else None,
    }
    if not args.no_cpu_baseline and world == 1:
        cpu_steps = args.cpu_steps or 200
        cs, done, cores, kind_s = cpu_run(args, 1, cpu_steps, 2, budget_s=15)
        line["cpu_baseline"] = {"value": cs, "unit": "steps/s", "cores": cores, "kind": kind_s,
                                "sample": "%d full steps of the same workload on the host cores, gradients "
                                          "already in host memory" % done}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    elif a.workload == "rsp":
        run_rsp(a)
    else:
        run_b200(a)
