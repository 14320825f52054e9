#!/usr/bin/env python
"""Gradient-sync + update benchmark (BASELINE.json metric: grad-sync+update steps/s).

One *step* = every parameter array of the model pushed from every GPU, reduced, updated once, and
the new weights present in every GPU's copy — the loop of the reference's own benchmark
tools/bandwidth/measure.py:112-121 (push all, pull all, wait), with the optimizer on the store.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload resnet50] [--impl reference]

Prints ONE JSON line (rank 0).  The headline (`value`, `roofline`, `e2e`, `parity`) is the C2
ResNet-50 configuration BASELINE.json's metric is quoted on; `configs` carries the same
measurements for the other BASELINE configurations (C3 mp-SGD / mp-Adam, C4 LAMB + global-norm
clip, C5 row-sparse).  `value` times the step with gradients already resident in HBM; `e2e` times
the same step through the kvstore plugin call with HOST buffers (pinned host gradients copied in,
updated weights copied back, both inside the timed region).  `parity` is a pre-flight OUTSIDE the
timed region: a few steps of this very configuration, every rank's weights and the gathered
optimizer states compared with the CPU oracle (CommCPU reduce over the ranks' gradients in rank
order, then the reference optimizer functor).  `--impl reference` times the reference's own CPU
path (CommCPU reduce + CPU optimizer functors from oracle/_ref, or the bit-identical C restatement
when that is absent) on the host cores.
"""
import argparse
import hashlib
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
EXTRA_CONFIGS = [("C3_mp_sgd", "bert_base_sgd"), ("C3_mp_adam", "bert_base_adam"), ("C4_lamb_clip", "bert_large"),
                 ("C5_rsp", "rsp")]
# tolerances of BASELINE.json's north star: fp32 weights / states per step; fp16-gradient paths
TOL32 = (1e-5, 1e-6)
TOL16 = (2e-3, 1e-6)


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
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle pre-flight")
    ap.add_argument("--parity-steps", type=int, default=3)
    ap.add_argument("--no-configs", action="store_true",
                    help="only the headline workload, no `configs` block for the other BASELINE configurations")
    ap.add_argument("--configs-budget", type=float, default=240.0,
                    help="seconds the extra configurations may take before the line is printed without them")
    ap.add_argument("--single-process", action="store_true",
                    help="ONE process driving --gpus N devices (b200kv_create_group), launched without torchrun")
    ap.add_argument("--no-views", action="store_true",
                    help="use caller-owned gradient/weight arrays (adds the pack / unpack / out copies) "
                         "instead of zero-copy views into the store")
    return ap.parse_args()


def parse_args_for_check(world):
    """Argument set tests/multi_gpu_check.py uses to drive run_dense / run_rsp for their pre-flight."""
    return argparse.Namespace(gpus=world, steps=3, warmup=3, workload="resnet50", impl="b200", e2e_steps=2,
                              cpu_steps=0, no_cpu_baseline=True, no_parity=False, parity_steps=2,
                              no_configs=True, configs_budget=240.0, no_views=False, single_process=False)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def kernel_source_hash():
    """Identity of the update kernel's source: a traffic figure captured with ncu is only quoted
    while the kernel it was captured from is the kernel being timed."""
    h = hashlib.sha256()
    for f in ("optim.cuh", "mta.h", "launch_impl.cuh"):
        with open(os.path.join(ROOT, "mxnet_maintenance_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def measured_traffic(workload):
    """dram bytes per launch from the newest profiles/traffic_*.json whose kernel hash matches."""
    pdir = os.path.join(ROOT, "profiles")
    best = None
    for name in sorted(os.listdir(pdir)) if os.path.isdir(pdir) else []:
        if name.startswith("traffic_") and name.endswith(".json"):
            with open(os.path.join(pdir, name)) as f:
                d = json.load(f)
            if d.get("kernel_source_sha16") == kernel_source_hash() and workload in d:
                best = (d[workload].get("traffic"), "profiles/%s (ncu --set full, one launch, kernel source %s)"
                        % (name, d["kernel_source_sha16"]))
    return best or (None, "no ncu capture of the current kernel source (%s) under profiles/" % kernel_source_hash())


def opt_kwargs(kind, mp):
    if kind == "sgd":
        return dict(learning_rate=0.1, momentum=0.9, wd=1e-4, rescale_grad=1.0 / 256, multi_precision=mp)
    if kind == "adam":
        return dict(learning_rate=1e-4, beta1=0.9, beta2=0.999, epsilon=1e-8, wd=0.01,
                    rescale_grad=1.0 / 256, multi_precision=mp)
    return dict(learning_rate=1e-3, beta1=0.9, beta2=0.999, epsilon=1e-6, wd=0.01, bias_correction=True,
                rescale_grad=1.0 / 256, multi_precision=mp)


def config_of(workload, ncopies):
    """The `config` object — identical, byte for byte, on the b200 arm and the reference arm."""
    shp_name, dt, kind, bpe, desc = WORKLOADS[workload]
    cfg = {"workload": desc, "optimizer": kind if workload != "rsp" else "sgd momentum (lazy row update)",
           "update_on_kvstore": True, "device_copies": ncopies}
    if workload == "rsp":
        cfg["l2"] = "inputs rotate over 4 sets; touched rows only partly fit the 126 MB L2"
    else:
        from mxnet_maintenance_b200 import shapes
        nelem = sum(shapes.numel(s) for s in shapes.CONFIGS[shp_name]())
        esz = 2 if dt == "f16" else 4
        ws = (esz * 2 + (16 if kind != "sgd" else 8) + (4 if dt == "f16" else 0)) * nelem / 1e6
        cfg["l2"] = ("working set %.0f MB per step exceeds the 126 MB L2; no flush needed" % ws if ws > 126 else
                     "working set %.0f MB per step fits the 126 MB L2" % ws)
    return cfg


def cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's CommCPU reduce + CPU optimizer functors on the host cores
# ------------------------------------------------------------------------------------------------
def cpu_run(workload, ncopies, steps, warmup, nthread, budget_s=None):
    """steps/s of the reference's CPU path: CommCPU::Reduce (comm.h:121-207: one device -> the source
    itself, no copy :127-129; else copy src[0] into the merge buffer, then ReduceSumCPU with
    MXNET_KVSTORE_REDUCTION_NTHREADS = 4 threads for arrays >= 1e6 elements), the optimizer functor
    with `nthread` OpenMP threads, CommCPU::Broadcast (:209-224)."""
    from mxnet_maintenance_b200 import shapes
    from oracle.oracle import Oracle, have_ref, adam_lr
    shp_name, dt, kind, _, _ = WORKLOADS[workload]
    shp = shapes.CONFIGS[shp_name]()
    backend = "ref" if have_ref() else "port"
    o = Oracle(backend, nthread=nthread)
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
    mbuf = [np.empty_like(x) for x in w] if ncopies > 1 else None
    port = o if backend == "port" else Oracle("port", nthread=nthread)

    def step(t):
        if ncopies == 1:
            merged = g[0]                     # comm.h:127-129: returned untouched
        else:
            merged = mbuf
            for k in range(n):
                np.copyto(merged[k], g[0][k])
                o.reduce_sum([merged[k]] + [g[c][k] for c in range(1, ncopies)], nthread=min(4, nthread))
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
            if ncopies == 1:                  # clip_global_norm scales in place: keep the inputs intact
                merged = [x.copy() for x in merged]
            port.clip_global_norm(merged, 1.0)
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
    return done / dt_s, done, ("reference" if backend == "ref" else "port")


def rsp_inputs(rank, step, num_rows, row_len, nids):
    rng = np.random.default_rng(1234 + 1000 * step + rank)
    raw = np.minimum(rng.zipf(1.05, nids) - 1, num_rows - 1).astype(np.int64)
    idx = np.unique(raw)
    val = (rng.uniform(-1, 1, (len(idx), row_len)) * 1e-2).astype(np.float32)
    pull_ids = np.minimum(rng.zipf(1.05, nids) - 1, num_rows - 1).astype(np.int64)
    return idx, val, pull_ids


def cpu_run_rsp(world, steps, warmup, nthread, budget_s=None):
    """ElementwiseSumRsp (ndarray_function.cc:59-176) + SGDMomDnsRspDnsKernel + Unique + row copy."""
    from mxnet_maintenance_b200 import shapes
    from oracle.oracle import Oracle
    num_rows, row_len = shapes.ROW_SPARSE_TABLE
    o = Oracle("port", nthread=nthread)
    table = np.zeros((num_rows, row_len), np.float32)
    mom = np.zeros_like(table)
    ins = [rsp_inputs(r, 0, num_rows, row_len, shapes.ROW_SPARSE_IDS) for r in range(world)]

    def step():
        midx, mval = o.rsp_sum([(i, v) for i, v, _ in ins])
        o.sgd_mom_rsp(table, mom, midx, mval, 0.1, 1e-4, 0.9)
        for _, _, pid in ins:
            u = o.unique(pid)
            table[u]                # full table: SparseRetainCopyRetainedRowsFromDns (direct row copy)
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    done = 0
    for _ in range(steps):
        step()
        done += 1
        if budget_s is not None and time.perf_counter() - t0 > budget_s and done >= 3:
            break
    return done / (time.perf_counter() - t0), done, "port"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    shp_name, dt, kind, bpe, desc = WORKLOADS[args.workload]
    ncopies = max(args.gpus, 1)
    warmup = max(args.warmup, 3)
    cores = os.cpu_count() or 1
    policy_threads = max(1, cores // 2)      # src/engine/openmp.cc:47-53: omp_get_num_procs() / 2 on x86
    if args.workload == "rsp":
        sps, done, kind_s = cpu_run_rsp(ncopies, args.steps, warmup, cores, budget_s=150)
        sps_policy, _, _ = cpu_run_rsp(ncopies, min(args.steps, 10), 1, policy_threads, budget_s=30)
    else:
        sps, done, kind_s = cpu_run(args.workload, ncopies, args.steps, warmup, cores, budget_s=150)
        sps_policy, _, _ = cpu_run(args.workload, ncopies, min(args.steps, 20), 2, policy_threads, budget_s=30)
    line = {
        "impl": "reference", "metric": "grad-sync+update steps/s", "value": sps, "unit": "steps/s",
        "n_gpus": args.gpus, "steps": done, "warmup": warmup, "ms_per_step": 1e3 / sps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": dt,
        "data": "synthetic", "config": config_of(args.workload, ncopies),
        "cpu_baseline": {"value": sps, "unit": "steps/s", "cores": cores, "kind": kind_s,
                         "sample": "%d full steps of the same workload" % done,
                         "omp_threads": cores, "cpu_model": cpu_model(),
                         "reference_thread_policy": {"omp_threads": policy_threads, "value": sps_policy,
                                                     "rule": "omp_get_num_procs()/2 (src/engine/openmp.cc:47-53); "
                                                             "CommCPU reduce: 4 threads for arrays >= 1e6 elements"}},
        "e2e": {"value": sps, "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "setup": "host cores: CommCPU reduce + CPU optimizer functors + broadcast memcpy, gradients already in "
                 "host memory (excludes the D2H / H2D copies the real 'local' store would add)",
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

    def snapshot(self, reset=True):
        if self.nv is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "error": getattr(self, "err", "")}
        s = sorted(self.samples)
        out = {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
               "reasons": sorted(self.reasons), "samples": len(s)}
        if reset:
            self.samples, self.reasons = [], set()
        return out


class Env:
    """Process-wide context: ranks, device, torch.distributed."""

    def __init__(self, args, init=True):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        assert self.world == max(args.gpus, 1), "launch with torchrun --nproc-per-node %d" % args.gpus
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1 and init:
            dist.init_process_group("nccl", device_id=self.dev)
        self.clocks = Clocks(self.local) if self.rank == 0 else None

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x):
        if self.world == 1:
            return x
        t = self.torch.tensor([x], device=self.dev, dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def all_true(self, flag):
        if self.world == 1:
            return bool(flag)
        t = self.torch.tensor([1 if flag else 0], device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MIN)
        return bool(t.item())


class ErrStat:
    """Running comparison of CUDA results with the oracle's."""

    def __init__(self, rtol, atol):
        self.rtol, self.atol = rtol, atol
        self.bit_exact, self.max_abs, self.max_rel, self.max_over_tol, self.n = True, 0.0, 0.0, 0.0, 0

    def add(self, got, ref):
        self.n += got.size
        if got.tobytes() == ref.tobytes():
            return
        self.bit_exact = False
        a, b = got.astype(np.float64).reshape(-1), ref.astype(np.float64).reshape(-1)
        d = np.abs(a - b)
        if not np.isfinite(d).all():
            self.max_over_tol = float("inf")
            return
        self.max_abs = max(self.max_abs, float(d.max()))
        self.max_over_tol = max(self.max_over_tol, float((d / (self.atol + self.rtol * np.abs(b))).max()))
        big = np.abs(b) > 1e-3
        if big.any():
            self.max_rel = max(self.max_rel, float((d[big] / np.abs(b[big])).max()))

    def result(self):
        return {"bit_exact": self.bit_exact, "max_abs_err": self.max_abs, "max_rel_err": self.max_rel,
                "max_err_over_tol": self.max_over_tol, "tol": [self.rtol, self.atol],
                "pass": self.bit_exact or self.max_over_tol <= 1.0, "elements_compared": self.n}


# ------------------------------------------------------------------------------------------------
# C5: row-sparse push (+ lazy update) and row_sparse_pull
# ------------------------------------------------------------------------------------------------
def run_rsp(args, env, steps, e2e_steps, cpu_baseline):
    import mxnet_maintenance_b200 as mb
    from mxnet_maintenance_b200 import shapes
    torch, dist, dev, world, rank = env.torch, env.dist, env.dev, env.world, env.rank
    num_rows, row_len = shapes.ROW_SPARSE_TABLE
    nids = shapes.ROW_SPARSE_IDS
    kv = mb.create("b200")
    kv.set_optimizer(mb.optimizer.SGD(learning_rate=0.1, momentum=0.9, wd=1e-4))
    kv.init_row_sparse(0, torch.zeros((num_rows, row_len), device=dev))
    nvar = 4   # rotate through a few input sets so consecutive steps differ
    host_sets = [[rsp_inputs(r, s, num_rows, row_len, nids) for r in range(world)] for s in range(nvar)] \
        if rank == 0 else None
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

    # ---- parity pre-flight: indices bit-exact, pulled rows bit-exact vs ElementwiseSumRsp + lazy SGD
    parity = None
    if not args.no_parity:
        psteps = min(args.parity_steps, nvar)
        idx_ok, st = True, ErrStat(0.0, 0.0)
        if rank == 0:
            from oracle.oracle import Oracle
            o = Oracle("port", nthread=os.cpu_count() or 1)
            table = np.zeros((num_rows, row_len), np.float32)
            mom = np.zeros_like(table)
        for s in range(psteps):
            step(s)
            kv.wait()
            if rank == 0:
                midx, mval = o.rsp_sum([(i, v) for i, v, _ in host_sets[s]])
                o.sgd_mom_rsp(table, mom, midx, mval, 0.1, 1e-4, 0.9)
                u = o.unique(host_sets[s][0][2])
                got_idx = out.indices.cpu().numpy()
                idx_ok = idx_ok and got_idx.tobytes() == u.tobytes()
                if got_idx.shape == u.shape:
                    st.add(out.data.cpu().numpy(), table[u])
                else:
                    st.bit_exact = False
                    st.max_over_tol = float("inf")
        if rank == 0:
            parity = dict(st.result(), checked=True, steps=psteps, indices_bit_exact=idx_ok,
                          exchange="local" if world == 1 else "owner-sharded table (row id % ranks): owners read their rows from the senders over NVLink",
                          oracle="ElementwiseSumRsp restatement (ndarray_function.cc:59-176) + reference "
                                 "SGDMomDnsRspDnsKernel + UniqueImpl<cpu>; rank 0's pulled rows compared")
            parity["pass"] = parity["pass"] and idx_ok

    for i in range(max(args.warmup, 3)):
        step(i)
    env.barrier()
    l0 = kv.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if env.clocks:
        env.clocks.snapshot()
        env.clocks.active = True
    e0.record()
    for i in range(steps):
        step(i)
    e1.record()
    env.barrier()
    if env.clocks:
        env.clocks.active = False
    ms = env.max_over_ranks(e0.elapsed_time(e1))
    launches = kv.launch_count() - l0
    # the same step with the engine-style pull (row count left on the device, no host wait inside)
    out_a = mb.RowSparse(None, None, (0, 0))
    for i in range(3):
        kv.push(0, [sets[i % nvar][0]])
        kv.row_sparse_pull(0, out=out_a, row_ids=sets[i % nvar][1], sync=False)
    env.barrier()
    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a0.record()
    for i in range(steps):
        kv.push(0, [sets[i % nvar][0]])
        kv.row_sparse_pull(0, out=out_a, row_ids=sets[i % nvar][1], sync=False)
    a1.record()
    env.barrier()
    ms_async = env.max_over_ranks(a0.elapsed_time(a1)) / steps
    # e2e: host ids + rows in, pulled rows out
    h = [(s[0].data.cpu().pin_memory(), s[0].indices.cpu().pin_memory(), s[1].cpu().pin_memory()) for s in sets]

    def e2e_step(i):
        v, ix, pid = h[i % nvar]
        g = mb.RowSparse(v.to(dev, non_blocking=True), ix.to(dev, non_blocking=True), (num_rows, row_len))
        kv.push(0, [g])
        kv.row_sparse_pull(0, out=out, row_ids=pid.to(dev, non_blocking=True))
        return out.data.cpu()
    e2e_step(0)
    env.barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        e2e_step(i)
    env.barrier()
    e2e_s = env.max_over_ranks(time.perf_counter() - t0)
    clocks = env.clocks.snapshot() if env.clocks else None
    del kv
    if rank != 0:
        return None
    ms_step = ms / steps
    peak, peak_src = peaks()
    nnr = float(np.mean([sum(len(i) for i, _, _ in hs) for hs in host_sets]))          # rows read, all ranks
    merged = float(np.mean([len(np.unique(np.concatenate([i for i, _, _ in hs]))) for hs in host_sets]))
    uniq_pull = float(np.mean([len(np.unique(hs[0][2])) for hs in host_sets]))
    # algorithmic bytes: reduce reads nnr*(8+4L) per copy, lazy update 20 B per
    # merged element, pull reads + writes U'*4L (+ ids)
    alg = nnr * (8 + 4 * row_len) + 20.0 * merged * row_len + 2 * uniq_pull * 4 * row_len + 8 * nids
    line = {
        "value": 1e3 / ms_step, "unit": "steps/s", "ms_per_step": ms_step, "steps": steps, "dtype": "f32",
        "config": dict(config_of("rsp", world), rows_pushed_per_step=nnr, merged_rows_per_step=merged,
                       unique_rows_pulled=uniq_pull),
        "roofline": {"bound": "hbm", "achieved": alg / (ms_step * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                     "frac": alg / (ms_step * 1e-3) / 1e9 / peak, "traffic": None,
                     "kernel": "row-sparse push (sort/merge + segmented reduce + lazy SGD) + row_sparse_pull "
                               "(sort + unique + gather): whole step",
                     "algorithmic_bytes_per_step": alg, "peak_source": peak_src,
                     "launches_per_step": launches / steps,
                     "note": "whole-step time incl. the row-count read-back that row_sparse_pull needs",
                     "ms_per_step_async_pull": ms_async,
                     "frac_async_pull": alg / (ms_async * 1e-3) / 1e9 / peak},
        "e2e": {"value": e2e_steps / e2e_s, "unit": "steps/s",
                "h2d_bytes_per_step": int(nnr / world * (8 + 4 * row_len) + 8 * nids),
                "d2h_bytes_per_step": int(uniq_pull * 4 * row_len), "steps": e2e_steps},
        "parity": parity, "gpu_launches": launches, "clocks": clocks,
    }
    if cpu_baseline:
        cores = os.cpu_count() or 1
        cs, done, kind_s = cpu_run_rsp(1, 20, 1, cores, budget_s=10)
        line["cpu_baseline"] = {"value": cs, "unit": "steps/s", "cores": cores, "kind": kind_s,
                                "sample": "%d steps: ElementwiseSumRsp + SGDMomDnsRspDnsKernel + Unique + row copy" % done}
    return line


# ------------------------------------------------------------------------------------------------
# dense configurations (C2, C3, C4)
# ------------------------------------------------------------------------------------------------
def parity_dense(env, kv, step, c, nsteps):
    """Pre-flight, outside every timed region: `nsteps` steps of THIS configuration with fresh random
    gradients on every rank; rank 0 rebuilds the expected result on the host — CommCPU reduce over
    the ranks' gradients in rank order (comm.h:357-392), then the reference optimizer functor — and
    compares the weights after every step and all optimizer states after the last one.  Every other
    rank checks that its weights are bit-identical to rank 0's."""
    torch, dist, dev, world, rank = env.torch, env.dist, env.dev, env.world, env.rank
    from oracle.oracle import Oracle, have_ref, adam_lr, clip_global_norm_f64, lamb_step_f64norms
    kind, mp, kw = c["kind"], c["mp"], c["kw"]
    flat_w, flat_g, mask, offs, shp = c["flat_w"], c["flat_g"], c["mask"], c["offs"], c["shp"]
    from mxnet_maintenance_b200 import shapes
    sizes = [shapes.numel(s) for s in shp]
    n = len(shp)
    st_w = ErrStat(*(TOL16 if mp else TOL32))
    st_s = ErrStat(*TOL32)
    same_everywhere = True
    if rank == 0:
        o = Oracle("ref" if have_ref() else "port", nthread=os.cpu_count() or 1)
        w = flat_w.cpu().numpy().copy()
        tot = w.size
        w32 = w.astype(np.float32) if mp else None
        s0, s1 = np.zeros(tot, np.float32), np.zeros(tot, np.float32)
    gen = torch.Generator(device=dev)
    gen.manual_seed(4321 + rank)
    for s in range(nsteps):
        fresh = ((torch.rand(flat_g.numel(), device=dev, generator=gen) * 2 - 1) * 1e-2).to(flat_g.dtype)
        flat_g.copy_(fresh * mask)
        del fresh
        if world > 1:
            glist = [torch.empty_like(flat_g) for _ in range(world)] if rank == 0 else None
            dist.gather(flat_g, glist, dst=0)
        else:
            glist = [flat_g]
        if rank == 0:
            hg = [g.cpu().numpy() for g in glist]
            del glist
            merged = o.reduce_sum([hg[0].copy()] + hg[1:]) if world > 1 else hg[0].copy()
            t = s + 1
            if kind == "sgd":
                o.multi_sgd([w], [merged], [s0], [w32] if mp else None, [kw["learning_rate"]], [kw["wd"]],
                            kw["momentum"], kw["rescale_grad"], -1.0)
            elif kind == "adam":
                lr = adam_lr(kw["learning_rate"], kw["beta1"], kw["beta2"], t)
                if mp:
                    o.mp_adam(w, merged, w32, s0, s1, lr, kw["wd"], beta1=kw["beta1"], beta2=kw["beta2"],
                              eps=kw["epsilon"], rescale=kw["rescale_grad"])
                else:
                    o.adam(w, merged, s0, s1, lr, kw["wd"], kw["beta1"], kw["beta2"], kw["epsilon"],
                           kw["rescale_grad"])
            else:
                per = [merged[a:a + m] for a, m in zip(offs, sizes)]
                clip_global_norm_f64(per, 1.0)
                for a, m, g_k in zip(offs, sizes, per):
                    lamb_step_f64norms(w[a:a + m], g_k, s0[a:a + m], s1[a:a + m], kw["learning_rate"], kw["wd"], t,
                                       kw["beta1"], kw["beta2"], kw["epsilon"], kw["rescale_grad"], True)
        step()
        kv.wait()
        if world > 1:
            ref0 = flat_w.clone()
            dist.broadcast(ref0, 0)
            same_everywhere = same_everywhere and bool(torch.equal(ref0, flat_w))
            del ref0
        if rank == 0:
            got = flat_w.cpu().numpy()
            for a, m in zip(offs, sizes):
                st_w.add(got[a:a + m], w[a:a + m])
    # optimizer states after the last step (collective: the state is sharded across the ranks)
    for k, (a, m) in enumerate(zip(offs, sizes)):
        for which in (0, 1, 2):
            have = kv.get_state(k, which)
            if rank != 0:
                continue
            want = {0: s0, 1: s1 if kind != "sgd" else None, 2: w32}[which]
            if have is None or want is None:
                if (have is None) != (want is None):
                    st_s.bit_exact, st_s.max_over_tol = False, float("inf")
                continue
            st_s.add(have.reshape(-1), want[a:a + m])
    same_everywhere = env.all_true(same_everywhere)
    if rank != 0:
        return None
    rw, rs = st_w.result(), st_s.result()
    return {"checked": True, "steps": nsteps, "exchange": kv.last_exchange(),
            "bit_exact": rw["bit_exact"] and rs["bit_exact"], "max_rel_err": max(rw["max_rel_err"], rs["max_rel_err"]),
            "max_abs_err": max(rw["max_abs_err"], rs["max_abs_err"]),
            "max_err_over_tol": max(rw["max_err_over_tol"], rs["max_err_over_tol"]),
            "tol": list(rw["tol"]), "state_tol": list(rs["tol"]),
            "pass": rw["pass"] and rs["pass"] and same_everywhere, "weights": rw, "states": rs,
            "all_ranks_bit_identical_weights": same_everywhere,
            "oracle": ("CommCPU reduce in rank order (comm.h:357-392) + " +
                       ("the reference's MultiLAMB step-1/step-2 arithmetic with float64-accumulated norms and "
                        "clip_global_norm (the CPU functor's sequential fp32 sums are off by ~1e-4 on 31M-element "
                        "tensors, SURVEY Appendix A.4)" if kind == "lamb" else
                        "reference functor %s (%s)" % ({"sgd": "MultiSGDKernel", "adam": "AdamUpdateKernel"}[kind],
                                                       "oracle/_ref" if have_ref() else "oracle port")))}


def run_dense(args, env, workload, steps, e2e_steps, cpu_baseline, psteps):
    import mxnet_maintenance_b200 as mb
    from mxnet_maintenance_b200 import shapes
    torch, dist, dev, world, rank = env.torch, env.dist, env.dev, env.world, env.rank
    shp_name, dt, kind, bpe, desc = WORKLOADS[workload]
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
    kw = opt_kwargs(kind, mp)
    kv.set_optimizer(O(**kw))
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
        for k, o in zip(keys, offs):
            assert (weights[k].data_ptr() - flat_w.data_ptr()) // esz == o
        tot = flat_w.numel()
        outs = [[w] for w in weights]
        del src_g
    mask = torch.zeros(tot, device=dev, dtype=tdt)
    for o, s in zip(offs, shp):
        mask[o:o + shapes.numel(s)] = 1

    # the step: one replayable pushpull of the whole gradient set (same arrays every step)
    step = kv.bind_pushpull(keys, grads, out=outs)

    parity = None
    if psteps > 0:
        parity = parity_dense(env, kv, step, dict(kind=kind, mp=mp, kw=kw, flat_w=flat_w, flat_g=flat_g,
                                                  mask=mask, offs=offs, shp=shp), psteps)
        env.barrier()
    del mask

    # ---- value: gradients resident in HBM -------------------------------------------------------
    warmup = max(args.warmup, 3)
    for _ in range(warmup):
        step()
    env.barrier()
    kv.set_profiling(True)
    kv.profile_read()
    l0 = kv.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if env.clocks:
        env.clocks.snapshot()
        env.clocks.active = True
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    env.barrier()
    ms = e0.elapsed_time(e1)
    phases = kv.profile_read_phases()
    kernel_ms, kernel_n = phases["update"]
    kv.set_profiling(False)
    launches = kv.launch_count() - l0
    ms = env.max_over_ranks(ms)
    ms_step = ms / steps
    sps = 1e3 / ms_step
    exchange = kv.last_exchange()

    # host enqueue cost per step, measured without ring back-pressure (6 steps after a sync)
    env.barrier()
    t_h0 = time.perf_counter()
    for _ in range(6):
        step()
    t_enq = (time.perf_counter() - t_h0) / 6
    env.barrier()

    # ---- e2e: host gradients in, host weights out, through the plugin call ------------------------
    # Pinned host buffers; H2D of step k+1 and D2H of step k ride separate streams (PCIe is full
    # duplex) while the update runs on the compute stream, ordered by events.
    h_g = torch.empty(tot, dtype=tdt).pin_memory()
    h_g.copy_(flat_g)
    h_w = torch.empty(tot, dtype=tdt).pin_memory()
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    main = torch.cuda.current_stream()
    ev_in = torch.cuda.Event()
    ev_upd = torch.cuda.Event()
    ev_out = torch.cuda.Event()

    def e2e_loop(n, with_step=True):
        ev_upd.record(main)
        ev_out.record(main)
        for _ in range(n):
            with torch.cuda.stream(s_in):
                s_in.wait_event(ev_upd)          # the previous update has consumed the gradients
                flat_g.copy_(h_g, non_blocking=True)
                ev_in.record(s_in)
            main.wait_event(ev_in)
            main.wait_event(ev_out)              # the previous weights have left the device
            if with_step:
                step()
            ev_upd.record(main)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_upd)
                h_w.copy_(flat_w, non_blocking=True)
                ev_out.record(s_out)
        main.wait_event(ev_out)

    e2e_loop(3)
    env.barrier()
    t0 = time.perf_counter()
    e2e_loop(e2e_steps)
    env.barrier()
    e2e_s = env.max_over_ranks(time.perf_counter() - t0)
    # the host link alone: the same two copies per step with no update between them
    e2e_loop(2, with_step=False)
    env.barrier()
    t0 = time.perf_counter()
    e2e_loop(max(e2e_steps // 2, 3), with_step=False)
    env.barrier()
    copy_s = env.max_over_ranks(time.perf_counter() - t0) / max(e2e_steps // 2, 3)
    if env.clocks:
        env.clocks.active = False
    e2e_sps = e2e_steps / e2e_s
    clocks = env.clocks.snapshot() if env.clocks else None
    step.release()
    del kv, step, weights, grads, outs, flat_w, flat_g, h_g, h_w
    torch.cuda.empty_cache()
    if rank != 0:
        return None

    # ---- roofline --------------------------------------------------------------------------------
    peak, peak_src = peaks()
    k_ms = kernel_ms / max(kernel_n, 1)
    if world == 1:
        alg = float(bpe) * nelem
        ach = alg / (k_ms * 1e-3) / 1e9
        traffic, tsrc = measured_traffic(workload) if not args.no_views else (None, "views disabled")
        roof = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": traffic, "traffic_source": tsrc,
                "kernel": "mta_kernel (fused multi-tensor %s update)" % kind,
                "algorithmic_bytes_per_step": alg, "bytes_per_element": bpe,
                "kernel_ms_per_step": k_ms, "kernel_share_of_step": k_ms / ms_step,
                "launches_per_step": launches / steps, "peak_source": peak_src,
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
        t_upd = float(bpe) * nelem / world / (peak * 1e9)
        fused = exchange in ("peer", "peer-push", "nvls") and kind != "lamb"
        # single fused kernel (SGD / Adam on the peer-memory path): time that kernel; otherwise
        # (NCCL path, or LAMB whose peer traffic is split over a reduce pass and the last update
        # pass) the link bytes are spread over the step, so the step time is the denominator
        t_den = (k_ms if fused else ms_step) * 1e-3
        ach = link_bytes / t_den / 1e9
        ach_step = link_bytes / (ms_step * 1e-3) / 1e9
        crossing = sbytes * (1.0 + 1.0 / world) if exchange == "nvls" else link_bytes
        roof = {"bound": "nvlink", "achieved": ach, "peak": 770.0, "unit": "GB/s", "frac": ach / 770.0,
                "traffic": None,
                "kernel": ("mta_kernel over peer memory: fused reduce-scatter + %s update + all-gather" % kind)
                          if exchange != "nccl" else "ncclReduceScatter + sharded update + ncclAllGather",
                "exchange": exchange,
                "link_bytes_per_direction_per_step": link_bytes, "kernel_ms_per_step": k_ms,
                "bytes_actually_crossing_each_link_direction": crossing,
                "frac_of_bytes_actually_crossing": crossing / t_den / 1e9 / 770.0,
                "step_achieved_GBps": ach_step, "step_frac_of_measured_peak": ach_step / 770.0,
                "step_frac_of_nominal_900": ach_step / 900.0,
                "step_lower_bound_ms_nominal": (link_bytes / 900e9 + t_upd) * 1e3,
                "frac_of_unoverlapped_bound": (link_bytes / 900e9 + t_upd) * 1e3 / ms_step,
                "phase_ms_per_step": {k: v[0] / max(v[1], 1) for k, v in phases.items()},
                "peak_source": "measured peer copy 770 GB/s per direction (B200_PROFILING.md); nominal 900",
                "note": "achieved = algorithmic collective bytes per direction (SURVEY 8d: reduce-scatter + "
                        "all-gather, 2*S*(N-1)/N) / fused-kernel time (CUDA events recorded by the library "
                        "around the kernel on rank 0); with NVLS the switch sums and replicates, so fewer "
                        "bytes actually cross the links — that figure is reported beside it"}
    nbytes = nelem * esz
    line = {
        "value": sps, "unit": "steps/s", "ms_per_step": ms_step, "steps": steps, "dtype": dt,
        "config": config_of(workload, world),
        "setup": {"arrays": "caller-owned (pack/unpack copies)" if args.no_views else
                            "zero-copy views into the store (b200kv_view)",
                  "parallelism": "dp%d: one process per GPU, reduce-scatter / sharded update / all-gather" % world
                                 if world > 1 else "single GPU, fused update in place", "exchange": exchange},
        "roofline": roof,
        "e2e": {"value": e2e_sps, "unit": "steps/s", "h2d_bytes_per_step": nbytes,
                "d2h_bytes_per_step": nbytes, "steps": e2e_steps,
                "per_gpu_GBps_each_direction": nbytes * e2e_sps / 1e9,
                "host_link_ceiling": {"steps_per_s": 1.0 / copy_s, "per_gpu_GBps_each_direction": nbytes / copy_s / 1e9,
                                      "what": "the same pinned H2D + D2H copies per step on all %d GPUs at once, "
                                              "no update between them" % world},
                "frac_of_host_link_ceiling": e2e_sps * copy_s,
                "what": "pinned host gradients -> device, kvstore pushpull (bound), updated weights -> pinned "
                        "host; H2D / update / D2H pipelined on three streams"},
        "parity": parity, "gpu_launches": launches, "host_enqueue_ms_per_step": 1e3 * t_enq, "clocks": clocks,
    }
    if cpu_baseline:
        cores = os.cpu_count() or 1
        cs, done, kind_s = cpu_run(workload, 1, args.cpu_steps or 200, 2, cores, budget_s=12)
        line["cpu_baseline"] = {"value": cs, "unit": "steps/s", "cores": cores, "kind": kind_s,
                                "omp_threads": cores, "cpu_model": cpu_model(),
                                "sample": "%d full steps of the same workload on the host cores, gradients "
                                          "already in host memory" % done}
    return line


def run_group(args):
    """--single-process: ONE process drives --gpus N devices through b200kv_create_group, the reference's
    own usage model for kvstore 'device' / 'nccl' (one array per device in every pushpull list).  Same
    step, same metric; with fewer than N visible GPUs the members become virtual ranks on GPU 0."""
    import torch
    import mxnet_maintenance_b200 as mb
    from mxnet_maintenance_b200 import shapes
    from oracle.oracle import Oracle, have_ref
    n = max(args.gpus, 2)
    real = torch.cuda.device_count() >= n
    devs = list(range(n)) if real else [0] * n
    shp_name, dt, kind, bpe, desc = WORKLOADS[args.workload]
    assert kind == "sgd" and dt == "f32", "--single-process measures the fp32 SGD workloads (resnet50, mlp)"
    shp = shapes.CONFIGS[shp_name]()
    sizes = [shapes.numel(s) for s in shp]
    nelem = sum(sizes)
    keys = list(range(len(shp)))
    kw = opt_kwargs(kind, False)
    kv = mb.B200(devices=devs)
    kv.set_optimizer(mb.optimizer.SGD(**kw))
    g0 = torch.Generator(device="cuda:0").manual_seed(99)
    w_init = [(torch.rand(s, device="cuda:0", generator=g0) * 2 - 1) for s in shp]
    kv.init(keys, w_init)
    wv = [[kv.weight_view(k, s, member=m) for m in range(n)] for k, s in zip(keys, shp)]
    gv = [[kv.grad_view(k, s, member=m) for m in range(n)] for k, s in zip(keys, shp)]
    step = kv.bind_pushpull(keys, gv, out=wv)
    gens = [torch.Generator(device="cuda:%d" % d).manual_seed(1234 + m) for m, d in enumerate(devs)]

    def fresh_grads():
        host = []
        for m in range(n):
            per = []
            for k, s in zip(keys, shp):
                g = (torch.rand(s, device=gv[k][m].device, generator=gens[m]) * 2 - 1) * 1e-2
                gv[k][m].copy_(g)
                per.append(g.reshape(-1).cpu().numpy())
            host.append(per)
        return host

    def sync_all():
        for d in sorted(set(devs)):
            torch.cuda.synchronize(d)
    # parity pre-flight: CommCPU reduce over the members' gradients in member order + MultiSGDKernel
    st = ErrStat(*TOL32)
    o = Oracle("ref" if have_ref() else "port", nthread=os.cpu_count() or 1)
    w = [x.reshape(-1).cpu().numpy().copy() for x in w_init]
    mom = [np.zeros(z, np.float32) for z in sizes]
    same = True
    for _ in range(0 if args.no_parity else args.parity_steps):
        hg = fresh_grads()
        sync_all()
        step()
        kv.wait()
        sync_all()
        merged = [o.reduce_sum([hg[m][k].copy() for m in range(n)]) for k in keys]
        o.multi_sgd(w, merged, mom, None, [kw["learning_rate"]] * len(keys), [kw["wd"]] * len(keys), kw["momentum"],
                    kw["rescale_grad"], -1.0)
        for k in keys:
            got = [wv[k][m].reshape(-1).cpu().numpy() for m in range(n)]
            same = same and all(x.tobytes() == got[0].tobytes() for x in got)
            st.add(got[0], w[k])
    if not args.no_parity:
        for k in keys:
            st.add(kv.get_state(k, 0).reshape(-1), mom[k])
    fresh_grads()
    sync_all()
    warmup = max(args.warmup, 3)
    for _ in range(warmup):
        step()
    kv.wait()
    sync_all()
    clocks = Clocks(0)
    l0 = kv.launch_count()
    kv.set_profiling(True)
    kv.profile_read()
    clocks.active = True
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(torch.cuda.current_stream(devs[0]))
    for _ in range(args.steps):
        step()
    e1.record(torch.cuda.current_stream(devs[0]))
    kv.wait()
    sync_all()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks.active = False
    # device time on member 0's stream; the members finish together (in-kernel barriers), and the wall
    # clock around the whole loop is reported beside it
    ms = max(e0.elapsed_time(e1), 0.0)
    kernel_ms, kernel_n = kv.profile_read()
    launches = kv.launch_count() - l0
    ms_step = ms / args.steps
    sbytes = float(nelem) * 4
    link = 2.0 * sbytes * (n - 1) / n
    k_ms = kernel_ms / max(kernel_n, 1)
    par = dict(st.result(), checked=not args.no_parity, steps=args.parity_steps, exchange=kv.last_exchange(),
               all_ranks_bit_identical_weights=same)
    par["pass"] = par["pass"] and same
    line = {"metric": "grad-sync+update steps/s", "value": 1e3 / ms_step, "unit": "steps/s", "n_gpus": n,
            "steps": args.steps, "warmup": warmup, "ms_per_step": ms_step, "wall_ms_per_step": wall_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": dt, "data": "synthetic",
            "config": config_of(args.workload, n),
            "setup": {"parallelism": "ONE process driving %d %s through b200kv_create_group (peer memory, no NCCL)"
                                     % (n, "GPUs" if real else "virtual ranks on GPU 0"),
                      "arrays": "zero-copy views into the store, one per device", "exchange": kv.last_exchange()},
            "roofline": {"bound": "nvlink" if real else "hbm", "achieved": link / (k_ms * 1e-3) / 1e9, "peak": 770.0,
                         "unit": "GB/s", "frac": link / (k_ms * 1e-3) / 1e9 / 770.0, "traffic": None,
                         "kernel": "mta_kernel over peer memory: fused reduce-scatter + sgd update + all-gather "
                                   "(one launch per device per step)",
                         "link_bytes_per_direction_per_step": link, "kernel_ms_per_step": k_ms,
                         "launches_per_step": launches / args.steps},
            "e2e": {"value": None, "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                    "what": "not measured in this mode"},
            "parity": par, "gpu_launches": launches, "clocks": clocks.snapshot()}
    print(json.dumps(line), flush=True)


def run_b200(args):
    env = Env(args)
    rank, world = env.rank, env.world
    done = threading.Event()
    state = {"line": None}

    def emit():
        if rank == 0 and state["line"] is not None:
            print(json.dumps(state["line"]), flush=True)

    def one(workload, steps, e2e_steps, cpu_baseline, psteps):
        if workload == "rsp":
            return run_rsp(args, env, min(steps, 100), e2e_steps, cpu_baseline)
        return run_dense(args, env, workload, steps, e2e_steps, cpu_baseline, psteps)

    psteps = 0 if args.no_parity else args.parity_steps
    head = one(args.workload, args.steps, args.e2e_steps, not args.no_cpu_baseline and world == 1, psteps)
    if rank == 0:
        line = {"metric": "grad-sync+update steps/s", "value": head["value"], "unit": "steps/s", "n_gpus": world,
                "steps": head["steps"], "warmup": max(args.warmup, 3), "ms_per_step": head["ms_per_step"],
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": head["dtype"],
                "data": "synthetic"}
        line.update({k: v for k, v in head.items() if k not in line})
        state["line"] = line
    if not args.no_configs and args.workload == "resnet50":
        # The other BASELINE configurations ride in the same line.  A watchdog prints the line without
        # them if they overrun their budget (a hung collective must not cost the headline).
        def watchdog():
            if not done.wait(args.configs_budget):
                if rank == 0:
                    state["line"].setdefault("configs", {})["_error"] = "extra configurations exceeded %.0f s" \
                        % args.configs_budget
                emit()
                os._exit(0)
        threading.Thread(target=watchdog, daemon=True).start()
        for name, wl in EXTRA_CONFIGS:
            t0 = time.perf_counter()
            try:
                # (the oracle side of the pre-flight gathers every rank's gradients on rank 0: fewer steps
                # for the big configurations when many ranks take part)
                r = one(wl, min(args.steps, 100), min(args.e2e_steps, 10), False, min(psteps, 2 if world < 4 else 1))
            except Exception as e:   # every rank raises alike (same code path); keep the headline
                r = {"error": "%s: %s" % (type(e).__name__, e)}
            if rank == 0:
                r = r or {}
                r["wall_s"] = time.perf_counter() - t0
                state["line"].setdefault("configs", {})[name] = r
        done.set()
    emit()
    if world > 1:
        env.dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    elif a.single_process:
        run_group(a)
    else:
        run_b200(a)
