"""One process, N real GPUs, NVSwitch multicast exchange (B200KV_NVLS=1 forces it below 4 devices):
the group's members allocate their weight segments / gradient buckets from one multicast arena
(csrc/mcast.cu) and run the same multimem.ld_reduce / multimem.st kernel as the one-process-per-GPU
mode.  Launched by tests/test_group_nvls_gpu.py in a process of its own (the library reads the
environment once).  The switch adds the N copies in its own order, so fp32 results are compared within
the north-star tolerance (rtol 1e-5, atol 1e-6) and fp16 SGD within the fp16 one; every device copy must
hold the same bits."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    from oracle.oracle import Oracle
    import test_group_gpu as tg
    n = int(sys.argv[1]) if len(sys.argv) > 1 else torch.cuda.device_count()
    devs = list(range(n))
    oracle = Oracle("port")
    want = "nvls"
    cases = [("sgd", dict()), ("adam", dict()), ("sgd", dict(mp=True)), ("sgd", dict(batch=False)),
             (None, dict()), ("lamb", dict())]
    for kind, kw in cases:
        kv, res, states = tg._run(oracle, devs, kind, adjudicate=True, **kw)
        used = kv.last_exchange()
        # (forced on, LAMB too reduces its shard through the switch; by default it keeps the peer loads)
        assert used == want, "%s %s: exchange %s, expected %s" % (kind, kw, used, want)
        if kw.get("mp"):
            tg._check(res, states, exact=False, rtol=2e-3, atol=1e-5)
        else:
            tg._check(res, states, exact=False)
        print("case", kind, kw, "exchange", used, "ok", flush=True)
    print("GROUP_NVLS_CHECK OK")


if __name__ == "__main__":
    main()
