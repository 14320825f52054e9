#!/bin/bash
# usage: tests/run_n.sh N workload...   -- one bench JSON line per workload at N GPUs
N=$1; shift
for WL in "$@"; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29540 \
    bench.py --gpus $N --steps 100 --warmup 10 --workload $WL --e2e-steps 5 2>&1 | grep '^{' | tail -1
done
