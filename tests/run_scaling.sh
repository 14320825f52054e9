#!/bin/bash
# usage: tests/run_scaling.sh N [workload]   -- prints one JSON line per variant (peer-memory plain / staged, NCCL)
N=$1; WL=${2:-resnet50}
run() { env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29540 bench.py --gpus $N --steps 100 --warmup 10 --workload $WL --e2e-steps 5 2>&1 | grep '^{' | tail -1; }
echo "== p2p plain";  run B200KV_STAGED=0
echo "== p2p staged"; run B200KV_STAGED=1
echo "== nccl";       run B200KV_P2P=0
