T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
port=29520
for w in resnet50 bert_base_sgd; do
for c in 1 2 0; do
  port=$((port+1))
  B200KV_NVLS_CTAS=$c timeout 300 $T --master-port $port bench.py --gpus 4 --workload $w --no-configs --no-cpu-baseline --e2e-steps 2 2>/dev/null | grep '^{' | python -c "
import sys, json
d=json.loads(sys.stdin.readline()); r=d['roofline']
print('$w', 'ctas=$c', 'ms', round(d['ms_per_step'],4), 'kernel', round(r['kernel_ms_per_step'],4), r['exchange'], d['parity']['pass'], d['parity'].get('max_err_over_tol'))
"
done; done
