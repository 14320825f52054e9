"""Launches tests/multi_gpu_check.py under torchrun on every GPU of the box (needs >= 2)."""
import os
import subprocess
import sys

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("exchange", ["default", "peer"])
def test_multi_gpu_parity(exchange):
    """default: the exchange the store picks (NVSwitch multicast from 4 GPUs: fp32 tolerance);
    peer: B200KV_NVLS=0, ranks summed in CommCPU order over peer memory: must be bit-exact."""
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs (run with gpurun --gpus 2)")
    n = 2 if n < 4 else (4 if n < 8 else 8)
    if exchange == "peer" and n < 4:
        pytest.skip("below 4 GPUs the default exchange already is the peer-memory path")
    env = dict(os.environ)
    if exchange == "peer":
        env["B200KV_NVLS"] = "0"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n),
           "--master-addr", "127.0.0.1", "--master-port", "29533",
           os.path.join(ROOT, "tests", "multi_gpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=1500, env=env)
    sys.stdout.write(r.stdout[-6000:])
    sys.stderr.write(r.stderr[-4000:])
    assert r.returncode == 0 and "MULTI_GPU_CHECK OK" in r.stdout
