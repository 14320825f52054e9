"""Group store (one process, several GPUs) over the NVSwitch multicast exchange: runs
tests/group_nvls_check.py with B200KV_NVLS=1 on two real devices (and on four when the box has them)."""
import os
import subprocess
import sys

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _multicast_supported(n):
    """CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED on the first n devices (unknown counts as yes: the check decides)."""
    try:
        from cuda.bindings import driver as cu
        cu.cuInit(0)
        for d in range(n):
            err, dev = cu.cuDeviceGet(d)
            err2, v = cu.cuDeviceGetAttribute(cu.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev)
            if int(err) == 0 and int(err2) == 0 and not v:
                return False
    except Exception:
        pass
    return True


@pytest.mark.parametrize("n", [2, 4])
def test_group_nvls(n):
    if torch.cuda.device_count() < n:
        pytest.skip("needs %d GPUs" % n)
    if not _multicast_supported(n):
        pytest.skip("no NVSwitch multicast on this box")
    env = dict(os.environ, B200KV_NVLS="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "group_nvls_check.py"), str(n)],
                       capture_output=True, text=True, timeout=900, env=env)
    sys.stdout.write(r.stdout[-4000:])
    sys.stderr.write(r.stderr[-4000:])
    assert r.returncode == 0 and "GROUP_NVLS_CHECK OK" in r.stdout
