"""BASELINE config 5 on real devices: alice / bob / carole on three GPUs of one box (needs >= 3 devices; the
single-GPU lease of the regular suite skips it, `gpurun --gpus 4 -- python -m pytest tests/test_gpu_party3.py -m gpu`
runs it).  One process per party under torchrun; reshares and openings over the peer-memory rings (NVLink) and, second
case, over the library's NCCL transport.  Each party has its own entropy, so the check is by value against the float
model, as in tests/test_gpu_party.py."""
import ctypes
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _device_count() -> int:
    try:
        cuda = ctypes.CDLL("libcuda.so.1")
        if cuda.cuInit(0) != 0:
            return 0
        n = ctypes.c_int(0)
        cuda.cuDeviceGetCount(ctypes.byref(n))
        return n.value
    except Exception:
        return 0


@pytest.mark.skipif(_device_count() < 3, reason="needs three GPUs (one per party)")
@pytest.mark.parametrize("transport", ["--peer-memory", "--native-nccl"])
def test_mlp_three_parties_on_three_gpus(transport):
    port = 29700 + os.getpid() % 200
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "3", "--master-addr", "127.0.0.1",
           "--master-port", str(port), "-m", "moose_b200.party", "--mlp", "4096", "--fused", transport, "--repeat", "2"]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1]
    rec = json.loads(line)
    assert rec["parties"] == ["player0", "player1", "player2"]
    assert rec["max_abs_err_first_rows"] < 5e-3
    assert rec["payload_bytes_per_run"] > 0 and rec["rows_per_s"] > 0
