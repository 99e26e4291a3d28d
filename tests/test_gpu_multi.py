"""Multi-GPU parity on hardware: launches tests/multi_gpu_worker.py under
torchrun with one process per GPU.  Needs at least two devices (`gpurun --gpus 2`); skipped on a
single-GPU box, where tests/test_gpu_shards.py and the emulated-shard tests of tests/test_gpu_solver.py
cover the same host logic on one device."""
import json
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2])
def test_multi_gpu_worker(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs, found {torch.cuda.device_count()}")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "multi_gpu_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-3000:] + "\n" + res.stderr[-6000:]
    line = [ln for ln in res.stdout.splitlines() if ln.startswith("{")][-1]
    report = json.loads(line)
    assert report["world"] == world
    assert report["function_distributed"] == "ok"
    assert any(k.startswith("solver_peer_float64") for k in report)
