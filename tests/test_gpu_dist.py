"""Multi-GPU parity (needs >= 2 CUDA devices; skipped otherwise): sharded engine vs oracle."""
import os
import subprocess
import sys

import pytest

from qiskit_b200 import _lib

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(world):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29700 + world),
           os.path.join(ROOT, "tests", "dist_gpu_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert res.stdout.count("DIST_GPU_OK") == world, res.stdout[-3000:]
    return res.stdout


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_engine_against_oracle(world):
    if _lib.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    print(_run(world))
