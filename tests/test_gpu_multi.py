"""n8 on REAL separate GPUs (VERDICT r01 "missing" #1): one process per device under torchrun + NCCL; the merged images of
the p2p (native pipelined peer-memory frame loop), tiles and samples partitionings are compared with the CPU oracle.
Skipped with fewer than two devices — run with `gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu`."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def device_count():
    from pathtracing_b200 import _lib
    return _lib.load().ptb_device_count()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_multi_gpu_partitionings_across_processes(world):
    if device_count() < world:
        pytest.skip(f"needs {world} CUDA devices")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(29600 + world), os.path.join(HERE, "multi_gpu_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    out = r.stdout + r.stderr
    assert r.returncode == 0 and "MULTI_OK" in out, out[-4000:]
