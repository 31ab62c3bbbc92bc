"""Multi-GPU tests (>= 2 devices): NVLink P2P collectives vs NCCL, send/recv, DDP on the sm_100a path.
Skipped on single-GPU boxes; run with `gpurun --gpus 2 -- python -m pytest tests/test_multi_gpu.py -m gpu`."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _torchrun(script, *args, nproc=2, timeout=900):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), script, *args]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)
    return r.returncode, r.stdout + r.stderr


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_p2p_collectives_and_ddp():
    n = min(_ngpu(), 8)
    rc, out = _torchrun(os.path.join(ROOT, "tests", "mp", "p2p_worker.py"), nproc=n)
    assert rc == 0, out[-6000:]
    assert out.count("P2P_WORKER_OK") == n


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_hello_world_nvlink():
    rc, out = _torchrun(os.path.join(ROOT, "examples", "hello_world", "hello_world.py"), "--backend", "nccl", nproc=2)
    assert rc == 0, out[-4000:]
    assert "worker_0 sent data to Rank 1" in out and "worker_1 has received data from rank 0" in out
