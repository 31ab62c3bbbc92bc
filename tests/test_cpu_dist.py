"""Multi-process CPU tests (gloo, world_size 2) of the distributed plumbing — BASELINE.json config 1 and the DDP
semantics of SURVEY.md section 4 item 2.  Each test launches torchrun in a subprocess on 127.0.0.1."""
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module", autouse=True)
def _built():
    sys.path.insert(0, ROOT)
    import build

    build.build()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _torchrun(script, *args, nproc=2, timeout=600, cwd=None):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), script, *args]
    env = dict(os.environ, OMP_NUM_THREADS="2")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=cwd or ROOT, env=env)
    return r.returncode, r.stdout + r.stderr


def test_hello_world_gloo_world2():
    rc, out = _torchrun(os.path.join(ROOT, "examples", "hello_world", "hello_world.py"), "--backend", "gloo")
    assert rc == 0, out
    assert "worker_0 sent data to Rank 1" in out
    assert "worker_1 has received data from rank 0" in out


def test_hello_world_requires_torchrun_env():
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE")}
    r = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "hello_world", "hello_world.py")],
                       capture_output=True, text=True, env=env, timeout=120)
    assert r.returncode != 0 and "environment variables" in r.stderr


def test_ddp_semantics_gloo_world2(tmp_path):
    rc, out = _torchrun(os.path.join(ROOT, "tests", "mp", "ddp_worker.py"), str(tmp_path))
    assert rc == 0, out[-4000:]
    assert out.count("DDP_WORKER_OK") == 2


def test_resnet_example_gloo_world2(tmp_path):
    (tmp_path / "saved_models").mkdir()
    rc, out = _torchrun(os.path.join(ROOT, "examples", "resnet", "main.py"), "--backend", "gloo", "--synthetic",
                        "--synthetic_size", "32", "--batch_size", "4", "--num_epochs", "1", "--steps_per_epoch", "2",
                        cwd=str(tmp_path))
    assert rc == 0, out[-4000:]
    assert "Epoch 0 completed" in out and (tmp_path / "saved_models" / "resnet_distributed.pth").exists()


def test_unet_example_gloo_world2(tmp_path):
    (tmp_path / "saved_models").mkdir()
    (tmp_path / "logs").mkdir()
    rc, out = _torchrun(os.path.join(ROOT, "examples", "unet", "train.py"), "--backend", "gloo", "--synthetic",
                        "--synthetic_size", "10", "--image_size", "32", "--batch_size", "2", "--num_epochs", "1",
                        cwd=str(tmp_path))
    assert rc == 0, out[-4000:]
    assert "FINAL DICE COEFFICIENT" in out and (tmp_path / "saved_models" / "model.pth").exists()
    logs = list((tmp_path / "logs").glob("training_log_*.log"))
    assert len(logs) >= 1 and "FINAL TRAINING RESULTS" in logs[0].read_text()


@pytest.mark.parametrize("kind", ["exit", "raise", "stall=150"])
def test_fault_injection_never_hangs(kind):
    """a rank that dies (exit), throws (raise) or goes silent (stall) in the middle of training must turn
    into a non-zero exit code of the job within the collective timeout, never into a hang or a clean exit."""
    import time

    env = dict(os.environ, OMP_NUM_THREADS="2", FAULT_TEST_TIMEOUT_S="45",
               DMPI_FAULT=f"rank=1,step=2,{kind if kind.startswith('stall') else 'kind=' + kind}")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "mp", "fault_worker.py")]
    t0 = time.time()
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=400, cwd=ROOT, env=env)
    out = r.stdout + r.stderr
    assert r.returncode != 0, out[-3000:]
    assert "FAULT_WORKER_DONE rank=0" not in out, out[-3000:]
    assert "FAULT_WORKER rank=0 finished step 1" in out and "FAULT_WORKER rank=0 finished step 2" not in out, out[-3000:]
    assert time.time() - t0 < 300  # the collective timeout of the survivors is 45 s
