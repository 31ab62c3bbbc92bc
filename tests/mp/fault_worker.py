"""Worker for tests/test_cpu_dist.py::test_fault_injection_* (torchrun, gloo, world_size 2): a few DDP steps that the
DMPI_FAULT hook interrupts on one rank; the survivors must fail (non-zero exit), not hang."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import datetime  # noqa: E402

import torch  # noqa: E402

import deeplearning_mpi_b200.nn as dnn  # noqa: E402
from deeplearning_mpi_b200 import optim  # noqa: E402
from deeplearning_mpi_b200.models import resnet18  # noqa: E402
from deeplearning_mpi_b200.parallel import dist  # noqa: E402
from deeplearning_mpi_b200.parallel.ddp import DistributedDataParallel  # noqa: E402

dist.init_process_group("gloo", timeout=datetime.timedelta(seconds=float(os.environ.get("FAULT_TEST_TIMEOUT_S", "60"))))
rank = dist.get_rank()
torch.manual_seed(0)
ddp = DistributedDataParallel(resnet18(num_classes=10))
opt = optim.SGD(ddp.parameters(), lr=0.01)
crit = dnn.CrossEntropyLoss()
try:
    for it in range(4):
        opt.zero_grad()
        crit(ddp(torch.randn(2, 3, 32, 32)), torch.randint(0, 10, (2,))).backward()
        opt.step()
        print(f"FAULT_WORKER rank={rank} finished step {it}", flush=True)
finally:
    dist.destroy_process_group()
print(f"FAULT_WORKER_DONE rank={rank}", flush=True)
