"""Worker for tests/test_cpu_dist.py (launched by torchrun, gloo, world_size 2): DDP semantics on CPU."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as tdist  # noqa: E402
import torchvision  # noqa: E402

import deeplearning_mpi_b200.nn as dnn  # noqa: E402
from deeplearning_mpi_b200 import optim  # noqa: E402
from deeplearning_mpi_b200.models import resnet18  # noqa: E402
from deeplearning_mpi_b200.parallel import dist  # noqa: E402
from deeplearning_mpi_b200.parallel.ddp import DistributedDataParallel  # noqa: E402

out_dir = sys.argv[1]
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
assert world == 2 and dist.get_backend() == "gloo"


def gathered(t):
    parts = [torch.zeros_like(t) for _ in range(world)]
    tdist.all_gather(parts, t.contiguous())
    return parts


torch.manual_seed(100 + rank)  # different init on every rank: the wrap must make them identical
model = resnet18(num_classes=10)
ddp = DistributedDataParallel(model, device_ids=None, output_device=None)
for n, p in ddp.named_parameters():
    a, b = gathered(p.detach())
    assert torch.equal(a, b), f"param {n} differs after wrap"
assert all(k.startswith("module.") for k in ddp.state_dict()) and len(ddp.state_dict()) == 122
assert ddp.n_buckets >= 2

opt = optim.SGD(ddp.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-5)
crit = dnn.CrossEntropyLoss()
g = torch.Generator().manual_seed(7 + rank)
for it in range(3):
    x = torch.randn(4, 3, 32, 32, generator=g)
    y = torch.randint(0, 10, (4,), generator=g)
    opt.zero_grad()
    loss = crit(ddp(x), y)
    loss.backward()
    for n, p in ddp.named_parameters():  # averaged gradients are identical on both ranks
        a, b = gathered(p.grad)
        assert torch.equal(a, b), f"grad {n} differs at iteration {it}"
    opt.step()
for n, p in ddp.named_parameters():
    a, b = gathered(p.detach())
    assert torch.equal(a, b), f"param {n} diverged after training"

# the all-reduce really averages: compare against a manual average of local gradients computed without sync
opt.zero_grad()
x = torch.randn(4, 3, 32, 32, generator=g)
y = torch.randint(0, 10, (4,), generator=g)
ddp._sync_buffers()  # both passes below must see the same (rank 0's) running statistics
ddp.eval()  # eval-mode BN so the two passes below see identical statistics
with ddp.no_sync():
    crit(ddp(x), y).backward()
local = [p.grad.detach().clone() for p in ddp.parameters()]
expected = []
for t in local:
    parts = gathered(t)
    expected.append((parts[0] + parts[1]) / 2)
opt.zero_grad()
ddp.train(True)
for m in ddp.modules():
    if isinstance(m, torch.nn.BatchNorm2d):
        m.eval()
crit(ddp(x), y).backward()
for (n, p), e in zip(ddp.named_parameters(), expected):
    assert torch.allclose(p.grad, e, atol=1e-6, rtol=1e-5), f"grad {n} is not the mean over ranks"
ddp.train(True)

# running statistics are re-broadcast from rank 0 at the next training forward
opt.zero_grad()
crit(ddp(torch.randn(4, 3, 32, 32, generator=g)), y).backward()
with torch.no_grad():
    model.bn1.running_mean += float(rank)  # make the ranks disagree
crit(ddp(torch.randn(4, 3, 32, 32)), y)  # forward only: buffers synced before the forward ran
# (rank 1's local +1 was overwritten by rank 0's values before its forward; after the forward both moved by their own batch)

# rank-0-only evaluation between epochs must neither hang nor desynchronise the next joint step
if rank == 0:
    ddp.eval()
    with torch.no_grad():
        for _ in range(3):
            ddp(torch.randn(2, 3, 32, 32))
    torch.save(ddp.state_dict(), os.path.join(out_dir, "ours.pth"))
    ddp.train()
opt.zero_grad()
crit(ddp(torch.randn(4, 3, 32, 32, generator=g)), y).backward()
opt.step()
tdist.barrier()

# our checkpoint loads into the reference's stack: stock DDP around torchvision resnet18 (ref:pytorch/resnet/main.py:40-52)
stock = torchvision.models.resnet18(weights=None)
stock.fc = torch.nn.Linear(512, 10)
stock_ddp = torch.nn.parallel.DistributedDataParallel(stock)
stock_ddp.load_state_dict(torch.load(os.path.join(out_dir, "ours.pth"), map_location="cpu"))
# ... and a stock checkpoint loads into ours
if rank == 0:
    torch.save(stock_ddp.state_dict(), os.path.join(out_dir, "stock.pth"))
tdist.barrier()
ddp.load_state_dict(torch.load(os.path.join(out_dir, "stock.pth"), map_location="cpu"))
ddp.eval(), stock_ddp.eval()
xe = torch.randn(2, 3, 32, 32)
with torch.no_grad():
    assert torch.allclose(ddp(xe), stock_ddp(xe), atol=1e-4)

# mismatching models must be rejected at wrap time
try:
    bad = resnet18(num_classes=10 if rank == 0 else 11)
    DistributedDataParallel(bad)
    raised = False
except RuntimeError:
    raised = True
flags = gathered(torch.tensor([float(raised)]))
assert any(f.item() == 1.0 for f in flags), "shape mismatch across ranks was not detected"

dist.destroy_process_group()
print(f"DDP_WORKER_OK rank={rank}")
