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

# control-plane helpers: rank 0's object reaches every rank; the segmentation dataset shares rank 0's palette scan
obj = dist.broadcast_object({"palette": [0, 255], "who": rank} if rank == 0 else None, src=0)
assert obj == {"palette": [0, 255], "who": 0}
import numpy as np  # noqa: E402
from PIL import Image  # noqa: E402

from deeplearning_mpi_b200.utils.data import SegmentationFolder  # noqa: E402

ds_root = os.path.join(out_dir, "seg")
if rank == 0:
    os.makedirs(os.path.join(ds_root, "images"))
    os.makedirs(os.path.join(ds_root, "masks"))
    rng = np.random.default_rng(0)
    for i in range(4):
        Image.fromarray(rng.integers(0, 255, (16, 16, 3), dtype=np.uint8)).save(os.path.join(ds_root, "images", f"a{i}.png"))
        Image.fromarray(((rng.random((16, 16)) > 0.5) * 200).astype(np.uint8)).save(os.path.join(ds_root, "masks", f"a{i}.png"))
tdist.barrier()
ds = SegmentationFolder.shared(os.path.join(ds_root, "images"), os.path.join(ds_root, "masks"), scale=0.5)
assert ds.mask_values == [0, 200] and len(ds) == 4 and ds[1]["mask"].shape == (8, 8)

# the bucket launch order was re-derived from the first backward pass and agrees on both ranks
assert sorted(ddp.launch_order) == list(range(ddp.n_buckets))
lo = gathered(torch.tensor(ddp.launch_order, dtype=torch.int64))
assert torch.equal(lo[0], lo[1]) and ddp._order_final

# frozen parameters: wrapped, trained around, never bucketed or changed
torch.manual_seed(7)
fz = resnet18(num_classes=10)
for p_ in fz.layer1.parameters():
    p_.requires_grad_(False)
before = [p_.detach().clone() for p_ in fz.layer1.parameters()]
fz_ddp = DistributedDataParallel(fz)
fz_opt = optim.SGD([p_ for p_ in fz_ddp.parameters() if p_.requires_grad], lr=0.05, momentum=0.9)
for _ in range(2):
    fz_opt.zero_grad()
    crit(fz_ddp(torch.randn(4, 3, 32, 32)), torch.randint(0, 10, (4,))).backward()
    fz_opt.step()
assert all(torch.equal(a, b.detach()) for a, b in zip(before, fz.layer1.parameters()))
assert len(fz_ddp._slots) == sum(1 for p_ in fz.parameters() if p_.requires_grad) < len(fz_ddp.arena.slots)
for n_, p_ in fz_ddp.named_parameters():
    a, b = gathered(p_.detach())
    assert torch.equal(a, b), f"param {n_} diverged with frozen layers"

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
