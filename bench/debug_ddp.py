import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as tdist
import deeplearning_mpi_b200.nn as dnn
from deeplearning_mpi_b200 import optim
from deeplearning_mpi_b200.models import resnet18
from deeplearning_mpi_b200.parallel import dist
from deeplearning_mpi_b200.parallel.ddp import DistributedDataParallel
dist.init_process_group("nccl"); rank, world = dist.get_rank(), dist.get_world_size()
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
tdist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
torch.manual_seed(100)
model = resnet18(num_classes=10).to(dev)
ddp = DistributedDataParallel(model, device_ids=[dev.index])
opt = optim.SGD(ddp.parameters(), lr=0.05, momentum=0.9)
g = torch.Generator().manual_seed(7 + rank)
xb = torch.randn(16, 3, 64, 64, generator=g).to(dev); yb = torch.randint(0, 10, (16,), generator=g).to(dev)
def local():
    opt.zero_grad()
    with ddp.no_sync():
        dnn.cross_entropy(ddp(xb), yb).backward()
    torch.cuda.synchronize(); return ddp.arena.flat_g.clone()
def rel(a, b): return ((a - b).norm() / (b.norm() + 1e-12)).item()
A = local(); B = local()
mean = A.clone(); tdist.all_reduce(mean); mean /= world
opt.zero_grad(); dnn.cross_entropy(ddp(xb), yb).backward(); torch.cuda.synchronize(); Cg = ddp.arena.flat_g.clone()
if rank == 0:
    print("A vs B (nondeterminism):", rel(A, B)); print("C vs mean(A):", rel(Cg, mean)); print("C vs A(local):", rel(Cg, A))
    for b, (lo, n) in enumerate(ddp.bucket_ranges):
        print(f" bucket {b}: off {lo} n {n} C-vs-mean {rel(Cg[lo:lo+n], mean[lo:lo+n]):.3e}  C-vs-local {rel(Cg[lo:lo+n], A[lo:lo+n]):.3e}")
    for s in ddp.arena.slots[:12]:
        print(f"   {s.name:32s} bucket {ddp.bucket_of[s.index]} C-vs-mean {rel(Cg[s.offset:s.offset+s.numel], mean[s.offset:s.offset+s.numel]):.3e}")
tdist.barrier(); tdist.destroy_process_group(); dist.destroy_process_group()
