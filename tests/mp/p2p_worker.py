"""Worker for tests/test_multi_gpu.py (torchrun, one rank per GPU): NVLink peer-to-peer collectives and DDP."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as tdist  # noqa: E402

import deeplearning_mpi_b200.nn as dnn  # noqa: E402
from deeplearning_mpi_b200 import optim  # noqa: E402
from deeplearning_mpi_b200.models import resnet18  # noqa: E402
from deeplearning_mpi_b200.parallel import dist  # noqa: E402
from deeplearning_mpi_b200.parallel.ddp import DistributedDataParallel  # noqa: E402

dist.init_process_group("nccl")
rank, world = dist.get_rank(), dist.get_world_size()
assert dist.get_backend() == "p2p", dist.get_backend()
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
# NCCL group only as the oracle for collective numerics
tdist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)

st = dist._require()
comm = st.scratch_comm()
stage = comm.tensor(0, comm.data_bytes).view(torch.float32)

# ---- all-reduce: every algorithm, sizes 1 KB .. 64 MB, unaligned tails, repeated calls (epoch / parity reuse)
torch.manual_seed(1234 + rank)
for nbytes in (1024, 4096 + 16, 65536, 1 << 20, (1 << 20) + 48, 25 << 20, 64 << 20):
    n = nbytes // 4
    for algo in ("twoshot", "oneshot", "oneshot_bf16", "nvls"):
        if algo.startswith("oneshot") and nbytes > comm.slot_bytes:
            continue
        if algo == "nvls" and not comm.has_multicast:
            continue
        for rep in range(3):
            src = torch.randn(n, device=dev)
            ref = src.clone()
            tdist.all_reduce(ref)
            ref /= world
            stage[:n].copy_(src)
            comm.allreduce(0, n * 4, 1.0 / world, algo=algo)
            torch.cuda.synchronize()
            got = stage[:n].clone()
            tol = 2e-2 if algo == "oneshot_bf16" else 1e-5
            err = (got - ref).abs().max().item()
            assert err <= tol * (ref.abs().max().item() + 1e-6), (algo, nbytes, rep, err)
            parts = [torch.empty_like(got) for _ in range(world)]
            tdist.all_gather(parts, got)
            assert all(torch.equal(parts[0], p) for p in parts), f"{algo} {nbytes}: ranks differ bitwise"

# bf16 buckets
n = 1 << 20
srcb = torch.randn(n, device=dev).bfloat16()
refb = srcb.float().clone()
tdist.all_reduce(refb)
stage_b = comm.tensor(0, n * 2).view(torch.bfloat16)
stage_b.copy_(srcb)
comm.allreduce(0, n * 2, 1.0, dtype=1, algo="twoshot")
torch.cuda.synchronize()
assert (stage_b.float() - refb).abs().max().item() <= 3e-2 * refb.abs().max().item()

# ---- flag-protocol stress: thousands of back-to-back collectives of random kind / size / block count, every rank
# delayed by its own random amount before each launch (sizes straddle the one-shot / two-shot / NVLS crossovers).
# Inputs are (rank + 1) * c, so the expected value is known without an oracle call; mismatches are counted on the
# device and read once at the end.
import random  # noqa: E402

rng_all = random.Random(4242)          # same sequence of operations on every rank
rng_me = random.Random(977 + rank)     # this rank's own delays
n_stress = int(os.environ.get("DMPI_STRESS_ITERS", "3000"))
bad = torch.zeros((), dtype=torch.int32, device=dev)
tri = world * (world + 1) / 2
for it in range(n_stress):
    kind = rng_all.choice(["oneshot", "oneshot", "twoshot", "nvls", "bcast"])
    if kind == "nvls" and not comm.has_multicast:
        kind = "twoshot"
    n = rng_all.choice([4, 64, 1024, 16384, 65532, 65536, 65540, 262144, 1 << 20]) + 4 * rng_all.randrange(0, 8)
    if kind == "oneshot" and n * 4 > comm.slot_bytes:
        n = comm.slot_bytes // 4
    nb = rng_all.choice([1, 3, 8, 16, 32, 64])
    c = float(1 + it % 7)
    if rng_me.random() < 0.5:
        torch.cuda._sleep(rng_me.randrange(0, 300000))   # up to ~150 us of skew
    if kind == "bcast":
        root = rng_all.randrange(world)
        stage[:n].fill_(c * (rank + 1))
        comm.broadcast(0, n * 4, root, nblocks=min(nb, 16))
        bad += (stage[:n] != c * (root + 1)).any().int()
    else:
        stage[:n].fill_(c * (rank + 1))
        comm.allreduce(0, n * 4, 1.0, algo=kind, nblocks=nb)
        bad += (stage[:n] != c * tri).any().int()
torch.cuda.synchronize()
assert bad.item() == 0, f"flag-protocol stress: {bad.item()} of {n_stress} collectives returned wrong data"

# ---- broadcast + generic all_reduce / broadcast on user tensors
t = torch.full((1000,), float(rank), device=dev)
dist.broadcast(t, src=world - 1)
assert torch.all(t == float(world - 1))
u = torch.arange(10, device=dev, dtype=torch.float32) + rank
dist.all_reduce(u)
assert torch.allclose(u, torch.arange(10, device=dev, dtype=torch.float32) * world + sum(range(world)))

# ---- send / recv: rank 0 -> everyone (hello_world pattern), then a ring with a 3 MB payload (multi-chunk)
x = torch.zeros(1, device=dev)
if rank == 0:
    x += 42
    for r in range(1, world):
        dist.send(x, dst=r)
else:
    dist.recv(x, src=0)
    assert x.item() == 42
big = torch.full((786433,), float(rank), device=dev)
recv_buf = torch.empty_like(big)
nxt, prv = (rank + 1) % world, (rank - 1) % world
if rank % 2 == 0:
    dist.send(big, dst=nxt)
    dist.recv(recv_buf, src=prv)
else:
    dist.recv(recv_buf, src=prv)
    dist.send(big, dst=nxt)
assert torch.all(recv_buf == float(prv))
dist.barrier()

# ---- DDP on the sm_100a path
torch.manual_seed(100 + rank)
model = resnet18(num_classes=10).to(dev)
ddp = DistributedDataParallel(model, device_ids=[dev.index], output_device=dev.index)


def all_equal(tensor, what):
    parts = [torch.empty_like(tensor) for _ in range(world)]
    tdist.all_gather(parts, tensor.contiguous())
    assert all(torch.equal(parts[0], p) for p in parts), what


for n_, p in ddp.named_parameters():
    all_equal(p.detach(), f"param {n_} differs after wrap")
opt = optim.SGD(ddp.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-5)
g = torch.Generator(device="cpu").manual_seed(7 + rank)
losses = []
for it in range(4):
    xb = torch.randn(16, 3, 64, 64, generator=g).to(dev)
    yb = torch.randint(0, 10, (16,), generator=g).to(dev)
    opt.zero_grad()
    loss = dnn.cross_entropy(ddp(xb), yb)
    loss.backward()
    torch.cuda.synchronize()
    all_equal(ddp.arena.flat_g, f"averaged gradients differ across ranks at iteration {it}")
    opt.step()
    losses.append(loss.item())
all_equal(ddp.arena.flat_p, "parameters diverged")
# BatchNorm buffers: rank 0's running statistics reach every rank (broadcast issued on the comm stream after the
# forward pass, joined by finalize / the next forward)
all_equal(ddp.arena.flat_b, "buffers differ across ranks after the per-iteration broadcast")
assert sorted(ddp.launch_order) == list(range(ddp.n_buckets)), ddp.launch_order

# ---- the reducer averages exactly: inject a known, rank-dependent gradient arena and run the bucket machinery
class _Trigger(torch.autograd.Function):
    @staticmethod
    def forward(ctx, t):
        return t.clone()

    @staticmethod
    def backward(ctx, grad):
        idx = torch.arange(ddp.arena.flat_g.numel(), device=dev, dtype=torch.float32)
        ddp.arena.flat_g.copy_((idx % 1000) * 0.001 + (rank + 1))
        for i in range(len(ddp.arena.slots)):  # slots are in bucket-fill order
            ddp._mark_ready(i)
        return grad


ddp._arm()
_Trigger.apply(torch.ones(1, device=dev, requires_grad=True)).sum().backward()
torch.cuda.synchronize()
idx = torch.arange(ddp.arena.flat_g.numel(), device=dev, dtype=torch.float32)
want = (idx % 1000) * 0.001 + sum(r + 1 for r in range(world)) / world
assert torch.allclose(ddp.arena.flat_g, want, atol=1e-5), "bucket all-reduce is not the mean over ranks"
all_equal(ddp.arena.flat_g, "injected gradients differ across ranks after the reduce")

# ---- real gradients: the synchronised gradient is the mean of the local (no_sync) ones, up to the run-to-run noise of
# this tiny, randomly initialised network (fp32 atomics reorder sums by ~1e-7, bf16 + small-batch BatchNorm amplify
# that chaotically: two identical local passes already differ by tens of percent here)
xb = torch.randn(16, 3, 64, 64, generator=g).to(dev)
yb = torch.randint(0, 10, (16,), generator=g).to(dev)


def local_grads():
    opt.zero_grad()
    with ddp.no_sync():
        dnn.cross_entropy(ddp(xb), yb).backward()
    torch.cuda.synchronize()
    return ddp.arena.flat_g.clone()


ga, gb = local_grads(), local_grads()
noise = ((ga - gb).norm() / (gb.norm() + 1e-12)).item()
expected = ga.clone()
tdist.all_reduce(expected)
expected /= world
opt.zero_grad()
dnn.cross_entropy(ddp(xb), yb).backward()
torch.cuda.synchronize()
rel = ((ddp.arena.flat_g - expected).norm() / (expected.norm() + 1e-12)).item()
assert rel <= 3 * noise + 2e-2, f"DDP gradient is not the mean over ranks (rel L2 err {rel}, run-to-run noise {noise})"

# rank-0-only eval, then a joint step
if rank == 0:
    ddp.eval()
    with torch.no_grad():
        ddp(xb)
    ddp.train()
opt.zero_grad()
dnn.cross_entropy(ddp(xb), yb).backward()
opt.step()
torch.cuda.synchronize()
all_equal(ddp.arena.flat_p, "parameters diverged after the rank-0-only eval")

# ---- loss curves: this framework's DDP against stock torch DDP (torchvision ResNet-18, cuDNN, fp32, NCCL) from the same
# weights, every rank on its own batches (multi-GPU form)
import torch.nn.functional as F  # noqa: E402
import torchvision  # noqa: E402

torch.manual_seed(77)                      # same initial weights on every rank and in both stacks
ours_m = resnet18(num_classes=10).to(dev)
stock_m = torchvision.models.resnet18(weights=None)
stock_m.fc = torch.nn.Linear(512, 10)
stock_m.load_state_dict(ours_m.state_dict())
stock_m = stock_m.to(dev).train()
ours_ddp = DistributedDataParallel(ours_m, device_ids=[dev.index], output_device=dev.index)
stock_ddp = torch.nn.parallel.DistributedDataParallel(stock_m, device_ids=[dev.index], output_device=dev.index)
o_opt = optim.SGD(ours_ddp.parameters(), lr=0.02, momentum=0.9, weight_decay=1e-5)
s_opt = torch.optim.SGD(stock_ddp.parameters(), lr=0.02, momentum=0.9, weight_decay=1e-5)
gen = torch.Generator().manual_seed(5)
proto = torch.randn(10, 3, 32, 32, generator=gen)            # shared class prototypes: a learnable task
gen_r = torch.Generator().manual_seed(900 + rank)            # rank-specific samples
lo, ls = [], []
for step_i in range(16):
    yb = torch.randint(0, 10, (64,), generator=gen_r)
    xb = (proto[yb] + 0.7 * torch.randn(64, 3, 32, 32, generator=gen_r)).to(dev)
    yb = yb.to(dev)
    o_opt.zero_grad()
    l1 = dnn.cross_entropy(ours_ddp(xb), yb)
    l1.backward()
    o_opt.step()
    s_opt.zero_grad()
    l2 = F.cross_entropy(stock_ddp(xb), yb)
    l2.backward()
    s_opt.step()
    lo.append(l1.item())
    ls.append(l2.item())
dl = [abs(a - b) for a, b in zip(lo, ls)]
assert lo[-1] < 0.6 * lo[0] and ls[-1] < 0.6 * ls[0], (lo, ls)
assert dl[0] < 2e-2 and sum(dl) / len(dl) < 0.12 and max(dl) < 0.35, f"rank {rank}: loss curves diverge {list(zip(lo, ls))}"
all_equal(ours_ddp.arena.flat_p, "parameters diverged in the loss-curve run")
del stock_ddp

# ---- bf16 wire format (DistributedDataParallel(grad_wire_dtype="bf16")): large buckets travel as bf16, sums in fp32
torch.manual_seed(300 + rank)
model16 = resnet18(num_classes=10).to(dev)
ddp16 = DistributedDataParallel(model16, device_ids=[dev.index], output_device=dev.index, grad_wire_dtype="bf16")


class _Trigger16(torch.autograd.Function):
    @staticmethod
    def forward(ctx, t):
        return t.clone()

    @staticmethod
    def backward(ctx, grad):
        idx16 = torch.arange(ddp16.arena.flat_g.numel(), device=dev, dtype=torch.float32)
        ddp16.arena.flat_g.copy_(((idx16 % 977) * 0.003 - 1.0) * (rank + 1))
        for i in range(len(ddp16.arena.slots)):
            ddp16._mark_ready(i)
        return grad


ddp16._arm()
_Trigger16.apply(torch.ones(1, device=dev, requires_grad=True)).sum().backward()
torch.cuda.synchronize()
idx16 = torch.arange(ddp16.arena.flat_g.numel(), device=dev, dtype=torch.float32)
want16 = ((idx16 % 977) * 0.003 - 1.0) * (sum(r + 1 for r in range(world)) / world)
err16 = (ddp16.arena.flat_g - want16).abs().max().item()
assert err16 <= 1.2e-2 * want16.abs().max().item(), f"bf16-wire all-reduce error {err16}"
all_equal(ddp16.arena.flat_g, "bf16-wire gradients differ across ranks")

tdist.barrier()
tdist.destroy_process_group()
dist.destroy_process_group()
print(f"P2P_WORKER_OK rank={rank} multicast={comm.has_multicast} losses={['%.3f' % l for l in losses]}")
