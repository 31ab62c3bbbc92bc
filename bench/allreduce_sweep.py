"""All-reduce sweep 1 KB - 1 GB (BASELINE.json config 5): this framework's NVLink P2P kernels vs NCCL, same buffers.
Launch with torchrun, one rank per GPU.  Device-timed (CUDA events), max over ranks, bus bandwidth = 2(W-1)/W * S / t.
SWEEP_DTYPE=bf16 sweeps bf16 buffers (two-shot / NVLS with fp32 accumulation; NCCL on a bf16 tensor).
Writes gpurun_out/allreduce_sweep_w<W>[_bf16].json on rank 0."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as tdist

from deeplearning_mpi_b200.parallel import dist

dist.init_process_group("nccl")
rank, world = dist.get_rank(), dist.get_world_size()
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
tdist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
st = dist._require()
max_bytes = int(os.environ.get("SWEEP_MAX_BYTES", str(1 << 30)))
BF16 = os.environ.get("SWEEP_DTYPE", "fp32") == "bf16"
NBS = [int(v) for v in os.environ.get("SWEEP_BLOCKS", "16,32,64").split(",")]
comm = st.new_comm(max_bytes)
stage = comm.tensor(0, max_bytes).view(torch.float32)
nccl_buf = torch.empty(max_bytes // 4, device=dev).view(torch.bfloat16 if BF16 else torch.float32)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fn, iters):
    ts = []
    for _ in range(iters):
        flush.zero_()
        tdist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    t = torch.tensor([sorted(ts)[len(ts) // 2]], device=dev)
    tdist.all_reduce(t, op=tdist.ReduceOp.MAX)
    return t.item() * 1e-3


rows = []
sizes = [1 << k for k in range(10, 31) if (1 << k) <= max_bytes]
sizes += [int(9.0 * 2**20), int(25.3 * 2**20), int(42.65 * 2**20), int(97.49 * 2**20), int(141.0 * 2**20)]
for nbytes in sorted(set(s - s % 16 for s in sizes if s <= max_bytes)):
    iters = 20 if nbytes <= (8 << 20) else 6
    res = {"bytes": nbytes}
    for algo in (("twoshot", "nvls") if BF16 else ("oneshot", "oneshot_bf16", "twoshot", "nvls")):
        if algo.startswith("oneshot") and nbytes > comm.slot_bytes:
            continue
        if algo == "nvls" and not comm.has_multicast:
            continue
        for nb in NBS:
            for _ in range(2):
                comm.allreduce(0, nbytes, 1.0 / world, dtype=int(BF16), algo=algo, nblocks=nb)
            t = timed(lambda: comm.allreduce(0, nbytes, 1.0 / world, dtype=int(BF16), algo=algo, nblocks=nb), iters)
            res[f"{algo}_b{nb}_us"] = t * 1e6
    view = nccl_buf[: nbytes // (2 if BF16 else 4)]
    for _ in range(2):
        tdist.all_reduce(view)
    t = timed(lambda: tdist.all_reduce(view), iters)
    res["nccl_us"] = t * 1e6
    best = min(v for k, v in res.items() if k.endswith("_us") and not k.startswith("nccl"))
    res["ours_best_us"] = best
    res["ours_busbw_GBs"] = 2 * (world - 1) / world * nbytes / (best * 1e-6) / 1e9
    res["nccl_busbw_GBs"] = 2 * (world - 1) / world * nbytes / (res["nccl_us"] * 1e-6) / 1e9
    rows.append(res)
    if rank == 0:
        nv = min([v for k, v in res.items() if k.startswith("nvls")], default=float("nan"))
        ts = min([v for k, v in res.items() if k.startswith("twoshot")], default=float("nan"))
        print(f"{nbytes:>11d} B  twoshot {ts:8.1f} nvls {nv:8.1f} | ours {best:9.1f} us ({res['ours_busbw_GBs']:7.1f} GB/s)   nccl {res['nccl_us']:9.1f} us "
              f"({res['nccl_busbw_GBs']:7.1f} GB/s)", flush=True)
if rank == 0:
    print("multicast (NVLS) mapping:", comm.has_multicast)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump({"world": world, "dtype": "bf16" if BF16 else "fp32", "rows": rows}, open(f"gpurun_out/allreduce_sweep_w{world}{'_bf16' if BF16 else ''}.json", "w"), indent=1)
tdist.barrier()
tdist.destroy_process_group()
dist.destroy_process_group()
