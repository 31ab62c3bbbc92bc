"""NVLink peer-to-peer bandwidth between the GPUs of one node, measured with this framework's own copy kernel over the
symmetric arena's peer mappings (the roofline denominator of the collectives).
Launch with torchrun, one rank per GPU.  Device-timed, best of 5, 256 MiB per transfer.

  one pair at a time : rank 0 reads from / writes to every peer while the others idle   (per-link figure)
  all ranks at once  : every rank reads from (rank + k) % W, k = 1 .. W-1, concurrently   (switch under full load)
Writes gpurun_out/p2p_bandwidth_w<W>.json on rank 0."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from deeplearning_mpi_b200 import _native as N
from deeplearning_mpi_b200.parallel import dist

dist.init_process_group("nccl")
rank, world = dist.get_rank(), dist.get_world_size()
st = dist._require()
dev = st.device
lib = N.cuda()
nbytes = int(os.environ.get("P2P_BYTES", str(256 << 20)))
comm = st.new_comm(2 * nbytes)                       # [0, nbytes): source region, [nbytes, 2 nbytes): sink region
local = lib.dmpi_arena_data(comm.handle, rank)
peers = [lib.dmpi_arena_data(comm.handle, r) for r in range(world)]


def timed(dst, src, blocks):
    best = 1e9
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        lib.dmpi_peer_copy(dst, src, nbytes, blocks, N.stream_ptr())
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return nbytes / (best * 1e-3) / 1e9


res = {"world": world, "bytes": nbytes, "pairs": [], "concurrent": []}
blocks = int(os.environ.get("P2P_BLOCKS", "64"))
res["local_copy_GBs_rw"] = None
if rank == 0:
    res["local_copy_GBs_rw"] = 2 * timed(local + nbytes, local, 148 * 4)   # read + write bytes, like MEASURED_PEAKS
for peer in range(1, world):
    dist.barrier()
    if rank == 0:
        rd = timed(local + nbytes, peers[peer], blocks)          # peer -> local
        wr = timed(peers[peer] + nbytes, local, blocks)          # local -> peer
        res["pairs"].append({"peer": peer, "read_GBs": rd, "write_GBs": wr})
        print(f"rank 0 <-> {peer}: read {rd:7.1f} GB/s   write {wr:7.1f} GB/s", flush=True)
    dist.barrier()
for k in range(1, world):
    dist.barrier()
    rd = timed(local + nbytes, peers[(rank + k) % world], blocks)
    vals = st.store.all_gather(f"p2pbw{k}", rank, world, repr(rd).encode()) if world > 1 else [repr(rd).encode()]
    vals = [float(v.decode()) for v in vals]
    if rank == 0:
        res["concurrent"].append({"shift": k, "min_GBs": min(vals), "mean_GBs": sum(vals) / len(vals)})
        print(f"all ranks read from rank+{k}: min {min(vals):7.1f}  mean {sum(vals) / len(vals):7.1f} GB/s per GPU", flush=True)
if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open(f"gpurun_out/p2p_bandwidth_w{world}.json", "w"), indent=1)
dist.barrier()
dist.destroy_process_group()
