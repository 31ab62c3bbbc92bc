"""Point-to-point hello world (capability of ref:pytorch/hello_world/hello_world.py).

Rank 0 sends a 1-element tensor to every other rank, which receives it.  ``--backend nccl`` runs over this
framework's NVLink put-with-flag kernels on one node (NCCL only as the multi-node fallback); ``--backend gloo`` runs on
CPU.  Launch with torchrun (or examples/hello_world/run.sh):

    torchrun --nproc_per_node 2 --nnodes 1 --master_addr 127.0.0.1 --master_port 29500 \
        examples/hello_world/hello_world.py --backend gloo
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _bootstrap  # noqa: F401,E402

import torch  # noqa: E402

from deeplearning_mpi_b200.parallel import dist  # noqa: E402

# Environment variables set by torchrun
try:
    LOCAL_RANK = int(os.environ["LOCAL_RANK"])
    WORLD_SIZE = int(os.environ["WORLD_SIZE"])
    WORLD_RANK = int(os.environ["RANK"])
except KeyError:
    raise KeyError("Please set correct environment variables (launch with torchrun)")


def run(backend: str) -> None:
    tensor = torch.zeros(1)
    if backend != "gloo":  # device tensors for the NVLink / NCCL path
        tensor = tensor.to(torch.device(f"cuda:{LOCAL_RANK}"))
    if WORLD_RANK == 0:
        tensor += 42.0
        for rank_recv in range(1, WORLD_SIZE):
            dist.send(tensor=tensor, dst=rank_recv)
            print("worker_{} sent data to Rank {}\n".format(0, rank_recv))
    else:
        dist.recv(tensor=tensor, src=0)
        print("worker_{} has received data from rank {}\n".format(WORLD_RANK, 0))
        assert float(tensor.item()) == 42.0, "payload corrupted"


def init_processes(backend: str) -> None:
    dist.init_process_group(backend, rank=WORLD_RANK, world_size=WORLD_SIZE)
    try:
        run(backend)
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument("--backend", type=str, default="nccl", choices=["nccl", "gloo", "p2p"])
    args = parser.parse_args()
    init_processes(backend=args.backend)
