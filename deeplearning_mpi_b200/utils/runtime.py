"""Small runtime helpers shared by the example trainers: env contract, seeding, log files, timers, config."""
from __future__ import annotations

import argparse
import os
import random
import time
from dataclasses import dataclass
from datetime import datetime
from typing import Optional

import numpy as np
import torch


def torchrun_env():
    """(LOCAL_RANK, WORLD_SIZE, RANK) from the torchrun environment; raises like the reference when absent
    (ref:pytorch/resnet/main.py:17-23, ref:pytorch/unet/train.py:20-25)."""
    try:
        return int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"])
    except KeyError as e:
        raise RuntimeError("Missing required environment variables for distributed training (launch with torchrun)") from e


def set_random_seeds(seed: int) -> None:
    """torch / numpy / random seeding (ref:pytorch/resnet/main.py:26-33).  The cuDNN flags the reference toggles are
    accepted for compatibility; cuDNN is not on this framework's hot path."""
    torch.backends.cudnn.deterministic = True
    torch.backends.cudnn.benchmark = True
    torch.manual_seed(seed)
    np.random.seed(seed)
    random.seed(seed)


def create_log_file(log_dir: str = "logs") -> str:
    """logs/training_log_<YYYYMMDD_HHMMSS>.log (ref:pytorch/unet/train.py:44-51)."""
    return os.path.join(log_dir, f"training_log_{datetime.now().strftime('%Y%m%d_%H%M%S')}.log")


def log_to_file(path: Optional[str], message: str) -> None:
    if path:
        with open(path, "a") as f:
            f.write(message + "\n")


def system_information(local_rank: int, world_size: int) -> str:
    name = torch.cuda.get_device_name(local_rank) if torch.cuda.is_available() else "cpu"
    return f"World size: {world_size}, Local rank: {local_rank}, GPU: {name}"


class CudaTimer:
    """Device-side timing with CUDA events on the current stream (primary metric path)."""

    def __init__(self):
        self.e0 = torch.cuda.Event(enable_timing=True)
        self.e1 = torch.cuda.Event(enable_timing=True)

    def __enter__(self):
        self.e0.record()
        return self

    def __exit__(self, *exc):
        self.e1.record()

    def ms(self) -> float:
        self.e1.synchronize()
        return self.e0.elapsed_time(self.e1)


class FaultInjector:
    """Test hook: ``DMPI_FAULT="rank=1,step=2,kind=exit"`` makes rank 1 fail at its third
    synchronising training forward.  kinds: ``exit`` (the process dies without cleanup, like a killed worker),
    ``raise`` (a Python exception, which the trainers let propagate to a non-zero exit code), ``stall=<seconds>``
    (the rank goes silent; peers must run into their collective timeout instead of hanging forever).
    Parsed once per process; inactive (and free) when the variable is unset."""

    def __init__(self, spec: Optional[str] = None):
        spec = os.environ.get("DMPI_FAULT", "") if spec is None else spec
        self.rank = self.step = -1
        self.kind, self.stall_s = "", 0.0
        for item in filter(None, (x.strip() for x in spec.split(","))):
            k, _, v = item.partition("=")
            if k == "rank":
                self.rank = int(v)
            elif k == "step":
                self.step = int(v)
            elif k == "kind":
                self.kind = v
            elif k == "stall":
                self.kind, self.stall_s = "stall", float(v)
            else:
                raise ValueError(f"DMPI_FAULT: unknown field {item!r}")
        if self.kind not in ("", "exit", "raise", "stall"):
            raise ValueError(f"DMPI_FAULT: unknown kind {self.kind!r}")
        self.active = self.kind != "" and self.rank >= 0 and self.step >= 0
        self.count = 0

    def tick(self, rank: int) -> None:
        """Called once per synchronising training forward of the DDP wrapper."""
        if not self.active:
            return
        step, self.count = self.count, self.count + 1
        if rank != self.rank or step != self.step:
            return
        if self.kind == "exit":
            os._exit(17)
        if self.kind == "raise":
            raise RuntimeError(f"DMPI_FAULT: injected failure on rank {rank} at step {step}")
        time.sleep(self.stall_s)


class nvtx_range:
    """NVTX range (shows up in nsys / ncu timelines); no-op without CUDA."""

    def __init__(self, name: str):
        self.name = name

    def __enter__(self):
        if torch.cuda.is_available():
            torch.cuda.nvtx.range_push(self.name)

    def __exit__(self, *exc):
        if torch.cuda.is_available():
            torch.cuda.nvtx.range_pop()


@dataclass
class TrainConfig:
    """One config object behind the reference's flag names."""
    num_epochs: int = 100
    batch_size: int = 128
    learning_rate: float = 0.1
    random_seed: int = 0
    model_dir: str = "saved_models"
    model_filename: str = "model.pth"
    resume: bool = False
    backend: str = "nccl"
    synthetic: bool = False
    steps_per_epoch: int = 0
    graph: bool = False

    @property
    def model_filepath(self) -> str:
        return os.path.join(self.model_dir, self.model_filename)

    @staticmethod
    def parser(defaults: "TrainConfig") -> argparse.ArgumentParser:
        p = argparse.ArgumentParser(formatter_class=argparse.ArgumentDefaultsHelpFormatter)
        p.add_argument("--num_epochs", type=int, default=defaults.num_epochs, help="Number of training epochs.")
        p.add_argument("--batch_size", type=int, default=defaults.batch_size, help="Training batch size for one process.")
        p.add_argument("--learning_rate", type=float, default=defaults.learning_rate, help="Learning rate.")
        p.add_argument("--random_seed", type=int, default=defaults.random_seed, help="Random seed.")
        p.add_argument("--model_dir", type=str, default=defaults.model_dir, help="Directory for saving models.")
        p.add_argument("--model_filename", type=str, default=defaults.model_filename, help="Model filename.")
        p.add_argument("--resume", action="store_true", help="Resume training from saved checkpoint.")
        p.add_argument("--backend", type=str, default=defaults.backend, choices=["nccl", "p2p", "gloo"])
        p.add_argument("--synthetic", action="store_true", help="Use synthetic data of the dataset's shape (no files needed).")
        p.add_argument("--steps_per_epoch", type=int, default=0, help="Cap the steps per epoch (0 = full epoch).")
        p.add_argument("--graph", action="store_true",
                       help="Replay the whole training iteration (forward, backward, gradient all-reduce, optimizer) as "
                            "ONE CUDA graph after a short eager warm-up: the launch-bound small-model regime runs "
                            "3-4x faster (utils/graph.py).  GPU backends only; fixed batch shape (odd batches run eagerly).")
        return p

    @classmethod
    def from_args(cls, ns: argparse.Namespace) -> "TrainConfig":
        fields = {f for f in cls.__dataclass_fields__}
        return cls(**{k: v for k, v in vars(ns).items() if k in fields})
