"""Rank-0 checkpoints in the reference's format: ``torch.save(ddp_model.state_dict(), path)``
with ``module.``-prefixed keys, fp32, torch-native layouts; loaded after wrapping with a cuda:0 -> cuda:LOCAL_RANK
map_location (ref:pytorch/resnet/main.py:48-52,139; ref:pytorch/unet/train.py:72-74,216,231).  Optionally also the
optimizer / epoch (the reference saves weights only)."""
from __future__ import annotations

import os
from typing import Optional

import torch


def save_checkpoint(model, path: str, optimizer=None, epoch: Optional[int] = None) -> None:
    """Model-only by default (drop-in for the reference).  With ``optimizer``/``epoch`` a second file ``path + '.train'``
    carries the training state so the main file stays loadable by the reference scripts."""
    d = os.path.dirname(path)
    if d and not os.path.isdir(d):
        raise OSError(f"The model directory '{d}' does not exist. Please create it before running the script.")
    tmp = path + ".tmp"
    torch.save(model.state_dict(), tmp)
    os.replace(tmp, path)  # never leave a half-written checkpoint behind
    if optimizer is not None or epoch is not None:
        torch.save({"optimizer": optimizer.state_dict() if optimizer is not None else None, "epoch": epoch}, path + ".train")


def load_checkpoint(model, path: str, local_rank: int = 0, optimizer=None) -> Optional[int]:
    map_location = {"cuda:0": f"cuda:{local_rank}"} if torch.cuda.is_available() else "cpu"
    state = torch.load(path, map_location=map_location)
    own = model.state_dict()
    # accept checkpoints written with or without the DDP 'module.' prefix
    if not any(k in own for k in state) and all(k.startswith("module.") for k in state):
        state = {k[len("module."):]: v for k, v in state.items()}
    elif not any(k in own for k in state) and all(("module." + k) in own for k in state):
        state = {"module." + k: v for k, v in state.items()}
    model.load_state_dict(state)
    epoch = None
    extra = path + ".train"
    if optimizer is not None and os.path.exists(extra):
        t = torch.load(extra, map_location=map_location)
        if t.get("optimizer") is not None:
            optimizer.load_state_dict(t["optimizer"])
        epoch = t.get("epoch")
    return epoch
