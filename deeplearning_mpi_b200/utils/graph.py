"""Whole-iteration CUDA-graph capture.

For small models (the reference's actual ResNet-18 / CIFAR-10 / batch-128 configuration) a training step is a few hundred
kernels of a few microseconds each, so the step time is set by CPU launch latency, not by the GPU.  Every kernel of this
framework is launched on the current stream with arguments that only depend on tensor addresses, the per-bucket
all-reduce kernels join the capture through the reducer's event edges, and the flag protocol keeps its epochs in
device memory — so forward + backward + gradient all-reduce + optimizer can be captured once and replayed with a
single launch per iteration.

    step = GraphedTrainStep(train_step, (images, labels))      # train_step(images, labels) -> loss tensor
    for images, labels in loader:
        loss = step(images, labels)                             # copies the batch into the static inputs, replays

Rules: fixed input shapes; the callable must be the complete iteration (zero_grad .. optimizer.step) built from this
framework's modules/optimizers; hyper-parameters that are kernel arguments (learning rate, momentum) are frozen at
capture time — call :meth:`recapture` after changing them.  Every rank of a DDP job must capture and replay the same
sequence.
"""
from __future__ import annotations

from typing import Callable, Sequence

import torch

from .. import _native as N
from ..ops import kernels as K


class GraphedTrainStep:
    def __init__(self, step_fn: Callable[..., torch.Tensor], example_inputs: Sequence[torch.Tensor], warmup: int = 3):
        self.step_fn = step_fn
        self.static_inputs = [t.detach().clone() for t in example_inputs]
        self.warmup = warmup
        self.graph = None
        self.static_out = None
        self.launches_per_replay = 0
        self._capture()

    def _capture(self) -> None:
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):  # eager warm-up on a side stream (cuFuncSetAttribute, plan creation, first-step paths)
            for _ in range(self.warmup):
                self.step_fn(*self.static_inputs)
        cur.wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        before = N.launches()
        with torch.cuda.graph(self.graph):
            self.static_out = self.step_fn(*self.static_inputs)
        self.launches_per_replay = N.launches() - before

    def recapture(self) -> None:
        self.graph = None
        self._capture()

    def __call__(self, *inputs: torch.Tensor) -> torch.Tensor:
        if any(tuple(d.shape) != tuple(s.shape) or d.dtype != s.dtype for d, s in zip(self.static_inputs, inputs)):
            # e.g. the short last batch of an epoch: run it eagerly (same kernels, launched one by one)
            return self.step_fn(*inputs)
        for dst, src in zip(self.static_inputs, inputs):
            if dst.data_ptr() != src.data_ptr():
                K.copy_(dst, src)
        self.graph.replay()
        N.count_launch(self.launches_per_replay)
        return self.static_out


class LazyGraphedStep:
    """``GraphedTrainStep`` that captures at its first call, from that call's inputs — for training loops that only see
    tensors once the data loader runs (examples/resnet/main.py --graph, examples/unet/train.py --graph)."""

    def __init__(self, step_fn: Callable[..., torch.Tensor], warmup: int = 3):
        self.step_fn, self.warmup, self._graphed = step_fn, warmup, None

    def __call__(self, *inputs: torch.Tensor) -> torch.Tensor:
        if self._graphed is None:
            # the warm-up steps inside the capture helper are real optimizer steps on this batch
            self._graphed = GraphedTrainStep(self.step_fn, inputs, warmup=self.warmup)
        return self._graphed(*inputs)
