"""Fused optimizers with the constructor signatures the reference uses:

    optim.SGD(ddp_model.parameters(), lr=..., momentum=0.9, weight_decay=1e-5)    ref:pytorch/resnet/main.py:114
    optim.Adam(model.parameters(), lr=...)                                        ref:pytorch/unet/train.py:160-161
    torch.nn.utils.clip_grad_norm_(model.parameters(), max_norm=1.0)              ref:pytorch/unet/train.py:194

When the parameters live in a :class:`~deeplearning_mpi_b200.utils.arena.ParamArena` on a CUDA device the whole
model is updated by ONE kernel over the flat fp32 arena (which also writes the bf16 mirror used by the
tcgen05 kernels); otherwise (CPU / foreign parameters) the same update rule runs through plain PyTorch so the
CPU tests can compare against ``torch.optim`` bit-for-bit semantics.
"""
from __future__ import annotations

from typing import Iterable, List, Optional

import torch

from .ops import kernels as K
from .utils.arena import _ARENAS, ParamArena


def _find_arena(params: List[torch.nn.Parameter]) -> Optional[ParamArena]:
    """The arena whose trainable parameters are exactly ``params`` (frozen ones may be passed along, as torch.optim
    allows: they have no gradient and are skipped), if any."""
    arena = None
    for obj in _ARENAS:
        a = obj()
        if a is None:
            continue
        if not all(id(p) in a.by_param for p in params):
            continue
        n_train = sum(1 for p in params if p.requires_grad)
        if n_train == a.n_trainable and len(params) in (a.n_trainable, len(a.slots)):
            arena = a
    return arena


class _FlatOptimizer:
    def __init__(self, params: Iterable[torch.nn.Parameter], defaults: dict):
        self.params = [p for p in params]
        if not self.params:
            raise ValueError("optimizer got an empty parameter list")
        self.defaults = defaults
        self.param_groups = [dict(defaults, params=self.params)]
        self.state: dict = {}
        self._arena: Optional[ParamArena] = None
        self._clip_sumsq: Optional[torch.Tensor] = None
        self._clip_max_norm = 0.0
        self.step_count = 0

    # -- arena discovery is lazy: the model may be wrapped / moved after the optimizer is built
    def _resolve_arena(self, cuda_only: bool = True) -> Optional[ParamArena]:
        if self._arena is not None:
            s = self._arena.slots[0]
            if s.param.data_ptr() != self._arena.flat_p.data_ptr() + s.offset * 4:
                self._arena = None
        if self._arena is None:
            self._arena = _find_arena(self.params)
        a = self._arena
        if a is not None and cuda_only and not a.is_cuda:
            return None
        return a

    def zero_grad(self, set_to_none: bool = False) -> None:
        a = self._resolve_arena(cuda_only=False)
        if a is not None:
            a.zero_grad()
            return
        for p in self.params:
            if p.grad is not None:
                if set_to_none:
                    p.grad = None
                else:
                    p.grad = p.grad.detach()
                    p.grad.zero_()

    def fuse_clip_grad_norm(self, max_norm: float) -> torch.Tensor:
        """Device-side clip_grad_norm_ folded into the next ``step()``: computes the global L2 norm with one
        reduction kernel and lets the update kernel apply min(1, max_norm/(norm+1e-6)).  Returns the norm (a
        device scalar; no host sync)."""
        a = self._resolve_arena()
        assert a is not None, "fused clipping needs arena-resident CUDA parameters"
        if self._clip_sumsq is None:
            self._clip_sumsq = torch.empty((), dtype=torch.float32, device=a.device)
        K.zero_(self._clip_sumsq)
        K.sumsq(a.flat_g, a.total_trainable, self._clip_sumsq)
        self._clip_max_norm = float(max_norm)
        return K.scalar_sqrt(self._clip_sumsq)

    def _take_clip(self):
        ss, mn = (self._clip_sumsq, self._clip_max_norm) if self._clip_max_norm > 0 else (None, 0.0)
        self._clip_max_norm = 0.0
        return ss, mn

    def state_dict(self) -> dict:
        return {"step": self.step_count, "defaults": dict(self.defaults),
                "state": {k: v.clone() if torch.is_tensor(v) else v for k, v in self.state.items()}}

    def load_state_dict(self, sd: dict) -> None:
        self.step_count = sd["step"]
        for k, v in sd["state"].items():
            if k in self.state and torch.is_tensor(v):
                self.state[k].copy_(v)
            else:
                self.state[k] = v


class SGD(_FlatOptimizer):
    def __init__(self, params, lr: float = 1e-3, momentum: float = 0.0, dampening: float = 0.0,
                 weight_decay: float = 0.0, nesterov: bool = False):
        if nesterov:
            raise NotImplementedError("nesterov momentum is not used by the reference and not implemented")
        super().__init__(params, dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay))

    @torch.no_grad()
    def step(self) -> None:
        g = self.param_groups[0]
        lr, mom, damp, wd = g["lr"], g["momentum"], g["dampening"], g["weight_decay"]
        a = self._resolve_arena()
        first = self.step_count == 0
        if a is not None:
            if mom != 0 and "momentum_buffer" not in self.state:
                self.state["momentum_buffer"] = K.zero_(torch.empty_like(a.flat_p))
            ss, mn = self._take_clip()
            K.sgd_step(a.flat_p, a.flat_g, self.state.get("momentum_buffer"), a.flat_s, a.total_trainable, lr, mom, damp, wd,
                       first, ss, mn)
            a.refresh_derived()
            a.mark_shadows_fresh()
        else:
            for i, p in enumerate(self.params):
                if p.grad is None:
                    continue
                d = p.grad
                if wd != 0:
                    d = d.add(p, alpha=wd)
                if mom != 0:
                    buf = self.state.get(i)
                    if buf is None:
                        buf = self.state[i] = d.clone()
                    else:
                        buf.mul_(mom).add_(d, alpha=1 - damp)
                    d = buf
                p.add_(d, alpha=-lr)
        self.step_count += 1


class Adam(_FlatOptimizer):
    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))

    @torch.no_grad()
    def step(self) -> None:
        g = self.param_groups[0]
        lr, (b1, b2), eps, wd = g["lr"], g["betas"], g["eps"], g["weight_decay"]
        self.step_count += 1
        t = self.step_count
        bc1, bc2 = 1 - b1 ** t, 1 - b2 ** t
        a = self._resolve_arena()
        if a is not None:
            if "exp_avg" not in self.state:
                self.state["exp_avg"] = K.zero_(torch.empty_like(a.flat_p))
                self.state["exp_avg_sq"] = K.zero_(torch.empty_like(a.flat_p))
                # the step count also lives on the device so that a captured CUDA graph of the step stays correct
                self.state["step_t"] = torch.full((), float(t - 1), dtype=torch.float32, device=a.device)
            K.scalar_add_(self.state["step_t"], 1.0)
            ss, mn = self._take_clip()
            K.adam_step_dev(a.flat_p, a.flat_g, self.state["exp_avg"], self.state["exp_avg_sq"], a.flat_s, a.total_trainable, lr,
                            b1, b2, eps, wd, self.state["step_t"], ss, mn)
            a.refresh_derived()
            a.mark_shadows_fresh()
        else:
            for i, p in enumerate(self.params):
                if p.grad is None:
                    continue
                d = p.grad if wd == 0 else p.grad.add(p, alpha=wd)
                st = self.state.setdefault(i, {"m": torch.zeros_like(p), "v": torch.zeros_like(p)})
                st["m"].mul_(b1).add_(d, alpha=1 - b1)
                st["v"].mul_(b2).addcmul_(d, d, value=1 - b2)
                denom = (st["v"].sqrt() / (bc2 ** 0.5)).add_(eps)
                p.addcdiv_(st["m"], denom, value=-lr / bc1)


def clip_grad_norm_(parameters, max_norm: float, norm_type: float = 2.0) -> torch.Tensor:
    """Drop-in for ``torch.nn.utils.clip_grad_norm_`` (L2 only on the fused path): scales the gradients in place
    and returns the total norm.  Arena-resident CUDA gradients take two kernels over the flat gradient arena."""
    params = [p for p in parameters]
    a = _find_arena(params)
    if a is not None and a.is_cuda and norm_type == 2.0:
        ss = K.zero_(torch.empty((), dtype=torch.float32, device=a.device))
        K.sumsq(a.flat_g, a.total_trainable, ss)
        K.scale_by_clip(a.flat_g, a.total_trainable, ss, max_norm)
        return K.scalar_sqrt(ss)
    return torch.nn.utils.clip_grad_norm_(params, max_norm, norm_type)
