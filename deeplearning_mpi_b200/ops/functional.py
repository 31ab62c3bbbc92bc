"""Autograd functions that bind the sm_100a kernels into PyTorch's autograd graph.

Activations between these functions are NHWC bf16.  Weight gradients are NOT returned to autograd: the
wgrad kernels accumulate straight into the parameter's slot of the flat gradient arena (= the DDP bucket),
after which ``arena.grad_ready(param)`` notifies the reducer — the "wgrad epilogue writes into the bucket
and bumps the ready counter" design of SURVEY.md K3 / N8.  Parameters are still passed as inputs so that
autograd knows the function needs a backward pass.
"""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.nn as nn

from . import conv as CV
from . import kernels as K


_RELU_MASK = os.environ.get("DMPI_NOMASK", "0") != "1"  # A/B switch for the ReLU bit mask of residual blocks


def _nb(t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    return t if t is None else t.detach()


# ------------------------------------------------------------------------------------------------------------------
# Weight gradients on a side stream.  wgrad(l) only feeds the optimizer / the gradient bucket, so it does not have to
# sit on the critical path dy(l) -> dgrad(l) -> BatchNorm backward(l-1) -> ...  It is launched on a second stream and
# co-runs with the HBM-bound BatchNorm backward kernels of the next layer (the conv kernels are tensor / L2 bound, so
# the two use different resources).  The parameter is reported ready (and the main stream made to wait for the side
# stream) one layer later; the last layer of the backward pass runs in order and joins everything.
# Measured (ResNet-50, 1 GPU, same box A/B): 12.72k / 12.89k img/s with the side stream vs 12.72k / 12.78k without —
# the persistent conv kernels hold an SM's whole shared memory, so little actually co-runs; kept as an opt-in switch.
_WGRAD_STREAM = os.environ.get("DMPI_WGRAD_STREAM", "0") == "1"
_side_streams = {}
_pending_wgrads = []  # (arena, weight, event) in launch order


def _side_stream(device) -> "torch.cuda.Stream":
    s = _side_streams.get(device)
    if s is None:
        s = torch.cuda.Stream(device=device)
        _side_streams[device] = s
    return s


def _flush_pending_wgrads(keep_last: int = 0) -> None:
    """Main stream waits for the parked weight gradients (all but the newest ``keep_last``) and reports them ready."""
    while len(_pending_wgrads) > keep_last:
        arena, weight, ev = _pending_wgrads.pop(0)
        torch.cuda.current_stream().wait_event(ev)
        arena.grad_ready(weight)


class GradLink:
    """Side channel between the two branches of a residual fork.

    Where a tensor feeds both the main branch (conv1 ...) and the identity branch of a block, autograd would sum the two
    branch gradients with a separate elementwise kernel (3 tensor passes).  Instead the branch that finishes first in
    the backward pass (the ``give_*`` role) parks its gradient here and returns None to autograd, and the dgrad of the
    main branch's first convolution (the ``take`` role) adds it inside its epilogue.  The giver always runs first:
    it is reachable from the block output without passing through the taker, and it was created later in the forward
    pass, so the autograd engine schedules it earlier."""

    __slots__ = ("grad",)

    def __init__(self):
        self.grad = None

    def give(self, g) -> None:  # a gradient tensor, or a callable that accumulates one onto the taker's result
        if self.grad is not None:
            raise RuntimeError("GradLink: a parked gradient was never consumed")
        self.grad = g

    def take(self):
        g = self.grad
        if g is None:
            raise RuntimeError("GradLink: the residual-branch gradient was not produced before the main branch's dgrad")
        self.grad = None
        return g


class ConvBnAct(torch.autograd.Function):
    """y = Conv2d(x) [+bias];  out = ReLU?(BatchNorm(y) [+ residual]).

    Replaces the Conv2d -> BatchNorm2d -> ReLU(inplace) chains of ref:pytorch/unet/model.py:8-14 and the
    conv/bn/relu/+identity patterns of torchvision BasicBlock/Bottleneck.
    """

    @staticmethod
    def forward(ctx, x, weight, bias, gamma, beta, residual, conv: nn.Conv2d, bn: nn.BatchNorm2d, arena, relu: bool,
                out_holder, stem_in_hw, link=None, link_role=None):
        # link / link_role: see GradLink.  "take": this conv's dgrad adds link.grad;  "give_dx": dx goes to the link
        # instead of to autograd;  "give_dres": the residual input's gradient goes to the link.
        # out_holder: None or a 1-element list holding a pre-allocated NHWC view (e.g. a channel slice of a UNet
        # concat buffer).  It is wrapped in a list so that autograd does not treat it as a differentiable input.
        out = out_holder[0] if out_holder else None
        R, S = conv.kernel_size
        stride, pad = conv.stride[0], conv.padding[0]
        training = bn.training
        stats = arena.bn_stats(bn) if training else None
        is_stem = stem_in_hw is not None
        if is_stem:
            oh, ow, _, _ = CV.smallc_geometry(stem_in_hw[0], stem_in_hw[1], R, S, stride, pad)
            y = CV.smallc_fprop(x, arena.stem_w[id(weight)], R, stride, (oh, ow), bias=_nb(bias), stats=stats)
        else:
            y = CV.conv2d_fprop(x, arena.w_natural(weight), R, S, stride, pad, bias=_nb(bias), stats=stats)
        if out is None:
            out = torch.empty_like(y)
        c = y.shape[-1]
        mean = torch.empty(c, dtype=torch.float32, device=y.device) if training else None
        invstd = torch.empty(c, dtype=torch.float32, device=y.device) if training else None
        track = bn.track_running_stats and bn.running_mean is not None
        momentum = bn.momentum if bn.momentum is not None else 0.1
        # residual + ReLU: the backward gate cannot be recomputed from y alone; one bit per element is 1/16 of the
        # bytes of re-reading `out`
        mask = None
        # (grad mode is off inside Function.forward: ask the context whether a backward pass can follow)
        if training and relu and residual is not None and any(ctx.needs_input_grad) and _RELU_MASK:
            mask = torch.empty((y.numel() // 8,), dtype=torch.uint8, device=y.device)
        K.bn_forward(y, residual, out, stats[0] if training else None, stats[1] if training else None, _nb(gamma),
                     _nb(beta), mean, invstd, bn.running_mean if track else None, bn.running_var if track else None,
                     bn.num_batches_tracked if (track and training) else None, bn.eps, momentum, training, relu,
                     relu_mask=mask)
        ctx.relu_mask = mask
        ctx.conv, ctx.bn, ctx.arena, ctx.relu, ctx.is_stem = conv, bn, arena, relu, is_stem
        ctx.has_res = residual is not None
        ctx.link, ctx.link_role = link, link_role
        ctx.params = (weight, bias, gamma, beta)
        ctx.in_hw = (x.shape[1], x.shape[2]) if not is_stem else stem_in_hw
        ctx.save_for_backward(x, y, out, mean, invstd)
        return out

    @staticmethod
    def backward(ctx, dout):
        x, y, out, mean, invstd = ctx.saved_tensors
        if mean is None:
            # the fused backward implements the batch-statistics formula only (it needs the forward's mean / invstd)
            raise RuntimeError("backward through an eval-mode BatchNorm is not supported by the fused kernels: call "
                               "model.train() for training steps (evaluation runs under torch.no_grad() in the "
                               "reference, ref:pytorch/resnet/main.py:62)")
        weight, bias, gamma, beta = ctx.params
        conv, bn, arena = ctx.conv, ctx.bn, ctx.arena
        R, S = conv.kernel_size
        stride, pad = conv.stride[0], conv.padding[0]
        if dout.stride(-1) != 1 or dout.dtype != torch.bfloat16:
            dout = dout.to(torch.bfloat16).contiguous()
        dy = torch.empty_like(y)
        dres = torch.empty_like(y) if ctx.has_res else None
        # Without a residual the ReLU gate is a function of y alone: recompute it instead of reading `out` again.
        from_y = ctx.relu and not ctx.has_res
        mask = ctx.relu_mask
        K.bn_backward(dout, y, out if (ctx.relu and not from_y and mask is None) else None, dy, dres, mean, invstd,
                      gamma.detach(), arena.bn_acc(bn), arena.grad_physical(gamma), arena.grad_physical(beta),
                      beta=beta.detach(), relu_from_y=from_y, relu_mask=mask)
        arena.grad_ready(gamma)
        arena.grad_ready(beta)
        if bias is not None:
            # d(bias) = sum(dy) is identically zero behind a training-mode BatchNorm; the slot stays as zeroed.
            arena.grad_physical(bias)
            arena.grad_ready(bias)
        dx = None
        if ctx.is_stem:
            _flush_pending_wgrads()
            dwin = arena.stem_dw[id(weight)]
            K.zero_(dwin)
            CV.smallc_wgrad(dy, x, dwin, R, stride)
            k, c = weight.shape[0], weight.shape[1]
            K.stem_unpack_grad(dwin, arena.grad_physical(weight), k, R, S, c)
        else:
            overlap = _WGRAD_STREAM and ctx.needs_input_grad[0]
            if overlap:
                # everything launched so far (the previous layer's wgrad included) stays ahead of this layer's dgrad
                _flush_pending_wgrads(keep_last=0)
                main = torch.cuda.current_stream()
                side = _side_stream(dy.device)
                side.wait_stream(main)
                with torch.cuda.stream(side):
                    CV.conv2d_wgrad(dy, x, arena.grad_physical(weight), R, S, stride, pad)
                    ev = side.record_event()
                dy.record_stream(side)
                x.record_stream(side)
                if not _pending_wgrads:
                    # safety net for graphs whose first layer also needs an input gradient (nothing "last" to join at)
                    torch.autograd.Variable._execution_engine.queue_callback(_flush_pending_wgrads)
                _pending_wgrads.append((arena, weight, ev))
            else:
                _flush_pending_wgrads()  # last layer of the backward pass: join the side stream, then run in order
                CV.conv2d_wgrad(dy, x, arena.grad_physical(weight), R, S, stride, pad)
            if ctx.needs_input_grad[0]:
                parked = ctx.link.take() if ctx.link_role == "take" else None
                if ctx.link_role == "give_dx" and stride == 2:
                    # strided downsample branch: park the dgrad itself; the main branch's dgrad runs it afterwards,
                    # accumulating onto its own (dense) result — no zero-filled sparse tensor, no add kernel
                    wt = arena.w_transposed(weight)
                    in_hw = ctx.in_hw
                    ctx.link.give(lambda dx_main: CV.conv2d_dgrad(dy, wt, R, S, stride, pad, in_hw, out=dx_main,
                                                                    accumulate=True))
                elif callable(parked):
                    dx = CV.conv2d_dgrad(dy, arena.w_transposed(weight), R, S, stride, pad, ctx.in_hw)
                    dx = parked(dx)
                else:
                    dx = CV.conv2d_dgrad(dy, arena.w_transposed(weight), R, S, stride, pad, ctx.in_hw, addend=parked)
        if not (_WGRAD_STREAM and ctx.needs_input_grad[0] and not ctx.is_stem):
            arena.grad_ready(weight)
        if ctx.link_role == "give_dx" and dx is not None:
            ctx.link.give(dx)
            dx = None
        elif ctx.link_role == "give_dres" and dres is not None:
            ctx.link.give(dres)
            dres = None
        return dx, None, None, None, None, dres, None, None, None, None, None, None, None, None


class ConvTranspose2x2(torch.autograd.Function):
    """ConvTranspose2d(k=2, s=2) writing into a channel slice of a pre-allocated NHWC buffer (zero-copy concat,
    ref:pytorch/unet/model.py:36-38,46-47)."""

    @staticmethod
    def forward(ctx, x, weight, bias, out_holder, arena):
        k = weight.shape[1]
        out = out_holder[0]
        CV.convt2x2_fwd(x, arena.w_transposed(weight), out, _nb(bias))
        ctx.arena, ctx.params = arena, (weight, bias)
        ctx.save_for_backward(x)
        ctx.k = k
        return out

    @staticmethod
    def backward(ctx, dout):
        (x,) = ctx.saved_tensors
        weight, bias = ctx.params
        arena = ctx.arena
        if dout.dtype != torch.bfloat16:
            dout = dout.to(torch.bfloat16)
        CV.convt2x2_wgrad(dout, x, arena.grad_physical(weight))
        arena.grad_ready(weight)
        if bias is not None:
            # every output pixel sees the bias once: its gradient is the per-channel sum of dout, added into the slot
            K.channel_stats(dout, arena.grad_physical(bias), None)
            arena.grad_ready(bias)
        dx = CV.convt2x2_dgrad(dout, arena.w_natural(weight)) if ctx.needs_input_grad[0] else None
        return dx, None, None, None, None


class UpsampleBilinear2x(torch.autograd.Function):
    """nn.Upsample(scale_factor=2, mode="bilinear", align_corners=True) on NHWC bf16, writing into a channel slice of
    the decoder's concat buffer (ref:pytorch/unet/model.py:39-40,46-47)."""

    @staticmethod
    def forward(ctx, x, out_holder):
        return K.upsample2x_bilinear_forward(x, out_holder[0] if out_holder else None)

    @staticmethod
    def backward(ctx, dout):
        if dout.dtype != torch.bfloat16:
            dout = dout.to(torch.bfloat16)
        return K.upsample2x_bilinear_backward(dout), None


class ConcatView(torch.autograd.Function):
    """Zero-copy torch.cat([up, skip], dim=channels): both producers already wrote their channel slice of ``buf``;
    this node only ties them into the autograd graph and splits the incoming gradient into two slice views."""

    @staticmethod
    def forward(ctx, up, skip, buf_holder, link=None):
        # link (GradLink, optional): the skip tensor also feeds a MaxPool; its gradient slice is parked in the link
        # and added inside the pooling backward kernel instead of by an autograd add over a strided view
        ctx.c_up = up.shape[-1]
        ctx.link = link
        return buf_holder[0].view(buf_holder[0].shape)

    @staticmethod
    def backward(ctx, g):
        g_skip = g[..., ctx.c_up:]
        if ctx.link is not None:
            ctx.link.give(g_skip)
            g_skip = None
        return g[..., : ctx.c_up], g_skip, None, None


class MaxPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, s, p, link=None):
        y, idx = K.maxpool_forward(x, k, s, p, want_idx=ctx.needs_input_grad[0])
        ctx.cfg = (tuple(x.shape), k, s, p)
        ctx.link = link
        ctx.save_for_backward(idx)
        return y

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        shape, k, s, p = ctx.cfg
        if dy.stride(-1) != 1:
            dy = dy.contiguous()
        addend = ctx.link.take() if ctx.link is not None else None
        if addend is not None and addend.dtype != torch.bfloat16:
            addend = addend.to(torch.bfloat16)
        return K.maxpool_backward(dy, idx, shape, k, s, p, addend=addend), None, None, None, None


class GlobalAvgPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        ctx.shape = tuple(x.shape)
        return K.avgpool_forward(x.contiguous())

    @staticmethod
    def backward(ctx, dy):
        return K.avgpool_backward(dy, ctx.shape)


class LinearFn(torch.autograd.Function):
    """bf16 [M, K] x fp32-master Linear -> fp32 logits [M, N] (nn.Linear of ref:pytorch/resnet/main.py:41)."""

    @staticmethod
    def forward(ctx, x, weight, bias, arena):
        y = CV.linear_fwd(x.contiguous(), arena.w_natural(weight), _nb(bias))
        ctx.arena, ctx.params = arena, (weight, bias)
        ctx.save_for_backward(x)
        return y

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        weight, bias = ctx.params
        arena = ctx.arena
        n_out, kk = weight.shape
        wt = arena.w_transposed(weight)  # [K, N_pad]
        npad = wt.shape[1]
        if dy.dtype != torch.float32 or dy.stride(1) != 1:
            dy = dy.float().contiguous()
        # one pass: bf16 zero-padded copy for the GEMMs + the bias gradient (column sums) into its arena slot
        dyp = K.linear_bwd_prep(dy, npad, arena.grad_physical(bias) if bias is not None else None)
        CV.linear_wgrad(dyp, x.contiguous(), arena.grad_physical(weight))
        arena.grad_ready(weight)
        if bias is not None:
            arena.grad_ready(bias)
        dx = CV.linear_dgrad(dyp, wt) if ctx.needs_input_grad[0] else None
        return dx, None, None, None


class Head1x1(torch.autograd.Function):
    """Conv2d(C, out_classes<=8, 1) on NHWC bf16 -> fp32 NCHW logits (ref:pytorch/unet/model.py:68)."""

    @staticmethod
    def forward(ctx, x, weight, bias, arena):
        n, h, w, c = x.shape
        kc = weight.shape[0]
        y = K.head_forward(x, weight.detach().reshape(kc, c), _nb(bias))  # [n*h*w, kc]
        ctx.arena, ctx.params = arena, (weight, bias)
        ctx.save_for_backward(x)
        if kc == 1:
            return y.view(n, 1, h, w)
        return y.view(n, h, w, kc).permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        weight, bias = ctx.params
        arena = ctx.arena
        n, h, w, c = x.shape
        kc = weight.shape[0]
        d2 = dy.permute(0, 2, 3, 1).reshape(n * h * w, kc).contiguous().float()
        dx = K.head_backward(d2, x, weight.detach().reshape(kc, c), arena.grad_physical(weight).view(kc, c),
                             arena.grad_physical(bias) if bias is not None else None)
        arena.grad_ready(weight)
        if bias is not None:
            arena.grad_ready(bias)
        return dx, None, None, None


class StageInput(torch.autograd.Function):
    """NCHW fp32 batch -> zero-padded NHWC8 bf16 staging buffer for the RGB stem (no gradient)."""

    @staticmethod
    def forward(ctx, x, pad, hp, wp):
        return K.nchw_to_nhwc8_padded(x.contiguous().float(), pad, hp, wp)

    @staticmethod
    def backward(ctx, g):
        return None, None, None, None


class SoftmaxCrossEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels):
        loss, d = K.softmax_xent(logits.contiguous(), labels)
        ctx.save_for_backward(d)
        return loss

    @staticmethod
    def backward(ctx, g):
        (d,) = ctx.saved_tensors
        return K.scale_dev(d, g), None


class BCEWithLogits(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, targets):
        shape = logits.shape
        loss, d = K.bce_logits(logits.contiguous().view(-1), targets.contiguous().view(-1).float())
        ctx.save_for_backward(d)
        ctx.shape = shape
        return loss

    @staticmethod
    def backward(ctx, g):
        (d,) = ctx.saved_tensors
        return K.scale_dev(d, g).view(ctx.shape), None
