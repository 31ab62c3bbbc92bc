"""Thin, allocation-explicit Python wrappers over the C ABI of the non-GEMM sm_100a kernels
(BatchNorm, pooling, losses, optimizers, layout).  No math happens here."""
from __future__ import annotations

from typing import Optional, Tuple

import os

import torch

from .. import _native as N


def _p(t: Optional[torch.Tensor]):
    return t.data_ptr() if t is not None else None


def _chk(rc: int, name: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{name} failed with code {rc}")


def _rows_ld(x: torch.Tensor) -> Tuple[int, int]:
    """[..., C] tensor with uniform pixel stride -> (rows, row stride)."""
    assert x.stride(-1) == 1
    c = x.shape[-1]
    rows = x.numel() // c
    if x.dim() == 2:
        return rows, x.stride(0)
    assert x.dim() == 4
    n, h, w, _ = x.shape
    sn, sh, sw, _ = x.stride()
    assert sh == w * sw and (n == 1 or sn == h * sh), "tensor must have a uniform pixel stride"
    return rows, sw


def zero_(t: torch.Tensor) -> torch.Tensor:
    """In-place clear of a dense tensor with the store kernel behind dmpi_memset_async (CPU tensors: Tensor.zero_)."""
    if not t.is_cuda:
        return t.zero_()
    assert t.is_contiguous()
    rc = N.cuda().dmpi_memset_async(_p(t), 0, t.numel() * t.element_size(), N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_memset_async")
    return t


def copy_(dst: torch.Tensor, src: torch.Tensor) -> torch.Tensor:
    """dst <- src on the current stream with the framework's own copy kernel when both are dense, same dtype and a
    multiple of 16 bytes (a memcpy NODE in a replayed CUDA graph costs several us of latency); Tensor.copy_ otherwise."""
    nbytes = src.numel() * src.element_size()
    if (dst.is_cuda and src.is_cuda and dst.dtype == src.dtype and dst.is_contiguous() and src.is_contiguous()
            and dst.shape == src.shape and nbytes % 16 == 0 and nbytes > 0 and dst.data_ptr() % 16 == 0
            and src.data_ptr() % 16 == 0 and dst.device == src.device):
        blocks = max(1, min(148 * 4, nbytes // (512 * 16 * 4)))
        rc = N.cuda().dmpi_peer_copy(_p(dst), _p(src), nbytes, blocks, N.stream_ptr())
        N.count_launch()
        _chk(rc, "dmpi_peer_copy")
        return dst
    return dst.copy_(src, non_blocking=True)


# ------------------------------------------------------------------------------------------ BatchNorm
_DETERMINISTIC = os.environ.get("DMPI_DETERMINISTIC_BN", "0" if os.environ.get("DMPI_DETERMINISTIC", "1") == "0" else "1") == "1"
_BN_DET_BYTES = 1 << 20
_bn_det = {}


def _bn_det_scratch(device) -> torch.Tensor:
    """Zero-initialised scratch of the deterministic BatchNorm backward reduction, one per (device, stream)."""
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    t = _bn_det.get(key)
    if t is None:
        t = torch.zeros(_BN_DET_BYTES // 4, dtype=torch.float32, device=device)
        _bn_det[key] = t
    return t


def bn_forward(y, res, out, ssum, ssq, gamma, beta, save_mean, save_invstd, running_mean, running_var, nbt,
               eps: float, momentum: float, training: bool, relu: bool, relu_mask=None) -> None:
    """``relu_mask`` (uint8 [M, C/8], optional) receives one bit per output element (value > 0) for the backward pass."""
    M, ldy = _rows_ld(y)
    _, ldo = _rows_ld(out)
    ldr = _rows_ld(res)[1] if res is not None else 0
    C = y.shape[-1]
    rc = N.cuda().dmpi_bn_forward_mask(_p(y), _p(res), _p(out), M, C, ldy, ldr, ldo, _p(ssum), _p(ssq), _p(gamma),
                                       _p(beta), _p(save_mean), _p(save_invstd), _p(running_mean), _p(running_var),
                                       _p(nbt), float(eps), float(momentum), int(training), int(relu), _p(relu_mask),
                                       N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_bn_forward")


def bn_backward(dout, y, act, dy, dres, mean, invstd, gamma, acc, dgamma, dbeta, beta=None, relu_from_y=False,
                relu_mask=None) -> None:
    """ReLU gate source, in order of preference: ``relu_mask`` (bits written by bn_forward), ``relu_from_y`` (recomputed
    from y, exact without a residual), ``act`` (the forward output)."""
    M, ld_dout = _rows_ld(dout)
    _, ldy = _rows_ld(y)
    ld_act = _rows_ld(act)[1] if act is not None else 0
    _, ld_dy = _rows_ld(dy)
    ld_dres = _rows_ld(dres)[1] if dres is not None else 0
    C = y.shape[-1]
    scratch = _bn_det_scratch(dout.device) if _DETERMINISTIC else None
    rc = N.cuda().dmpi_bn_backward_det(_p(dout), _p(y), _p(act), _p(dy), _p(dres), M, C, ld_dout, ldy, ld_act, ld_dy,
                                       ld_dres, _p(mean), _p(invstd), _p(gamma), _p(beta), int(relu_from_y),
                                       _p(relu_mask), _p(acc), _p(dgamma), _p(dbeta), _p(scratch),
                                       _BN_DET_BYTES if scratch is not None else 0, N.stream_ptr())
    N.count_launch(2)
    _chk(rc, "dmpi_bn_backward")


def channel_stats(y, ssum, ssq) -> None:
    M, ld = _rows_ld(y)
    rc = N.cuda().dmpi_bn_stats(_p(y), M, y.shape[-1], ld, _p(ssum), _p(ssq), N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_bn_stats")


# ------------------------------------------------------------------------------------------ pooling
def maxpool_forward(x, k: int, s: int, p: int, want_idx: bool):
    n, h, w, c = x.shape
    oh = (h + 2 * p - k) // s + 1
    ow = (w + 2 * p - k) // s + 1
    y = torch.empty((n, oh, ow, c), dtype=torch.bfloat16, device=x.device)
    idx = torch.empty((n, oh, ow, c), dtype=torch.uint8, device=x.device) if want_idx else None
    rc = N.cuda().dmpi_maxpool_forward(_p(x), _p(y), _p(idx), n, h, w, c, oh, ow, k, s, p, x.stride(0), x.stride(1),
                                       x.stride(2), y.stride(0), y.stride(1), y.stride(2), N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_maxpool_forward")
    return y, idx


def maxpool_backward(dy, idx, in_shape, k: int, s: int, p: int, addend=None):
    """dx of MaxPool2d; ``addend`` (bf16 NHWC, any n/h/w strides) is a second gradient of the same tensor that is added
    in the same pass (the UNet skip connection)."""
    n, h, w, c = in_shape
    _, oh, ow, _ = dy.shape
    dx = torch.empty((n, h, w, c), dtype=torch.bfloat16, device=dy.device)
    if addend is None:
        rc = N.cuda().dmpi_maxpool_backward(_p(dy), _p(idx), _p(dx), n, h, w, c, oh, ow, k, s, p, dy.stride(0),
                                            dy.stride(1), dy.stride(2), dx.stride(0), dx.stride(1), dx.stride(2),
                                            N.stream_ptr())
    else:
        assert tuple(addend.shape) == (n, h, w, c) and addend.dtype == torch.bfloat16 and addend.stride(3) == 1
        rc = N.cuda().dmpi_maxpool_backward_add(_p(dy), _p(idx), _p(dx), n, h, w, c, oh, ow, k, s, p, dy.stride(0),
                                                dy.stride(1), dy.stride(2), dx.stride(0), dx.stride(1), dx.stride(2),
                                                _p(addend), addend.stride(0), addend.stride(1), addend.stride(2),
                                                N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_maxpool_backward")
    return dx


def avgpool_forward(x):
    n, h, w, c = x.shape
    assert x.is_contiguous()
    y = torch.empty((n, c), dtype=torch.bfloat16, device=x.device)
    rc = N.cuda().dmpi_avgpool_forward(_p(x), _p(y), n, h * w, c, N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_avgpool_forward")
    return y


def avgpool_backward(dy, in_shape):
    n, h, w, c = in_shape
    dx = torch.empty((n, h, w, c), dtype=torch.bfloat16, device=dy.device)
    rc = N.cuda().dmpi_avgpool_backward(_p(dy.contiguous()), _p(dx), n, h * w, c, N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_avgpool_backward")
    return dx


def upsample2x_bilinear_forward(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """NHWC bf16 [n, h, w, c] -> [n, 2h, 2w, c] (align_corners=True); ``out`` may be a channel slice of a wider buffer."""
    n, h, w, c = x.shape
    assert x.dtype == torch.bfloat16 and x.stride(3) == 1
    if out is None:
        out = torch.empty((n, 2 * h, 2 * w, c), dtype=torch.bfloat16, device=x.device)
    assert tuple(out.shape) == (n, 2 * h, 2 * w, c) and out.stride(3) == 1
    rc = N.cuda().dmpi_upsample2x_bilinear_forward(_p(x), _p(out), n, h, w, c, x.stride(0), x.stride(1), x.stride(2),
                                                   out.stride(0), out.stride(1), out.stride(2), N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_upsample2x_bilinear_forward")
    return out


def upsample2x_bilinear_backward(dy: torch.Tensor) -> torch.Tensor:
    n, oh, ow, c = dy.shape
    assert dy.dtype == torch.bfloat16 and dy.stride(3) == 1 and oh % 2 == 0 and ow % 2 == 0
    dx = torch.empty((n, oh // 2, ow // 2, c), dtype=torch.bfloat16, device=dy.device)
    rc = N.cuda().dmpi_upsample2x_bilinear_backward(_p(dy), _p(dx), n, oh // 2, ow // 2, c, dy.stride(0), dy.stride(1),
                                                    dy.stride(2), dx.stride(0), dx.stride(1), dx.stride(2),
                                                    N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_upsample2x_bilinear_backward")
    return dx


# ------------------------------------------------------------------------------------------ losses / head
def softmax_xent(logits: torch.Tensor, labels: torch.Tensor, grad_scale: float = 1.0):
    """Returns (loss scalar fp32 tensor, dlogits fp32 [M, N]) for mean-reduced cross-entropy."""
    m, n = logits.shape
    assert logits.dtype == torch.float32 and logits.stride(1) == 1 and labels.dtype == torch.int64
    loss = zero_(torch.empty((), dtype=torch.float32, device=logits.device))
    d = torch.empty((m, n), dtype=torch.float32, device=logits.device)
    rc = N.cuda().dmpi_softmax_xent(_p(logits), _p(labels), _p(d), _p(loss), m, n, logits.stride(0), n, n,
                                    float(grad_scale), N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_softmax_xent")
    return loss, d


def bce_logits(x: torch.Tensor, t: torch.Tensor, grad_scale: float = 1.0):
    assert x.dtype == torch.float32 and t.dtype == torch.float32 and x.is_contiguous() and t.is_contiguous()
    loss = zero_(torch.empty((), dtype=torch.float32, device=x.device))
    d = torch.empty_like(x)
    rc = N.cuda().dmpi_bce_logits(_p(x), _p(t), _p(d), _p(loss), x.numel(), float(grad_scale), N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_bce_logits")
    return loss, d


def scale_dev(src: torch.Tensor, scale: torch.Tensor) -> torch.Tensor:
    """src * scale for a 0-dim fp32 device tensor ``scale`` (the gradient flowing into a scalar loss), out of place."""
    assert src.dtype == torch.float32 and src.is_contiguous()
    if scale.dtype != torch.float32 or not scale.is_cuda:
        scale = scale.to(device=src.device, dtype=torch.float32)
    dst = torch.empty_like(src)
    rc = N.cuda().dmpi_scale_dev(_p(src), _p(dst), src.numel(), _p(scale), N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_scale_dev")
    return dst


def linear_bwd_prep(dy: torch.Tensor, npad: int, dbias: Optional[torch.Tensor]) -> torch.Tensor:
    """fp32 logit gradient [M, N] -> bf16 [M, npad] (zero padded) for the tcgen05 GEMMs; dbias += column sums."""
    assert dy.dtype == torch.float32 and dy.dim() == 2 and dy.stride(1) == 1
    m, n = dy.shape
    out = torch.empty((m, npad), dtype=torch.bfloat16, device=dy.device)
    rc = N.cuda().dmpi_linear_bwd_prep(_p(dy), dy.stride(0), m, n, _p(out), npad, _p(dbias), N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_linear_bwd_prep")
    return out


def head_forward(x, w, b):
    """x NHWC bf16 [..., C]; w fp32 [K, C]; -> fp32 [rows, K]."""
    M, ldx = _rows_ld(x)
    K, Cc = w.shape
    y = torch.empty((M, K), dtype=torch.float32, device=x.device)
    rc = N.cuda().dmpi_head_forward(_p(x), _p(w), _p(b), _p(y), M, Cc, K, ldx, N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_head_forward")
    return y


def head_backward(dy, x, w, dw, db):
    M, ldx = _rows_ld(x)
    K, Cc = w.shape
    dx = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
    rc = N.cuda().dmpi_head_backward(_p(dy), _p(x), _p(w), _p(dx), _p(dw), _p(db), M, Cc, K, ldx, Cc, N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_head_backward")
    return dx


# ------------------------------------------------------------------------------------------ optimizers
def sgd_step(p, g, mom, shadow, n, lr, momentum, dampening, wd, first_step, sumsq=None, max_norm=0.0) -> None:
    rc = N.cuda().dmpi_sgd_step(_p(p), _p(g), _p(mom), _p(shadow), n, lr, momentum, dampening, wd, int(first_step),
                                _p(sumsq), float(max_norm), N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_sgd_step")


def adam_step(p, g, m, v, shadow, n, lr, b1, b2, eps, wd, bc1, bc2, sumsq=None, max_norm=0.0) -> None:
    rc = N.cuda().dmpi_adam_step(_p(p), _p(g), _p(m), _p(v), _p(shadow), n, lr, b1, b2, eps, wd, bc1, bc2, _p(sumsq),
                                 float(max_norm), N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_adam_step")


def adam_step_dev(p, g, m, v, shadow, n, lr, b1, b2, eps, wd, step_t, sumsq=None, max_norm=0.0) -> None:
    """Adam with the step count read from the device tensor ``step_t`` (fp32 scalar) — CUDA-graph safe."""
    rc = N.cuda().dmpi_adam_step_dev(_p(p), _p(g), _p(m), _p(v), _p(shadow), n, lr, b1, b2, eps, wd, _p(step_t),
                                     _p(sumsq), float(max_norm), N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_adam_step_dev")


def sumsq(g, n, out) -> None:
    rc = N.cuda().dmpi_sumsq(_p(g), n, _p(out), N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_sumsq")


def scalar_add_(t: torch.Tensor, value: float) -> torch.Tensor:
    """t += value for a one-element fp32 device tensor (own kernel, graph-capturable)."""
    rc = N.cuda().dmpi_scalar_op(_p(t), _p(t), 0, float(value), N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_scalar_op")
    return t


def scalar_sqrt(t: torch.Tensor) -> torch.Tensor:
    """sqrt of a one-element fp32 device tensor into a new tensor (own kernel)."""
    out = torch.empty_like(t)
    rc = N.cuda().dmpi_scalar_op(_p(out), _p(t), 1, 0.0, N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_scalar_op")
    return out


def scale_by_clip(g, n, sumsq_t, max_norm) -> None:
    rc = N.cuda().dmpi_scale_by_clip(_p(g), n, _p(sumsq_t), float(max_norm), N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_scale_by_clip")


def cast_bf16(src, dst, n) -> None:
    rc = N.cuda().dmpi_cast_bf16(_p(src), _p(dst), n, N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_cast_bf16")


# ------------------------------------------------------------------------------------------ layout
def nchw_to_nhwc8_padded(x: torch.Tensor, pad: int, hp: int, wp: int, out: Optional[torch.Tensor] = None):
    n, c, h, w = x.shape
    assert x.dtype == torch.float32 and x.is_contiguous() and c <= 8
    if out is None:
        out = torch.empty((n, hp, wp, 8), dtype=torch.bfloat16, device=x.device)
    rc = N.cuda().dmpi_nchw_to_nhwc8_padded(_p(x), _p(out), n, c, h, w, pad, hp, wp, N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_nchw_to_nhwc8_padded")
    return out


def nchw_to_nhwc(x: torch.Tensor):
    n, c, h, w = x.shape
    assert x.dtype == torch.float32 and x.is_contiguous()
    out = torch.empty((n, h, w, c), dtype=torch.bfloat16, device=x.device)
    rc = N.cuda().dmpi_nchw_to_nhwc(_p(x), _p(out), n, c, h * w, N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_nchw_to_nhwc")
    return out


def nhwc_to_nchw(x: torch.Tensor):
    n, h, w, c = x.shape
    _, ld = _rows_ld(x)
    out = torch.empty((n, c, h, w), dtype=torch.float32, device=x.device)
    rc = N.cuda().dmpi_nhwc_to_nchw(_p(x), _p(out), n, c, h * w, ld, N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_nhwc_to_nchw")
    return out


def transpose_table(table_dev: torch.Tensor, n_entries: int, total_tiles: int) -> None:
    rc = N.cuda().dmpi_transpose_table(_p(table_dev), n_entries, total_tiles, N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_transpose_table")


def stem_pack(w_master: torch.Tensor, out: torch.Tensor, K: int, R: int, S: int, Cc: int) -> None:
    rc = N.cuda().dmpi_stem_pack(_p(w_master), _p(out), K, R, S, Cc, N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_stem_pack")


def stem_unpack_grad(dwin: torch.Tensor, dw: torch.Tensor, K: int, R: int, S: int, Cc: int) -> None:
    rc = N.cuda().dmpi_stem_unpack_grad(_p(dwin), _p(dw), K, R, S, Cc, N.stream_ptr())
    N.count_launch()
    _chk(rc, "dmpi_stem_unpack_grad")
