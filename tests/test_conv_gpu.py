"""Numerics of the tcgen05 implicit-GEMM / wgrad kernels against plain PyTorch fp32 references
(same bf16-rounded inputs, fp32 math)."""
import os

os.environ.setdefault("DMPI_PATCH", "2")  # exercise the patch-mode kernel on every eligible 3x3 case, not only C = 64
os.environ.setdefault("DMPI_WGRAD_PATCH", "2")

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _dev():
    return torch.device("cuda:0")


def _rand(*shape, scale=1.0):
    return (torch.randn(*shape, device=_dev()) * scale).to(torch.bfloat16)


def _close(got, ref, tol, what):
    got = got.float()
    ref = ref.float()
    err = (got - ref).abs().max().item()
    den = ref.abs().max().item() + 1e-6
    print(f"[{what}] max_abs_err={err:.4e} ref_max={den:.4e} rel={err / den:.4e}")
    assert torch.isfinite(got).all(), f"{what}: non-finite output"
    assert err / den < tol, f"{what}: rel err {err / den:.4e} > {tol}"


def _torch_setup():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(0)


def _w2d(w):  # [K, C, R, S] -> [K, R*S*C]
    return w.permute(0, 2, 3, 1).contiguous().view(w.shape[0], -1)


def _wt2d(w):  # [K, C, R, S] -> [C, R*S*K]
    return w.permute(1, 2, 3, 0).contiguous().view(w.shape[1], -1)


CONV_CASES = [
    # n, h, w, c, k, R, stride, pad
    (3, 28, 28, 128, 128, 3, 1, 1),   # patch mode, streamed filter taps, 2 channel blocks
    (1, 128, 128, 64, 64, 3, 1, 1),   # patch mode, row-segment tiles (128 x 1), resident filter
    (1, 56, 40, 64, 256, 3, 1, 1),    # patch mode, non-square image, two N tiles
    (2, 224, 224, 64, 64, 3, 1, 1),   # patch mode, many tiles per CTA
    (4, 14, 14, 256, 512, 1, 1, 0),
    (2, 56, 56, 64, 64, 3, 1, 1),
    (3, 28, 28, 128, 128, 3, 2, 1),
    (2, 28, 28, 256, 512, 1, 2, 0),
    (5, 7, 7, 512, 512, 3, 1, 1),
    (2, 14, 14, 256, 256, 3, 2, 1),
    (1, 32, 32, 192, 64, 3, 1, 1),
]


@pytest.mark.parametrize("case", CONV_CASES)
def test_fprop(case):
    from deeplearning_mpi_b200.ops import conv as C

    _torch_setup()
    n, h, w, c, k, R, stride, pad = case
    x = _rand(n, h, w, c)
    wt = _rand(k, c, R, R, scale=(c * R * R) ** -0.5)
    ssum = torch.zeros(k, device=_dev())
    ssq = torch.zeros(k, device=_dev())
    y = C.conv2d_fprop(x, _w2d(wt), R, R, stride, pad, stats=(ssum, ssq))
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), wt.float(), stride=stride, padding=pad).permute(0, 2, 3, 1)
    _close(y, ref, 1e-2, f"fprop{case}")
    yf = y.float()
    _close(ssum, yf.sum(dim=(0, 1, 2)), 2e-3, f"stats_sum{case}")
    _close(ssq, (yf * yf).sum(dim=(0, 1, 2)), 2e-3, f"stats_sq{case}")


def test_fprop_bias_relu_slice():
    from deeplearning_mpi_b200.ops import conv as C

    _torch_setup()
    n, h, w, c, k = 2, 16, 16, 128, 64
    buf_in = _rand(n, h, w, c + 64)
    x = buf_in[..., 64:]  # channel slice as input
    wt = _rand(k, c, 3, 3, scale=(c * 9) ** -0.5)
    bias = torch.randn(k, device=_dev())
    buf_out = torch.zeros(n, h, w, k + 128, dtype=torch.bfloat16, device=_dev())
    out = buf_out[..., 128:]
    C.conv2d_fprop(x, _w2d(wt), 3, 3, 1, 1, out=out, bias=bias, relu=True)
    ref = F.relu(F.conv2d(x.float().permute(0, 3, 1, 2), wt.float(), bias, padding=1)).permute(0, 2, 3, 1)
    _close(out, ref, 1e-2, "fprop_bias_relu_slice")
    assert buf_out[..., :128].abs().max().item() == 0.0


@pytest.mark.parametrize("case", CONV_CASES)
def test_dgrad(case):
    from deeplearning_mpi_b200.ops import conv as C

    _torch_setup()
    n, h, w, c, k, R, stride, pad = case
    oh = (h + 2 * pad - R) // stride + 1
    ow = (w + 2 * pad - R) // stride + 1
    dy = _rand(n, oh, ow, k)
    wt = _rand(k, c, R, R, scale=(k * R * R) ** -0.5)
    dx = C.conv2d_dgrad(dy, _wt2d(wt), R, R, stride, pad, (h, w))
    ref = torch.nn.grad.conv2d_input((n, c, h, w), wt.float(), dy.float().permute(0, 3, 1, 2), stride=stride,
                                     padding=pad).permute(0, 2, 3, 1)
    _close(dx, ref, 1e-2, f"dgrad{case}")


@pytest.mark.parametrize("case", [(4, 14, 14, 256, 512, 1, 1, 0), (2, 56, 56, 64, 64, 3, 1, 1),
                                  (1, 32, 32, 192, 64, 3, 1, 1), (3, 28, 28, 128, 256, 1, 1, 0)])
def test_dgrad_addend(case):
    """Residual-branch gradient added in the dgrad epilogue (flat 1x1 and boxed 3x3 layouts)."""
    from deeplearning_mpi_b200.ops import conv as C

    _torch_setup()
    n, h, w, c, k, R, stride, pad = case
    dy = _rand(n, h, w, k)
    wt = _rand(k, c, R, R, scale=(k * R * R) ** -0.5)
    add = _rand(n, h, w, c)
    dx = C.conv2d_dgrad(dy, _wt2d(wt), R, R, stride, pad, (h, w), addend=add)
    ref = torch.nn.grad.conv2d_input((n, c, h, w), wt.float(), dy.float().permute(0, 3, 1, 2), stride=stride,
                                     padding=pad).permute(0, 2, 3, 1) + add.float()
    _close(dx, ref, 1e-2, f"dgrad_addend{case}")


def test_dgrad_stride2_accumulate():
    """Strided 1x1 (downsample) and 3x3 dgrad accumulated onto an existing tensor through the parity views."""
    from deeplearning_mpi_b200.ops import conv as C

    _torch_setup()
    for (n, h, c, k, R, pad) in [(2, 28, 256, 512, 1, 0), (2, 14, 128, 128, 3, 1)]:
        oh = (h + 2 * pad - R) // 2 + 1
        dy = _rand(n, oh, oh, k)
        wt = _rand(k, c, R, R, scale=(k * R * R) ** -0.5)
        base = _rand(n, h, h, c)
        out = base.clone()
        C.conv2d_dgrad(dy, _wt2d(wt), R, R, 2, pad, (h, h), out=out, accumulate=True)
        ref = torch.nn.grad.conv2d_input((n, c, h, h), wt.float(), dy.float().permute(0, 3, 1, 2), stride=2,
                                         padding=pad).permute(0, 2, 3, 1) + base.float()
        _close(out, ref, 1e-2, f"dgrad_s2_accumulate{(n, h, c, k, R)}")


def test_fprop_many_tiles_per_cta():
    """More tiles than SMs x pipeline depth: exercises the persistent loop, both accumulator stages, both staging
    buffers and the register-resident statistics for every tile width (BN = 64, 128, 256)."""
    from deeplearning_mpi_b200.ops import conv as C

    _torch_setup()
    for (n, h, c, k) in [(64, 56, 64, 64), (64, 28, 64, 256), (32, 28, 128, 512)]:
        x = _rand(n, h, h, c)
        wt = _rand(k, c, 1, 1, scale=c ** -0.5)
        ssum = torch.zeros(k, device=_dev())
        ssq = torch.zeros(k, device=_dev())
        y = C.conv2d_fprop(x, _w2d(wt), 1, 1, 1, 0, stats=(ssum, ssq))
        ref = (x.float().view(-1, c) @ wt.float().view(k, c).t()).view(n, h, h, k)
        _close(y, ref, 1e-2, f"fprop_many{(n, h, c, k)}")
        yf = y.float()
        _close(ssum, yf.sum(dim=(0, 1, 2)), 2e-3, f"stats_sum_many{(n, h, c, k)}")
        _close(ssq, (yf * yf).sum(dim=(0, 1, 2)), 2e-3, f"stats_sq_many{(n, h, c, k)}")


@pytest.mark.parametrize("case", CONV_CASES)
def test_wgrad(case):
    from deeplearning_mpi_b200.ops import conv as C

    _torch_setup()
    n, h, w, c, k, R, stride, pad = case
    oh = (h + 2 * pad - R) // stride + 1
    ow = (w + 2 * pad - R) // stride + 1
    x = _rand(n, h, w, c)
    dy = _rand(n, oh, ow, k)
    dw = torch.zeros(k, R, R, c, device=_dev())
    C.conv2d_wgrad(dy, x, dw, R, R, stride, pad)
    ref = torch.nn.grad.conv2d_weight(x.float().permute(0, 3, 1, 2), (k, c, R, R), dy.float().permute(0, 3, 1, 2),
                                      stride=stride, padding=pad).permute(0, 2, 3, 1)
    _close(dw, ref, 1e-2, f"wgrad{case}")
    # accumulation semantics: a second call doubles the gradient
    C.conv2d_wgrad(dy, x, dw, R, R, stride, pad)
    _close(dw, 2 * ref, 1e-2, f"wgrad_accum{case}")


def test_linear():
    from deeplearning_mpi_b200.ops import conv as C

    _torch_setup()
    m, kk, nout = 256, 2048, 1000
    x = _rand(m, kk)
    w = _rand(nout, kk, scale=kk ** -0.5)
    b = torch.randn(nout, device=_dev())
    y = C.linear_fwd(x, w, b)
    _close(y, x.float() @ w.float().t() + b, 1e-2, "linear_fwd")
    npad = 1024
    dy = torch.zeros(m, npad, dtype=torch.bfloat16, device=_dev())
    dy[:, :nout] = _rand(m, nout)
    wt = torch.zeros(kk, npad, dtype=torch.bfloat16, device=_dev())
    wt[:, :nout] = w.t()
    dx = C.linear_dgrad(dy, wt)
    _close(dx, dy[:, :nout].float() @ w.float(), 1e-2, "linear_dgrad")
    dw = torch.zeros(nout, kk, device=_dev())
    C.linear_wgrad(dy, x, dw)
    _close(dw, dy[:, :nout].float().t() @ x.float(), 1e-2, "linear_wgrad")


def test_linear_small():
    from deeplearning_mpi_b200.ops import conv as C

    _torch_setup()
    m, kk, nout = 128, 512, 10
    x = _rand(m, kk)
    w = _rand(nout, kk, scale=kk ** -0.5)
    y = C.linear_fwd(x, w, None)
    _close(y, x.float() @ w.float().t(), 1e-2, "linear_small_fwd")


def test_convt2x2():
    from deeplearning_mpi_b200.ops import conv as C

    _torch_setup()
    n, h, w, c, k = 2, 16, 16, 128, 128
    x = _rand(n, h, w, c)
    wt = _rand(c, k, 2, 2, scale=c ** -0.5)  # torch layout [in, out, kh, kw]
    bias = torch.randn(k, device=_dev())
    wq = wt.permute(2, 3, 1, 0).contiguous().view(4 * k, c)        # [(i,j,k), c]
    wtt = wt.permute(0, 2, 3, 1).contiguous().view(c, 4 * k)       # [c, (i,j,k)]
    cat = torch.zeros(n, 2 * h, 2 * w, k + 64, dtype=torch.bfloat16, device=_dev())
    out = cat[..., :k]
    C.convt2x2_fwd(x, wq, out, bias)
    ref = F.conv_transpose2d(x.float().permute(0, 3, 1, 2), wt.float(), bias, stride=2).permute(0, 2, 3, 1)
    _close(out, ref, 1e-2, "convt_fwd")
    assert cat[..., k:].abs().max().item() == 0.0
    dout = _rand(n, 2 * h, 2 * w, k)
    dx = C.convt2x2_dgrad(dout, wtt)
    ref_dx = F.conv2d(dout.float().permute(0, 3, 1, 2), wt.float(), stride=2).permute(0, 2, 3, 1)
    _close(dx, ref_dx, 1e-2, "convt_dgrad")
    dw = torch.zeros(c, 4 * k, device=_dev())
    C.convt2x2_wgrad(dout, x, dw)
    # dw[c, i, j, k] = sum x[n,h,w,c] dout[n,2h+i,2w+j,k]
    d6 = dout.float().view(n, h, 2, w, 2, k)
    ref_dw = torch.einsum("nhwc,nhiwjk->cijk", x.float(), d6).reshape(c, 4 * k)
    _close(dw, ref_dw, 1e-2, "convt_wgrad")


@pytest.mark.parametrize("cfg", [(2, 64, 64, 7, 2, 3, 64), (2, 32, 32, 3, 1, 1, 64)])
def test_smallc(cfg):
    from deeplearning_mpi_b200.ops import conv as C

    _torch_setup()
    n, h, w, R, stride, pad, k = cfg
    x = torch.randn(n, 3, h, w, device=_dev()).to(torch.bfloat16)
    wt = _rand(k, 3, R, R, scale=(3 * R * R) ** -0.5)
    oh, ow, hp, wp = C.smallc_geometry(h, w, R, R, stride, pad)
    xp = torch.zeros(n, hp, wp, 8, dtype=torch.bfloat16, device=_dev())
    xp[:, pad:pad + h, pad:pad + w, :3] = x.permute(0, 2, 3, 1)
    wwin = torch.zeros(k, R, 8, 8, dtype=torch.bfloat16, device=_dev())
    wwin[:, :, :R, :3] = wt.permute(0, 2, 3, 1)
    wwin = wwin.view(k, R * 64)
    ssum = torch.zeros(k, device=_dev())
    ssq = torch.zeros(k, device=_dev())
    y = C.smallc_fprop(xp, wwin, R, stride, (oh, ow), stats=(ssum, ssq))
    ref = F.conv2d(x.float(), wt.float(), stride=stride, padding=pad).permute(0, 2, 3, 1)
    _close(y, ref, 1e-2, f"smallc_fprop{cfg}")
    dy = _rand(n, oh, ow, k)
    dww = torch.zeros(k, R * 64, device=_dev())
    C.smallc_wgrad(dy, xp, dww, R, stride)
    ref_dw = torch.nn.grad.conv2d_weight(x.float(), (k, 3, R, R), dy.float().permute(0, 3, 1, 2), stride=stride,
                                         padding=pad)
    got = dww.view(k, R, 8, 8)[:, :, :R, :3].permute(0, 3, 1, 2)
    _close(got, ref_dw, 1e-2, f"smallc_wgrad{cfg}")


def test_umma_row_shifted_descriptors():
    """Hardware contract the patch-mode kernel relies on: a tcgen05 shared-memory descriptor whose start address is
    shifted by whole 128-byte rows inside a 1024-byte swizzle atom (base-offset field 0) addresses exactly the rows a
    TMA load with 128B swizzle put there, for K-major and MN-major operands."""
    from deeplearning_mpi_b200 import _native as N

    _torch_setup()
    rows = 192
    x = _rand(rows, 64)
    w = _rand(64, 64)
    y = _rand(64, 128)
    lib = N.cuda()
    for mode in (0, 2):
        for shift in (0, 1, 3, 7, 8, 9, 17, 58, 59, 60):
            out = torch.zeros(128, 64, device=_dev())
            src = y if mode else w
            rc = lib.dmpi_umma_probe(x.data_ptr(), src.data_ptr(), out.data_ptr(), rows, shift, mode, N.stream_ptr())
            assert rc == 0
            torch.cuda.synchronize()
            ref = (y.float().t() @ x[shift:shift + 64].float()) if mode else (x[shift:shift + 128].float() @ w.float().t())
            _close(out, ref, 1e-2, f"umma_shift mode={mode} shift={shift}")


def test_fprop_statistics_are_bit_reproducible():
    """The per-channel sums of the conv epilogue are added in a fixed order (last-CTA reduction of per-CTA partials):
    repeated launches must agree bit for bit, for one and several N tiles, CTA pairs, patch mode and the 4-view
    transposed convolution (several column groups per channel)."""
    from deeplearning_mpi_b200.ops import conv as C

    _torch_setup()
    for (n, h, c, k, R, pad) in [(64, 56, 64, 64, 1, 0), (32, 28, 128, 512, 1, 0), (16, 28, 256, 256, 3, 1),
                                 (16, 56, 64, 64, 3, 1), (8, 14, 512, 1024, 1, 0)]:
        x = _rand(n, h, h, c)
        wt = _rand(k, c, R, R, scale=(c * R * R) ** -0.5)
        w2d = _w2d(wt)
        runs = []
        for _ in range(3):
            ssum = torch.zeros(k, device=_dev())
            ssq = torch.zeros(k, device=_dev())
            y = C.conv2d_fprop(x, w2d, R, R, 1, pad, stats=(ssum, ssq))
            torch.cuda.synchronize()
            runs.append((ssum.clone(), ssq.clone()))
        yf = y.float()
        _close(runs[0][0], yf.sum(dim=(0, 1, 2)), 2e-3, f"det_stats_sum{(n, h, c, k, R)}")
        _close(runs[0][1], (yf * yf).sum(dim=(0, 1, 2)), 2e-3, f"det_stats_sq{(n, h, c, k, R)}")
        for r in runs[1:]:
            assert torch.equal(r[0], runs[0][0]) and torch.equal(r[1], runs[0][1]), f"not reproducible: {(n, h, c, k, R)}"


def test_fprop_statistics_atomics_mode():
    """DMPI_DETERMINISTIC=0 path: the per-channel sums are published with fp32 atomics instead of the last-CTA
    reduction; same values up to summation order."""
    from deeplearning_mpi_b200.ops import conv as C

    _torch_setup()
    saved = C._DETERMINISTIC
    try:
        C._DETERMINISTIC = False
        for (n, h, c, k, R, pad) in [(64, 56, 64, 64, 1, 0), (16, 28, 256, 256, 3, 1), (8, 14, 512, 1024, 1, 0)]:
            x = _rand(n, h, h, c)
            wt = _rand(k, c, R, R, scale=(c * R * R) ** -0.5)
            ssum = torch.zeros(k, device=_dev())
            ssq = torch.zeros(k, device=_dev())
            y = C.conv2d_fprop(x, _w2d(wt), R, R, 1, pad, stats=(ssum, ssq))
            yf = y.float()
            _close(ssum, yf.sum(dim=(0, 1, 2)), 2e-3, f"atomic_stats_sum{(n, h, c, k, R)}")
            _close(ssq, (yf * yf).sum(dim=(0, 1, 2)), 2e-3, f"atomic_stats_sq{(n, h, c, k, R)}")
    finally:
        C._DETERMINISTIC = saved


TUNE_CASES = [
    (2, 56, 56, 64, 64, 3, 1, 1),      # N = 64: per-tap vs resident-filter patch mode
    (2, 28, 28, 128, 256, 3, 1, 1),    # N tiles 64 / 128 / 256, one CTA or pairs, streamed patch mode
    (4, 14, 14, 256, 512, 1, 1, 0),    # flat 1x1: every N tile / pair combination
    (1, 130, 136, 128, 128, 3, 1, 1),  # wide image: row-segment patch tiles with pairs
    (3, 7, 7, 512, 512, 3, 1, 1),      # tiny M: wave quantisation decides
    (1, 40, 136, 192, 64, 3, 1, 1),    # UNet decoder shape: 3 channel blocks into 64 filters (64-wide tile with CTA pairs)
    (1, 40, 72, 64, 192, 3, 1, 1),     # 192 filters: one partly filled 256-wide tile vs two 128-wide ones
]


@pytest.mark.parametrize("case", TUNE_CASES)
def test_every_tile_configuration(case):
    """Every configuration the autotuner may pick (ops/conv.py::tile_candidates) is numerically the same convolution:
    fprop with BatchNorm statistics and dgrad with a residual addend, each forced in turn."""
    from deeplearning_mpi_b200.ops import conv as C

    _torch_setup()
    n, h, w, c, k, R, stride, pad = case
    x = _rand(n, h, w, c)
    wt = _rand(k, c, R, R, scale=(c * R * R) ** -0.5)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), wt.float(), stride=stride, padding=pad).permute(0, 2, 3, 1)
    dy = _rand(n, h, w, k)
    add = _rand(n, h, w, c)
    ref_dx = torch.nn.grad.conv2d_input((n, c, h, w), wt.float(), dy.float().permute(0, 3, 1, 2), stride=stride,
                                        padding=pad).permute(0, 2, 3, 1) + add.float()
    cands = C.tile_candidates(k, R, R, stride, pad, c, w, h)
    assert len(cands) >= 2
    for bn, cg, pt in cands:
        with C.force_config(bn, cg, pt is not None):
            ssum, ssq = torch.zeros(k, device=_dev()), torch.zeros(k, device=_dev())
            y = C.conv2d_fprop(x, _w2d(wt), R, R, stride, pad, stats=(ssum, ssq))
            _close(y, ref, 1e-2, f"fprop{case} bn={bn} cg={cg} patch={pt}")
            _close(ssum, y.float().sum(dim=(0, 1, 2)), 2e-3, f"stats bn={bn} cg={cg} patch={pt}")
    for bn, cg, pt in C.tile_candidates(c, R, R, stride, pad, k, w, h):
        with C.force_config(bn, cg, pt is not None):
            dx = C.conv2d_dgrad(dy, _wt2d(wt), R, R, stride, pad, (h, w), addend=add)
            _close(dx, ref_dx, 1e-2, f"dgrad+add{case} bn={bn} cg={cg} patch={pt}")


def test_autotune_picks_and_caches():
    from deeplearning_mpi_b200.ops import conv as C

    if not C._AUTOTUNE:
        pytest.skip("DMPI_AUTOTUNE=0")
    _torch_setup()
    x = _rand(4, 14, 14, 256)
    wt = _rand(512, 256, 1, 1, scale=256 ** -0.5)
    C._tuned.clear()
    y1 = C.conv2d_fprop(x, _w2d(wt), 1, 1)
    keys = [k for k in C._tuned if k[0] == "fprop"]
    assert len(keys) == 1 and C._tuned[keys[0]] in C.tile_candidates(512, 1, 1, 1, 0, 256, 14, 14)
    y2 = C.conv2d_fprop(x, _w2d(wt), 1, 1)
    assert torch.equal(y1, y2)


SPLITK_CASES = [
    (128, 2, 2, 256, 256, 3, 1, 1),    # ResNet-18 / 32x32 layer3: 4 tiles, K = 2304
    (128, 1, 1, 512, 512, 3, 1, 1),    # layer4: ONE tile of pixels, K = 4608
    (128, 4, 4, 128, 256, 3, 2, 1),    # stride-2 entry of layer3 (four parity views)
    (64, 2, 2, 1024, 64, 1, 1, 0),     # flat 1x1 into a 64-wide tile
]


@pytest.mark.parametrize("case", SPLITK_CASES)
def test_splitk_tiny_layers(case, monkeypatch):
    """Layers with a handful of output tiles run split-K (several CTAs per tile, partial tiles added in split order):
    same numbers as the fp32 reference, BatchNorm statistics intact, residual addend in the dgrad epilogue, and
    bit-identical from launch to launch."""
    from deeplearning_mpi_b200.ops import conv as C

    monkeypatch.setattr(C, "_SPLITK", True)  # opt-in path (off by default: see ops/conv.py)
    _torch_setup()
    n, h, w, c, k, R, stride, pad = case
    oh = (h + 2 * pad - R) // stride + 1
    ow = (w + 2 * pad - R) // stride + 1
    assert C._pick_splitk(k, R * R * (c // 64), C._m_tiles(ow, oh, n, R == 1 and stride == 1)) is not None
    x = _rand(n, h, w, c)
    wt = _rand(k, c, R, R, scale=(c * R * R) ** -0.5)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), wt.float(), stride=stride, padding=pad).permute(0, 2, 3, 1)
    outs = []
    for _ in range(2):
        ssum, ssq = torch.zeros(k, device=_dev()), torch.zeros(k, device=_dev())
        y = C.conv2d_fprop(x, _w2d(wt), R, R, stride, pad, stats=(ssum, ssq))
        outs.append((y.clone(), ssum.clone(), ssq.clone()))
    _close(outs[0][0], ref, 1e-2, f"splitk fprop{case}")
    yf = outs[0][0].float()
    _close(outs[0][1], yf.sum(dim=(0, 1, 2)), 2e-3, "splitk stats sum")
    _close(outs[0][2], (yf * yf).sum(dim=(0, 1, 2)), 2e-3, "splitk stats sq")
    assert all(torch.equal(a, b) for a, b in zip(outs[0], outs[1])), "split-K result differs between two launches"
    if stride == 1:
        dy = _rand(n, oh, ow, k)
        add = _rand(n, h, w, c)
        dx = C.conv2d_dgrad(dy, _wt2d(wt), R, R, stride, pad, (h, w), addend=add)
        ref_dx = torch.nn.grad.conv2d_input((n, c, h, w), wt.float(), dy.float().permute(0, 3, 1, 2), stride=stride,
                                            padding=pad).permute(0, 2, 3, 1) + add.float()
        _close(dx, ref_dx, 1e-2, f"splitk dgrad+add{case}")
