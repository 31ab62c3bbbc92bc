"""Lowering of Conv2d / ConvTranspose2d / Linear (forward, data-gradient, weight-gradient) onto the two
tcgen05 kernels in ``csrc/kernels/conv`` (``dmpi_igemm_run`` and ``dmpi_wgrad_run``).

Conventions
-----------
* Activations are ``torch.bfloat16`` tensors of logical shape ``[N, H, W, C]`` with unit channel stride
  (NHWC).  Channel *slices* of a wider buffer are fine — that is how the UNet skip "concat" is made
  zero-copy (ref:pytorch/unet/model.py:47 does a ``torch.cat`` copy instead).
* Conv weights reach the kernels as a K-major bf16 matrix ``[K_out, R*S*C]`` (fprop) or
  ``[C, R*S*K_out]`` (dgrad).  fp32 gradients are accumulated in ``[K_out, R, S, C]`` order, which is
  exactly the physical layout of a ``channels_last`` ``[K_out, C, R, S]`` parameter.

This module only builds descriptors and caches plans; all math runs in the native kernels.
"""
from __future__ import annotations

import ctypes as C
import os
import functools
from typing import Optional, Sequence, Tuple

import torch

from .. import _native as N

__all__ = [
    "conv2d_fprop", "conv2d_dgrad", "conv2d_wgrad", "linear_fwd", "linear_dgrad", "linear_wgrad",
    "convt2x2_fwd", "convt2x2_dgrad", "convt2x2_wgrad", "smallc_fprop", "smallc_wgrad", "choose_box",
]

_plans: dict = {}


def _check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed (rc={rc}): {N.conv_last_error()}")


@functools.lru_cache(maxsize=None)
def choose_box(ow: int, oh: int, n: int, rows: int = 128, exact: bool = False) -> Tuple[int, int, int]:
    """Pick the (w, h, n) pixel box of one M tile: product <= rows (== rows when ``exact``), maximising the
    fraction of useful rows over the whole tensor."""
    best, best_eff, best_rank = None, -1.0, (0, 0)
    wc = sorted({min(ow, rows)} | {-(-ow // k) for k in range(1, 17)} | {1 << i for i in range(0, 8)})
    for bw in wc:
        if bw < 1 or bw > rows or bw > 256:
            continue
        hmax = rows // bw
        hc = sorted({min(oh, hmax)} | {-(-oh // k) for k in range(1, 17)} | {1 << i for i in range(0, 8)})
        for bh in hc:
            if bh < 1 or bh > hmax:
                continue
            nmax = rows // (bw * bh)
            for bn in sorted({min(n, nmax), nmax} | {1 << i for i in range(0, 8)}):
                if bn < 1 or bn > nmax or bn > 256:
                    continue
                if exact and bw * bh * bn != rows:
                    continue
                tiles = -(-ow // bw) * -(-oh // bh) * -(-n // bn)
                eff = (ow * oh * n) / (tiles * rows)
                # ties: prefer the most compact spatial patch (halo reuse), then the widest rows
                rank = (bw * bh, bw)
                if eff > best_eff + 1e-9 or (abs(eff - best_eff) <= 1e-9 and rank > best_rank):
                    best, best_eff, best_rank = (bw, bh, bn), eff, rank
    if best is None:
        raise ValueError(f"no pixel box for {ow}x{oh}x{n} rows={rows}")
    return best


def _t4(dims: Sequence[int], strides: Sequence[int], offset: int = 0) -> N.Tensor4:
    t = N.Tensor4()
    for i in range(4):
        t.dims[i] = int(dims[i])
    for i in range(3):
        t.strides[i] = int(strides[i])
    t.offset = int(offset)
    return t


def _nhwc_meta(x: torch.Tensor):
    assert x.dim() == 4 and x.dtype == torch.bfloat16 and x.stride(3) == 1, (x.shape, x.dtype, x.stride())
    n, h, w, c = x.shape
    sn, sh, sw, _ = x.stride()
    return n, h, w, c, sn, sh, sw


def _is_flat(n, h, w, sn, sh, sw) -> bool:
    return sh == w * sw and (n == 1 or sn == h * sh)


_BN256_MIN_K = int(os.environ.get("DMPI_BN256_MIN_K", "256"))
_BN256_ODD = os.environ.get("DMPI_BN256_ODD", "1") == "1"  # A/B switch of the rule for widths such as 192


def _pick_bn(n_out: int, ktot: int = 1 << 30) -> int:
    """N tile of the implicit GEMM.  256-wide tiles halve the A-operand traffic per FLOP (the 3x3 and deep 1x1 layers
    are bound by L2 -> SM operand bandwidth: 83 -> 56 us on 14x14x256 3x3), but with a short reduction (K <= 128) the
    tile is epilogue-paced and two 128-wide tiles pipeline better (58 vs 73 us on 28x28 128 -> 512)."""
    if n_out <= 64:
        return 64
    if n_out % 256 == 0 and ktot >= _BN256_MIN_K:
        return 256
    # widths such as 192 (the UNet decoder's 64 -> [128 | 64] concat gradient): two 128-wide tiles and one 256-wide tile
    # pad to the same 256 columns, but the wide tile reads the pixel operand once (tuner: 866 vs 1 099 us at 512x512)
    if _BN256_ODD and n_out > 128 and ktot >= _BN256_MIN_K and -(-n_out // 256) * 256 <= -(-n_out // 128) * 128:
        return 256
    return 128


def _views_for_stride(h, w, c, sn, sh, sw, n, stride):
    """One strided view per input parity class (stride 2) or the plain view (stride 1)."""
    if stride == 1:
        return [_t4((c, w, h, n), (sw, sh, sn))], lambda a, b: 0
    views = []
    for a in range(2):
        for b in range(2):
            views.append(_t4((c, (w - b + 1) // 2, (h - a + 1) // 2, n), (2 * sw, 2 * sh, sn), a * sh + b * sw))
    return views, lambda a, b: a * 2 + b


def _get_plan(key, build):
    ent = _plans.get(key)
    if ent is None:
        desc = build()
        handle = N.cuda().dmpi_conv_plan_create(C.byref(desc))
        if not handle:
            raise RuntimeError("dmpi_conv_plan_create failed")
        ent = (handle, desc)
        _plans[key] = ent
    return ent[0]


# DMPI_DETERMINISTIC: 0 = fp32 atomics everywhere; 1 (default) = order-independent BatchNorm statistics and backward
# reductions (bit-identical forward, gradients equal to ~1e-7; costs ~2 % of a ResNet-50 step); 2 = additionally the
# split-K weight gradients are added in split order (bit-identical gradients, what cudnn.deterministic asks of the
# reference; costs another ~5 %: one partial tile per CTA written and re-read).
_DET_LEVEL = os.environ.get("DMPI_DETERMINISTIC", "1")
_DETERMINISTIC = os.environ.get("DMPI_DETERMINISTIC_CONV", "0" if _DET_LEVEL == "0" else "1") == "1"
_DETERMINISTIC_WGRAD = os.environ.get("DMPI_DETERMINISTIC_WGRAD", "1" if _DET_LEVEL == "2" else "0") == "1"
_STAT_SCRATCH_BYTES = 512 << 10  # >= (148 CTAs * 2 * 256 columns + 1) * 4
_stat_scratch = {}


def _stat_scratch_for_current_stream(device) -> torch.Tensor:
    """Zero-initialised scratch for the deterministic BatchNorm statistics, one per (device, stream): launches on one
    stream are ordered, so they can share it (the kernel leaves its ticket counter at zero)."""
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    t = _stat_scratch.get(key)
    if t is None:
        t = torch.zeros(_STAT_SCRATCH_BYTES // 4, dtype=torch.float32, device=device)
        _stat_scratch[key] = t
    return t


_WGRAD_DET_BYTES = 32 << 20  # one partial tile per CTA of a one-wave launch (<= 29 MiB on 148 SMs)
_wgrad_scratch = {}


def _det_wgrad_plan(plan, device):
    """Attaches (or detaches) the scratch that makes a weight-gradient launch add its split-K partials in a fixed
    order (deterministic) instead of with red.add; one scratch per (device, stream)."""
    if _DETERMINISTIC_WGRAD:
        key = (device.index, torch.cuda.current_stream(device).cuda_stream)
        t = _wgrad_scratch.get(key)
        if t is None:
            t = torch.empty(_WGRAD_DET_BYTES // 4, dtype=torch.float32, device=device)
            _wgrad_scratch[key] = t
        N.cuda().dmpi_conv_plan_set_stat_scratch(plan, t.data_ptr(), _WGRAD_DET_BYTES)
    else:
        N.cuda().dmpi_conv_plan_set_stat_scratch(plan, None, 0)
    return plan


# Split-K for layers with only a handful of output tiles (the 2x2 / 1x1 feature maps at the end of ResNet-18 on 32x32
# inputs): each CTA otherwise streams its whole reduction at the bandwidth one SM gets (55-120 GB/s), 20-35 us per launch
# whatever the tile shape (DMPI_AUTOTUNE_LOG).  With splits the K range is spread over up to 8 CTAs per tile; the CTA
# that finishes a tile last adds the fp32 partial tiles in split order (deterministic, bit-identical run to run).
# OPT-IN (DMPI_SPLITK=1): correct (tests/test_conv_gpu.py::test_splitk_tiny_layers) but, as built, SLOWER end to end -
# ResNet-18 32x32 batch 128: 1.95 ms/step with it, 1.40 ms without (same box): parking and re-reading the partial tiles
# plus 64-144 resident 200 KB CTAs per tiny launch cost more than the streaming they save.  Kept as the starting point
# for a cluster / DSMEM reduction; off by default.
_SPLITK = os.environ.get("DMPI_SPLITK", "0") == "1"
_SK_BYTES = 16 << 20  # 4 KiB of tickets + up to 148 partial tiles of 128 x 128 floats
_sk_scratch = {}


def _pick_splitk(n_out: int, num_kb: int, m_tiles: int):
    """(bn, splits) for a split-K launch, or None."""
    if not _SPLITK or _forced is not None or num_kb < 16 or n_out % 64:
        return None
    bn = 128 if n_out >= 128 else 64
    units = m_tiles * -(-n_out // bn)
    if units > 24:
        return None
    splits = min(8, 148 // units, num_kb // 4)
    while splits > 1 and (splits - 1) * -(-num_kb // splits) >= num_kb:
        splits -= 1
    return (bn, splits) if splits >= 2 else None


def _attach_splitk_scratch(plan, device) -> None:
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    t = _sk_scratch.get(key)
    if t is None:
        t = torch.zeros(_SK_BYTES, dtype=torch.uint8, device=device)  # tickets must start at zero; the kernel keeps them so
        _sk_scratch[key] = t
    N.cuda().dmpi_conv_plan_set_splitk_scratch(plan, t.data_ptr(), _SK_BYTES)


def _m_tiles(ow: int, oh: int, n: int, flat: bool) -> int:
    if flat:
        return -(-(n * oh * ow) // 128)
    bw, bh, bnn = choose_box(ow, oh, n)
    return -(-ow // bw) * -(-oh // bh) * -(-n // bnn)


def _run_igemm(plan, a, b, out, bias, stats, addend=None, add_strides=(0, 0, 0)):
    s0 = stats[0].data_ptr() if stats is not None else None
    s1 = stats[1].data_ptr() if stats is not None else None
    if stats is not None:
        if _DETERMINISTIC:
            scratch = _stat_scratch_for_current_stream(a.device)
            N.cuda().dmpi_conv_plan_set_stat_scratch(plan, scratch.data_ptr(), _STAT_SCRATCH_BYTES)
        else:
            N.cuda().dmpi_conv_plan_set_stat_scratch(plan, None, 0)
    if addend is None:
        rc = N.cuda().dmpi_igemm_run(plan, a.data_ptr(), b.data_ptr(), out.data_ptr(),
                                     bias.data_ptr() if bias is not None else None, s0, s1, N.stream_ptr())
    else:
        rc = N.cuda().dmpi_igemm_run_add(plan, a.data_ptr(), b.data_ptr(), out.data_ptr(),
                                         bias.data_ptr() if bias is not None else None, s0, s1, addend.data_ptr(),
                                         add_strides[0], add_strides[1], add_strides[2], N.stream_ptr())
    N.count_launch()
    _check(rc, "dmpi_igemm_run")


_CTA_GROUP = int(os.environ.get("DMPI_CTA_GROUP", "0"))  # 0: automatic, 1 / 2: forced (where legal)
_CG2_MIN_K = int(os.environ.get("DMPI_CG2_MIN_K", "256"))


def _pick_cg(bn: int, ktot: int) -> int:
    """CTA pairs (tcgen05.mma.cta_group::2): each CTA loads half of the weight tile, so the operand traffic per FLOP
    drops by 1/4 (BN=128) to 1/3 (BN=256).  Worth it where the main loop, not the epilogue, paces the kernel."""
    if bn < 128 or _CTA_GROUP == 1:
        return 1
    if _CTA_GROUP == 2:
        return 2
    return 2 if ktot >= _CG2_MIN_K else 1


_PATCH = int(os.environ.get("DMPI_PATCH", "1"))
_PATCH_MIN_EFF = float(os.environ.get("DMPI_PATCH_MIN_EFF", "0.85"))


def choose_patch_tile(ow: int, oh: int, halo: int = 1) -> Optional[Tuple[int, int, float]]:
    """Tile (columns, rows) of the patch-mode implicit GEMM for a stride-1 R x S convolution, or None.

    The accumulator rows run over the padded tile width pw = tw + 2*halo, so tw output columns x th output rows fit when
    (th-1)*pw + tw <= 128; the efficiency is the fraction of accumulator rows that are real output pixels."""
    best = None
    for tw in sorted({min(ow, 128)} | {-(-ow // k) for k in range(1, 9)}):
        if tw < 8 or tw > 128:
            continue
        pw = tw + 2 * halo
        th = min(oh, (128 - tw) // pw + 1)
        if (th - 1) * pw + tw > 128:
            continue
        tiles = -(-ow // tw) * -(-oh // th)
        eff = ow * oh / (tiles * 128.0)
        if best is None or eff > best[2] + 1e-9:
            best = (tw, th, eff)
    return best


_PATCH_PAIR = int(os.environ.get("DMPI_PATCH_PAIR", "1"))  # patch mode together with CTA pairs for N >= 128


def _patch_geom(R, S, stride, pad, c_in, ow, oh, bn, cg):
    """Patch-mode tile for a 3x3 stride-1 layer when the geometry and the kernel's shared-memory ring allow it (no
    policy): two patch stages plus at least two filter-tap tiles must fit (igemm.cu: kRingBytes)."""
    if R != S or R != 3 or stride != 1 or pad != 1 or c_in % 64 != 0:
        return None
    t = choose_patch_tile(ow, oh, pad)
    if t is None or t[2] < _PATCH_MIN_EFF:
        return None
    if cg == 2:
        ring = {64: 180, 128: 144, 256: 160}[bn] * 1024
        tap_bytes = bn * 64
    else:
        ring = {64: 192, 128: 160, 256: 144}[bn] * 1024
        tap_bytes = bn * 128
    pw = t[0] + 2 * pad
    rows = max(128 + 2 * pad * pw + 2 * pad, pw * (t[1] + 2 * pad))
    stride_b = -(-rows * 128 // 1024) * 1024
    if ring - 2 * stride_b < 2 * tap_bytes:
        return None
    return t


def _patch_ok(R, S, stride, pad, c_in, ow, oh, n_out=64, cg=1):
    """Heuristic choice (used when a shape has not been autotuned).  With N <= 64 the whole filter stays resident in
    shared memory (56x56x64: 88 vs 136 us).  With wider N the filter taps stream; that only pays together with CTA pairs
    (each CTA loads its own patch and half of every tap tile, ~2.3x less L2 -> SM traffic than per-tap loads), so the
    single-CTA streamed variant is used only when forced (DMPI_PATCH=2)."""
    if not _PATCH:
        return None
    bn = _pick_bn(n_out, 9 * c_in)
    if _PATCH != 2 and n_out > 64:
        # bench/igemm_patch_ab.py: with CTA pairs the 128-wide tile gains 8-15 % on large images (256x256: 318 -> 282 us)
        # and ties on small ones; the 256-wide tile loses 8-12 % (its per-tap loads were already amortised)
        if not (cg == 2 and _PATCH_PAIR and bn == 128 and ow >= 128):
            return None
    return _patch_geom(R, S, stride, pad, c_in, ow, oh, bn, cg)


# --------------------------------------------------------------------------------------- autotuning
# The reference runs with torch.backends.cudnn.benchmark = True (ref:pytorch/resnet/main.py:29): cuDNN times its
# algorithms on the first call of every shape and keeps the fastest.  Same here: the first eager call of a convolution
# shape times every legal tile configuration of the implicit GEMM (N tile 64 / 128 / 256, one CTA or a CTA pair, per-tap
# loads or patch mode) on scratch outputs and the shape keeps the winner - which also settles wave quantisation (196
# tiles on 148 SMs) and the L2-bound / epilogue-bound trade-offs that a closed-form rule gets wrong somewhere.
# Opt-in (DMPI_AUTOTUNE=1): on the three benchmark models the closed-form rules below already sit at the tuner's
# optimum (A/B on one box: ResNet-50 18.88 vs 18.95 ms, UNet 30.41 vs 30.56 ms, ResNet-18 32x32 1.370 vs 1.349 ms per
# step), and without it the tile shapes - hence the summation order of the BatchNorm statistics - are identical from
# run to run, which keeps results bit-identical across processes.  Nothing is timed while a stream is being captured.
_AUTOTUNE = os.environ.get("DMPI_AUTOTUNE", "0") == "1"
_tuned: dict = {}
_forced = None  # (bn, cg, use_patch) set by force_config(): tests / A-B scripts


class force_config:
    """Context manager that pins the implicit-GEMM tile configuration: (bn, cg, patch) with patch True / False."""

    def __init__(self, bn: int, cg: int, patch: bool):
        self.cfg = (bn, cg, patch)

    def __enter__(self):
        global _forced
        self.prev, _forced = _forced, self.cfg
        return self

    def __exit__(self, *exc):
        global _forced
        _forced = self.prev


def tile_candidates(n_out, R, S, stride, pad, c_in, ow, oh):
    """Every (bn, cg, patch tile or None) the kernel can run for this layer."""
    out = []
    for bn in (64, 128, 256):
        if bn > 64 and bn > n_out and bn // 2 >= n_out:
            continue  # a narrower tile already covers all filters
        for cg in (1, 2):
            out.append((bn, cg, None))
            t = _patch_geom(R, S, stride, pad, c_in, ow, oh, bn, cg) if _PATCH else None
            if t is not None:
                out.append((bn, cg, t))
    return out


def _can_time(device) -> bool:
    return _AUTOTUNE and _forced is None and device.type == "cuda" and not torch.cuda.is_current_stream_capturing()


def _choose(skey, heuristic, cands, run, device):
    """Configuration for shape ``skey``: forced > tuned > (tune now) > heuristic.  ``run(cfg)`` executes one launch on
    scratch outputs."""
    if _forced is not None:
        bn, cg, want_patch = _forced
        for c in cands:
            if c[0] == bn and c[1] == cg and (c[2] is not None) == bool(want_patch):
                return c
        raise ValueError(f"forced tile configuration {_forced} is not legal for this layer (candidates: {cands})")
    got = _tuned.get(skey)
    if got is not None:
        return got
    if not _can_time(device) or len(cands) < 2:
        return heuristic
    best, best_t = heuristic, float("inf")
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    seen = []
    for c in cands:
        try:
            run(c)  # plan creation + first launch
            ts = []
            for _ in range(3):
                ev[0].record()
                run(c)
                ev[1].record()
                ev[1].synchronize()
                ts.append(ev[0].elapsed_time(ev[1]))
            t = min(ts)
        except RuntimeError:
            continue  # a configuration the kernel refuses (e.g. patch ring too small): not a candidate
        seen.append((c[0], c[1], c[2] is not None, round(t * 1e3, 1)))
        if t < best_t * (0.97 if c != heuristic else 1.0):  # the closed-form choice wins ties (3 %)
            best, best_t = c, min(t, best_t) if c == heuristic else t
    _tuned[skey] = best
    if os.environ.get("DMPI_AUTOTUNE_LOG"):
        print(f"[dmpi autotune] {skey[:1] + skey[1:5] + skey[8:13]}: {best} ({best_t * 1e3:.1f} us; heuristic {heuristic}) "
              f"all (bn, cg, patch, us): {seen}", flush=True)
    return best


# --------------------------------------------------------------------------------------- fprop
def conv2d_fprop(x: torch.Tensor, w2d: torch.Tensor, R: int, S: int, stride: int = 1, pad: int = 0,
                 out: Optional[torch.Tensor] = None, bias: Optional[torch.Tensor] = None,
                 stats: Optional[Tuple[torch.Tensor, torch.Tensor]] = None, relu: bool = False) -> torch.Tensor:
    """y[n,oh,ow,k] = sum_{r,s,c} x[n, oh*stride+r-pad, ow*stride+s-pad, c] * w2d[k, (r*S+s)*C + c] (+bias)."""
    n, h, w, c, sn, sh, sw = _nhwc_meta(x)
    assert c % 64 == 0, "igemm needs C_in % 64 == 0 (use smallc_fprop for the RGB stem)"
    kout = w2d.shape[0]
    assert w2d.shape[1] == R * S * c and w2d.dtype == torch.bfloat16 and w2d.is_contiguous()
    oh = (h + 2 * pad - R) // stride + 1
    ow = (w + 2 * pad - S) // stride + 1
    if out is None:
        out = torch.empty((n, oh, ow, kout), dtype=torch.bfloat16, device=x.device)
    on, ohh, oww, oc, osn, osh, osw = _nhwc_meta(out)
    assert (on, ohh, oww, oc) == (n, oh, ow, kout)
    flat = (R == 1 and S == 1 and stride == 1 and pad == 0 and _is_flat(n, h, w, sn, sh, sw)
            and _is_flat(n, oh, ow, osn, osh, osw))
    bn = _pick_bn(kout, R * S * c)
    cg = _pick_cg(bn, R * S * c)
    ptile = _patch_ok(R, S, stride, pad, c, ow, oh, kout, cg)
    heur = (bn, cg, ptile)
    skey = ("fprop", n, h, w, c, sn, sh, sw, kout, R, S, stride, pad, osn, osh, osw, bias is not None, stats is not None, relu)

    sk = _pick_splitk(kout, R * S * (c // 64), _m_tiles(ow, oh, n, flat)) if (_tuned.get(skey) is None) else None

    def go(cfg, out, stats):
        bn, cg, ptile = cfg
        splits = 0
        if sk is not None and cfg == heur:
            bn, cg, ptile, splits = sk[0], 1, None, sk[1]
        on, ohh, oww, oc, osn, osh, osw = _nhwc_meta(out)
        key = ("fprop", n, h, w, c, sn, sh, sw, kout, R, S, stride, pad, osn, osh, osw, bias is not None,
               stats is not None, relu, bn, cg, ptile, splits)

        def build():
            d = N.ConvDesc()
            if flat:
                m = n * h * w
                d.a_nviews = 1
                d.a_view[0] = _t4((c, m, 1, 1), (sw, m * sw, m * sw))
                d.box_w, d.box_h, d.box_n = 128, 1, 1
                d.grid_w, d.grid_h, d.grid_n = m, 1, 1
                d.ntaps = 1
                d.tap_view[0] = 0
                d.o_nviews = 1
                d.o_view[0] = _t4((kout, m, 1, 1), (osw, m * osw, m * osw))
            else:
                views, vid = _views_for_stride(h, w, c, sn, sh, sw, n, stride)
                d.a_nviews = len(views)
                for i, v in enumerate(views):
                    d.a_view[i] = v
                if ptile is not None:
                    bw, bh, bnn = ptile[0], ptile[1], 1
                    d.patch, d.halo = 1, pad
                else:
                    bw, bh, bnn = choose_box(ow, oh, n)
                d.box_w, d.box_h, d.box_n = bw, bh, bnn
                d.grid_w, d.grid_h, d.grid_n = ow, oh, n
                d.ntaps = R * S
                assert d.ntaps <= N.MAX_TAPS
                t = 0
                for r in range(R):
                    for s in range(S):
                        if stride == 1:
                            d.tap_view[t], d.tap_dh[t], d.tap_dw[t] = 0, r - pad, s - pad
                        else:
                            a, b = (r - pad) % 2, (s - pad) % 2
                            d.tap_view[t] = vid(a, b)
                            d.tap_dh[t], d.tap_dw[t] = (r - pad - a) // 2, (s - pad - b) // 2
                        d.tap_koff[t] = t * c
                        t += 1
                d.o_nviews = 1
                d.o_view[0] = _t4((kout, ow, oh, n), (osw, osh, osn))
            d.cblocks = c // 64
            d.b_rows, d.b_k = kout, R * S * c
            d.nblk_per_oview = 0
            d.n_out, d.bn = kout, bn
            d.cta_group = cg
            d.bias_mod = kout if bias is not None else 0
            d.stat_mod = kout if stats is not None else 0
            d.relu = int(relu)
            d.out_mode = 0
            d.splits = splits
            return d

        plan = _get_plan(key, build)
        if splits:
            _attach_splitk_scratch(plan, x.device)
        _run_igemm(plan, x, w2d, out, bias, stats)

    cfg = _tuned.get(skey) if _forced is None else None
    if cfg is None:
        cfg = heur
        if _forced is not None or _can_time(x.device):
            scratch = []

            def trial(c_):  # launches on scratch outputs (same strides): timing never touches the caller's tensors
                if not scratch:
                    scratch.append(torch.empty_strided(out.shape, out.stride(), dtype=out.dtype, device=out.device))
                    scratch.append((torch.zeros_like(stats[0]), torch.zeros_like(stats[1])) if stats is not None else None)
                go(c_, scratch[0], scratch[1])

            cfg = _choose(skey, heur, tile_candidates(kout, R, S, stride, pad, c, ow, oh), trial, x.device)
    go(cfg, out, stats)
    return out


# --------------------------------------------------------------------------------------- dgrad
def conv2d_dgrad(dy: torch.Tensor, wt2d: torch.Tensor, R: int, S: int, stride: int, pad: int,
                 in_hw: Tuple[int, int], out: Optional[torch.Tensor] = None,
                 addend: Optional[torch.Tensor] = None, accumulate: bool = False) -> torch.Tensor:
    """dx[n,ih,iw,c] = sum_{r,s,k} dy[n,(ih+pad-r)/stride,(iw+pad-s)/stride,k] * wt2d[c, (r*S+s)*K + k] (+ addend).

    ``addend`` (stride-1 only): a bf16 NHWC tensor of dx's shape added in the kernel epilogue before rounding — the
    gradient arriving at the same tensor through the residual branch, so no separate add kernel runs.
    ``accumulate`` (stride-2 only, needs ``out``): the result is added onto what ``out`` already holds — each parity
    class reads its own strided view of ``out`` as the addend; positions that receive no tap are left untouched."""
    n, oh, ow, k, sn, sh, sw = _nhwc_meta(dy)
    assert k % 64 == 0
    c = wt2d.shape[0]
    assert wt2d.shape[1] == R * S * k and wt2d.is_contiguous()
    ih, iw = in_hw
    assert not accumulate or (out is not None and stride == 2)
    if out is None:
        out = (torch.zeros if (stride == 2 and R == 1) else torch.empty)(
            (n, ih, iw, c), dtype=torch.bfloat16, device=dy.device)
    _, _, _, _, osn, osh, osw = _nhwc_meta(out)
    ktot = (R * S if stride == 1 else max(1, (R * S) // 4)) * k
    bn = _pick_bn(c, ktot)
    cg = _pick_cg(bn, ktot)
    if stride == 1:
        flat = R == 1 and S == 1 and pad == 0 and _is_flat(n, oh, ow, sn, sh, sw) and _is_flat(n, ih, iw, osn, osh, osw)
        heur = (bn, cg, _patch_ok(R, S, 1, pad, k, iw, ih, c, cg))
        skey = ("dgrad1", n, oh, ow, k, sn, sh, sw, c, R, S, pad, osn, osh, osw, addend is not None)
        add_strides = None
        if addend is not None:
            an, ah, aw, ac, asn, ash, asw = _nhwc_meta(addend)
            assert (an, ah, aw, ac) == (n, ih, iw, c), (addend.shape, (n, ih, iw, c))
            if flat:
                assert _is_flat(n, ih, iw, asn, ash, asw)
                m_ = n * ih * iw
                add_strides = (m_ * asw, m_ * asw, asw)  # same (c, m, 1, 1) view convention as the flat output
            else:
                add_strides = (asn, ash, asw)

        sk = _pick_splitk(c, R * S * (k // 64), _m_tiles(iw, ih, n, flat)) if (_tuned.get(skey) is None) else None

        def go(cfg, out):
            bn, cg, ptile = cfg
            splits = 0
            if sk is not None and cfg == heur:
                bn, cg, ptile, splits = sk[0], 1, None, sk[1]
            key = ("dgrad1", n, oh, ow, k, sn, sh, sw, c, R, S, pad, osn, osh, osw, bn, cg, ptile, splits)

            def build():
                d = N.ConvDesc()
                d.a_nviews = 1
                if flat:
                    m = n * oh * ow
                    d.a_view[0] = _t4((k, m, 1, 1), (sw, m * sw, m * sw))
                    d.box_w, d.box_h, d.box_n = 128, 1, 1
                    d.grid_w, d.grid_h, d.grid_n = m, 1, 1
                    d.ntaps = 1
                    d.o_view[0] = _t4((c, m, 1, 1), (osw, m * osw, m * osw))
                else:
                    d.a_view[0] = _t4((k, ow, oh, n), (sw, sh, sn))
                    if ptile is not None:
                        bw, bh, bnn = ptile[0], ptile[1], 1
                        d.patch, d.halo = 1, pad
                    else:
                        bw, bh, bnn = choose_box(iw, ih, n)
                    d.box_w, d.box_h, d.box_n = bw, bh, bnn
                    d.grid_w, d.grid_h, d.grid_n = iw, ih, n
                    d.ntaps = R * S
                    t = 0
                    for r in range(R):
                        for s in range(S):
                            d.tap_view[t], d.tap_dh[t], d.tap_dw[t], d.tap_koff[t] = 0, pad - r, pad - s, t * k
                            t += 1
                    d.o_view[0] = _t4((c, iw, ih, n), (osw, osh, osn))
                d.o_nviews = 1
                d.cblocks = k // 64
                d.b_rows, d.b_k = c, R * S * k
                d.n_out, d.bn = c, bn
                d.cta_group = cg
                d.splits = splits
                return d

            plan = _get_plan(key, build)
            if splits:
                _attach_splitk_scratch(plan, dy.device)
            if addend is not None:
                _run_igemm(plan, dy, wt2d, out, None, None, addend, add_strides)
            else:
                _run_igemm(plan, dy, wt2d, out, None, None)

        cfg = _tuned.get(skey) if _forced is None else None
        if cfg is None:
            cfg = heur
            if _forced is not None or _can_time(dy.device):
                scratch = []

                def trial(c_):
                    if not scratch:
                        scratch.append(torch.empty_strided(out.shape, out.stride(), dtype=out.dtype, device=out.device))
                    go(c_, scratch[0])

                cfg = _choose(skey, heur, tile_candidates(c, R, S, 1, pad, k, iw, ih), trial, dy.device)
        go(cfg, out)
        return out

    assert stride == 2 and addend is None
    # Parity decomposition: input pixels with (ih%2, iw%2) == (a, b) only see taps r == (a+pad) mod 2,
    # s == (b+pad) mod 2; on that sub-grid the gather is a dense stride-1 convolution over dy.
    for a in range(2):
        for b in range(2):
            rs = [r for r in range(R) if (a + pad - r) % 2 == 0]
            ss = [s for s in range(S) if (b + pad - s) % 2 == 0]
            sub_h, sub_w = (ih - a + 1) // 2, (iw - b + 1) // 2
            if sub_h <= 0 or sub_w <= 0:
                continue
            if not rs or not ss:
                if not (R == 1) and not accumulate:
                    out[:, a::2, b::2, :].zero_()
                continue
            key = ("dgrad2", a, b, n, oh, ow, k, sn, sh, sw, c, R, S, pad, ih, iw, osn, osh, osw, bn, cg)

            def build(a=a, b=b, rs=rs, ss=ss, sub_h=sub_h, sub_w=sub_w):
                d = N.ConvDesc()
                d.a_nviews = 1
                d.a_view[0] = _t4((k, ow, oh, n), (sw, sh, sn))
                bw, bh, bnn = choose_box(sub_w, sub_h, n)
                d.box_w, d.box_h, d.box_n = bw, bh, bnn
                d.grid_w, d.grid_h, d.grid_n = sub_w, sub_h, n
                t = 0
                for r in rs:
                    for s in ss:
                        d.tap_view[t] = 0
                        d.tap_dh[t] = (a + pad - r) // 2
                        d.tap_dw[t] = (b + pad - s) // 2
                        d.tap_koff[t] = (r * S + s) * k
                        t += 1
                d.ntaps = t
                d.o_nviews = 1
                d.o_view[0] = _t4((c, sub_w, sub_h, n), (2 * osw, 2 * osh, osn), a * osh + b * osw)
                d.cblocks = k // 64
                d.b_rows, d.b_k = c, R * S * k
                d.n_out, d.bn = c, bn
                d.cta_group = cg
                return d

            if accumulate:
                view = out[:, a::2, b::2, :]  # same pixels and strides as the output view of this parity class
                _run_igemm(_get_plan(key, build), dy, wt2d, out, None, None, view,
                           (view.stride(0), view.stride(1), view.stride(2)))
            else:
                _run_igemm(_get_plan(key, build), dy, wt2d, out, None, None)
    return out


# --------------------------------------------------------------------------------------- wgrad
def _wgrad_bn(c: int) -> int:
    if c >= 256 and c % 256 == 0:
        return 256
    if c >= 128 and c % 128 == 0:
        return 128
    return 64


def conv2d_wgrad(dy: torch.Tensor, x: torch.Tensor, dw: torch.Tensor, R: int, S: int, stride: int, pad: int) -> None:
    """dw[k, r, s, c] += sum_{n,oh,ow} dy[n,oh,ow,k] * x[n, oh*stride+r-pad, ow*stride+s-pad, c]  (dw fp32)."""
    n, oh, ow, k, sn, sh, sw = _nhwc_meta(dy)
    xn, h, w, c, xsn, xsh, xsw = _nhwc_meta(x)
    assert xn == n and c % 64 == 0 and dw.dtype == torch.float32
    bn = _wgrad_bn(c)
    if (_WGRAD_PATCH and R == 3 and S == 3 and stride == 1 and pad == 1 and k % 64 == 0
            and (_WGRAD_PATCH == 2 or (k <= _WGRAD_PATCH_MAXK and n * oh * ow >= _WGRAD_PATCH_MIN_PIXELS))):
        # k <= 64 is where the tap-group kernel wastes half of every M = 128 MMA (bench/wgrad_ab.py: 512x512 192->64
        # 1772 -> 858 us, 128->64 811 -> 575 us; at k = 128 the two kernels tie or the tap-group one wins)
        return _conv2d_wgrad_patch(dy, x, dw, pad)
    key = ("wgrad", n, oh, ow, k, sn, sh, sw, h, w, c, xsn, xsh, xsw, R, S, stride, pad, bn)

    def build():
        d = N.ConvDesc()
        views, vid = _views_for_stride(h, w, c, xsn, xsh, xsw, n, stride)
        d.a_nviews = len(views)
        for i, v in enumerate(views):
            d.a_view[i] = v
        bw, bh, bnn = choose_box(ow, oh, n, rows=64, exact=True)
        d.box_w, d.box_h, d.box_n = bw, bh, bnn
        d.grid_w, d.grid_h, d.grid_n = ow, oh, n
        d.ntaps = R * S
        t = 0
        for r in range(R):
            for s in range(S):
                if stride == 1:
                    d.tap_view[t], d.tap_dh[t], d.tap_dw[t] = 0, r - pad, s - pad
                else:
                    a, b = (r - pad) % 2, (s - pad) % 2
                    d.tap_view[t] = vid(a, b)
                    d.tap_dh[t], d.tap_dw[t] = (r - pad - a) // 2, (s - pad - b) // 2
                d.tap_woff[t] = t * c
                t += 1
        d.o_nviews = 1
        d.o_view[0] = _t4((k, ow, oh, n), (sw, sh, sn))
        d.k_out, d.c_total, d.w_ld = k, c, R * S * c
        d.bn = bn
        d.splits = 0
        return d

    rc = N.cuda().dmpi_wgrad_run(_det_wgrad_plan(_get_plan(key, build), dw.device), dy.data_ptr(), x.data_ptr(),
                                 dw.data_ptr(), N.stream_ptr())
    N.count_launch()
    _check(rc, "dmpi_wgrad_run")


_WGRAD_PATCH = int(os.environ.get("DMPI_WGRAD_PATCH", "1"))  # 0 off, 1 automatic, 2 every eligible 3x3
_WGRAD_PATCH_MAXK = int(os.environ.get("DMPI_WGRAD_PATCH_MAXK", "64"))
_WGRAD_PATCH_MIN_PIXELS = 1 << 16  # below this the launch is latency-bound and the tap-group kernel is 2 % ahead


def _conv2d_wgrad_patch(dy: torch.Tensor, x: torch.Tensor, dw: torch.Tensor, pad: int) -> None:
    """3x3 stride-1 weight gradient from one shared-memory input patch per pixel tile (csrc/kernels/conv/wgrad_patch.cu):
    full-width tiles of a few rows for narrow images, 128-pixel row segments for wide ones."""
    n, oh, ow, k, sn, sh, sw = _nhwc_meta(dy)
    _, h, w, c, xsn, xsh, xsw = _nhwc_meta(x)
    if ow <= 127:
        bw = ow
        bh = max(1, min(oh, 128 // (ow + 2 * pad)))
    else:
        bw, bh = 128, 1
    key = ("wgrad_patch", n, oh, ow, k, sn, sh, sw, h, w, c, xsn, xsh, xsw, pad, bw, bh)

    def build():
        d = N.ConvDesc()
        d.a_nviews = 1
        d.a_view[0] = _t4((c, w, h, n), (xsw, xsh, xsn))
        d.box_w, d.box_h, d.box_n = bw, bh, 1
        d.grid_w, d.grid_h, d.grid_n = ow, oh, n
        d.patch, d.halo = 1, pad
        d.ntaps = 9
        t = 0
        for r in range(3):
            for s_ in range(3):
                d.tap_view[t], d.tap_dh[t], d.tap_dw[t] = 0, r - pad, s_ - pad
                d.tap_woff[t] = t * c
                t += 1
        d.o_nviews = 1
        d.o_view[0] = _t4((k, ow, oh, n), (sw, sh, sn))
        d.k_out, d.c_total, d.w_ld = k, c, 9 * c
        d.splits = int(os.environ.get("DMPI_WGRAD_PATCH_SPLITS", "0"))
        return d

    rc = N.cuda().dmpi_wgrad_patch_run(_det_wgrad_plan(_get_plan(key, build), dw.device), dy.data_ptr(), x.data_ptr(), dw.data_ptr(),
                                       N.stream_ptr())
    N.count_launch()
    _check(rc, "dmpi_wgrad_patch_run")


# --------------------------------------------------------------------------------------- linear
def linear_fwd(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x [M, K] bf16, w [N_out, K] bf16 -> fp32 [M, N_out] (logits stay fp32 for the loss)."""
    m, kk = x.shape
    nout = w.shape[0]
    assert kk % 64 == 0 and x.is_contiguous() and w.is_contiguous()
    out = torch.empty((m, nout), dtype=torch.float32, device=x.device)
    bn = 64 if nout <= 64 else 128
    key = ("linfwd", m, kk, nout, bias is not None, bn)

    def build():
        d = N.ConvDesc()
        d.a_nviews = 1
        d.a_view[0] = _t4((kk, m, 1, 1), (kk, m * kk, m * kk))
        d.box_w, d.box_h, d.box_n = 128, 1, 1
        d.grid_w, d.grid_h, d.grid_n = m, 1, 1
        d.ntaps, d.cblocks = 1, kk // 64
        d.b_rows, d.b_k = nout, kk
        d.o_nviews = 0
        d.n_out, d.bn = nout, bn
        d.bias_mod = nout if bias is not None else 0
        d.out_mode, d.out_ld = 1, nout
        return d

    _run_igemm(_get_plan(key, build), x, w, out, bias, None)
    return out


def linear_dgrad(dy: torch.Tensor, wt: torch.Tensor) -> torch.Tensor:
    """dy [M, N_pad] bf16 (N_pad % 64 == 0), wt [K, N_pad] bf16 -> dx [M, K] bf16."""
    m, npad = dy.shape
    kk = wt.shape[0]
    assert npad % 64 == 0 and wt.shape[1] == npad
    y4 = dy.view(1, 1, m, npad)
    out = torch.empty((1, 1, m, kk), dtype=torch.bfloat16, device=dy.device)
    conv2d_fprop(y4, wt, 1, 1, out=out)
    return out.view(m, kk)


def linear_wgrad(dy: torch.Tensor, x: torch.Tensor, dw: torch.Tensor) -> None:
    """dw[N_out, K] += dy[M, N_pad]^T x[M, K]   (rows of dy past N_out are ignored)."""
    m, npad = dy.shape
    kk = x.shape[1]
    nout = dw.shape[0]
    bn = _wgrad_bn(kk)
    key = ("linwgrad", m, npad, kk, nout, bn)

    def build():
        d = N.ConvDesc()
        d.a_nviews = 1
        d.a_view[0] = _t4((kk, m, 1, 1), (kk, m * kk, m * kk))
        d.box_w, d.box_h, d.box_n = 64, 1, 1
        d.grid_w, d.grid_h, d.grid_n = m, 1, 1
        d.ntaps = 1
        d.o_nviews = 1
        d.o_view[0] = _t4((npad, m, 1, 1), (npad, m * npad, m * npad))
        d.k_out, d.c_total, d.w_ld = nout, kk, kk
        d.bn = bn
        return d

    rc = N.cuda().dmpi_wgrad_run(_det_wgrad_plan(_get_plan(key, build), dw.device), dy.data_ptr(), x.data_ptr(),
                                 dw.data_ptr(), N.stream_ptr())
    N.count_launch()
    _check(rc, "dmpi_wgrad_run")


# --------------------------------------------------------------------------------------- ConvTranspose2d k=2 s=2
def convt2x2_fwd(x: torch.Tensor, wq: torch.Tensor, out: torch.Tensor, bias: Optional[torch.Tensor]) -> torch.Tensor:
    """out[n, 2h+i, 2w+j, k] = sum_c x[n,h,w,c] * wq[(i*2+j)*K + k, c] + bias[k].

    A plain GEMM [pixels, C] x [C, 4K]; the epilogue scatters each (i, j) quadrant through a strided output
    view, so writing into the upper channel slice of the UNet concat buffer costs nothing extra
    (ref:pytorch/unet/model.py:36-38,46-47)."""
    n, h, w, c, sn, sh, sw = _nhwc_meta(x)
    kq = wq.shape[0]
    k = kq // 4
    assert c % 64 == 0 and k % 64 == 0 and wq.shape[1] == c
    on, oh, ow, oc, osn, osh, osw = _nhwc_meta(out)
    assert (on, oh, ow, oc) == (n, 2 * h, 2 * w, k)
    bn = 64 if k % 128 else 128
    key = ("convt", n, h, w, c, sn, sh, sw, k, osn, osh, osw, bias is not None, bn)

    def build():
        d = N.ConvDesc()
        d.a_nviews = 1
        d.a_view[0] = _t4((c, w, h, n), (sw, sh, sn))
        bw, bh, bnn = choose_box(w, h, n)
        d.box_w, d.box_h, d.box_n = bw, bh, bnn
        d.grid_w, d.grid_h, d.grid_n = w, h, n
        d.ntaps, d.cblocks = 1, c // 64
        d.b_rows, d.b_k = kq, c
        d.o_nviews = 4
        for i in range(2):
            for j in range(2):
                d.o_view[i * 2 + j] = _t4((k, w, h, n), (2 * osw, 2 * osh, osn), i * osh + j * osw)
        d.nblk_per_oview = k // bn
        d.n_out, d.bn = kq, bn
        d.bias_mod = k if bias is not None else 0
        return d

    _run_igemm(_get_plan(key, build), x, wq, out, bias, None)
    return out


def convt2x2_dgrad(dout: torch.Tensor, wt: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """dx[n,h,w,c] = sum_{i,j,k} dout[n,2h+i,2w+j,k] * wt[c, (i*2+j)*K + k]."""
    n, oh, ow, k, sn, sh, sw = _nhwc_meta(dout)
    c = wt.shape[0]
    h, w = oh // 2, ow // 2
    assert k % 64 == 0 and wt.shape[1] == 4 * k
    if out is None:
        out = torch.empty((n, h, w, c), dtype=torch.bfloat16, device=dout.device)
    _, _, _, _, osn, osh, osw = _nhwc_meta(out)
    bn = _pick_bn(c, 4 * k)
    key = ("convt_dgrad", n, oh, ow, k, sn, sh, sw, c, osn, osh, osw, bn)

    def build():
        d = N.ConvDesc()
        d.a_nviews = 4
        for i in range(2):
            for j in range(2):
                d.a_view[i * 2 + j] = _t4((k, w, h, n), (2 * sw, 2 * sh, sn), i * sh + j * sw)
        bw, bh, bnn = choose_box(w, h, n)
        d.box_w, d.box_h, d.box_n = bw, bh, bnn
        d.grid_w, d.grid_h, d.grid_n = w, h, n
        d.ntaps, d.cblocks = 4, k // 64
        for t in range(4):
            d.tap_view[t], d.tap_koff[t] = t, t * k
        d.b_rows, d.b_k = c, 4 * k
        d.o_nviews = 1
        d.o_view[0] = _t4((c, w, h, n), (osw, osh, osn))
        d.n_out, d.bn = c, bn
        return d

    _run_igemm(_get_plan(key, build), dout, wt, out, None, None)
    return out


def convt2x2_wgrad(dout: torch.Tensor, x: torch.Tensor, dw: torch.Tensor) -> None:
    """dw[c, (i*2+j)*K + k] += sum_{n,h,w} x[n,h,w,c] * dout[n,2h+i,2w+j,k]   (dw fp32 [C, 4K])."""
    n, oh, ow, k, sn, sh, sw = _nhwc_meta(dout)
    xn, h, w, c, xsn, xsh, xsw = _nhwc_meta(x)
    bn = _wgrad_bn(k)
    key = ("convt_wgrad", n, oh, ow, k, sn, sh, sw, c, xsn, xsh, xsw, bn)

    def build():
        d = N.ConvDesc()
        d.a_nviews = 4
        for i in range(2):
            for j in range(2):
                d.a_view[i * 2 + j] = _t4((k, w, h, n), (2 * sw, 2 * sh, sn), i * sh + j * sw)
        bw, bh, bnn = choose_box(w, h, n, rows=64, exact=True)
        d.box_w, d.box_h, d.box_n = bw, bh, bnn
        d.grid_w, d.grid_h, d.grid_n = w, h, n
        d.ntaps = 4
        for t in range(4):
            d.tap_view[t], d.tap_woff[t] = t, t * k
        d.o_nviews = 1
        d.o_view[0] = _t4((c, w, h, n), (xsw, xsh, xsn))
        d.k_out, d.c_total, d.w_ld = c, k, 4 * k
        d.bn = bn
        return d

    # M side = x (rows of dw are input channels), N side = the four dout quadrants.
    rc = N.cuda().dmpi_wgrad_run(_det_wgrad_plan(_get_plan(key, build), dw.device), x.data_ptr(), dout.data_ptr(),
                                 dw.data_ptr(), N.stream_ptr())
    N.count_launch()
    _check(rc, "dmpi_wgrad_run")


# --------------------------------------------------------------------------------------- RGB stem (C_in = 3)
def smallc_geometry(h: int, w: int, R: int, S: int, stride: int, pad: int):
    """Padded NHWC8 staging buffer geometry for the small-channel trick (see smallc_fprop)."""
    oh = (h + 2 * pad - R) // stride + 1
    ow = (w + 2 * pad - S) // stride + 1
    hp = h + 2 * pad
    wp = max(w + 2 * pad, stride * (ow - 1) + 8)
    return oh, ow, hp, wp


def smallc_fprop(xp: torch.Tensor, wwin: torch.Tensor, R: int, stride: int, out_hw: Tuple[int, int],
                 out: Optional[torch.Tensor] = None, bias: Optional[torch.Tensor] = None,
                 stats: Optional[Tuple[torch.Tensor, torch.Tensor]] = None) -> torch.Tensor:
    """First-layer convolution with C_in <= 8 (RGB): 7x7/s2 (ResNet) and 3x3/s1 (UNet).

    ``xp`` is the zero-padded input in NHWC8 (``[N, Hp, Wp, 8]`` bf16).  Because W and C are adjacent in
    memory, the S taps x 8 channels of one filter row are one contiguous 64-element window; a tensor map
    with an *overlapping* W stride (stride*8 elements) presents those windows as the "channels" of a
    [pixels, 64] operand, so each filter row is one K block and the stem runs on the same tcgen05 kernel.
    ``wwin`` is the matching weight repack ``[K_out, R*64]`` (zeros where s >= S or c >= C_in)."""
    n, hp, wp, c8 = xp.shape
    assert c8 == 8 and xp.is_contiguous() and xp.dtype == torch.bfloat16
    oh, ow = out_hw
    kout = wwin.shape[0]
    assert wwin.shape[1] == R * 64
    if out is None:
        out = torch.empty((n, oh, ow, kout), dtype=torch.bfloat16, device=xp.device)
    _, _, _, _, osn, osh, osw = _nhwc_meta(out)
    bn = _pick_bn(kout)
    key = ("smallc", n, hp, wp, R, stride, oh, ow, kout, osn, osh, osw, bias is not None, stats is not None, bn)

    def build():
        d = N.ConvDesc()
        srow, simg = wp * 8, hp * wp * 8
        if stride == 1:
            d.a_nviews = 1
            d.a_view[0] = _t4((64, ow, hp, n), (8, srow, simg))
        else:
            d.a_nviews = 2
            for a in range(2):
                d.a_view[a] = _t4((64, ow, (hp - a + 1) // 2, n), (16, 2 * srow, simg), a * srow)
        bw, bh, bnn = choose_box(ow, oh, n)
        d.box_w, d.box_h, d.box_n = bw, bh, bnn
        d.grid_w, d.grid_h, d.grid_n = ow, oh, n
        d.ntaps, d.cblocks = R, 1
        for r in range(R):
            if stride == 1:
                d.tap_view[r], d.tap_dh[r] = 0, r
            else:
                d.tap_view[r], d.tap_dh[r] = r % 2, r // 2
            d.tap_dw[r] = 0
            d.tap_koff[r] = r * 64
        d.b_rows, d.b_k = kout, R * 64
        d.o_nviews = 1
        d.o_view[0] = _t4((kout, ow, oh, n), (osw, osh, osn))
        d.n_out, d.bn = kout, bn
        d.bias_mod = kout if bias is not None else 0
        d.stat_mod = kout if stats is not None else 0
        return d

    _run_igemm(_get_plan(key, build), xp, wwin, out, bias, stats)
    return out


def smallc_wgrad(dy: torch.Tensor, xp: torch.Tensor, dwwin: torch.Tensor, R: int, stride: int) -> None:
    """dwwin[k, r*64 + (s*8+c)] += sum dy[n,oh,ow,k] * window(xp)[n,oh,ow,r][s*8+c]   (fp32 [K_out, R*64])."""
    n, oh, ow, k, sn, sh, sw = _nhwc_meta(dy)
    _, hp, wp, _ = xp.shape
    key = ("smallc_wgrad", n, oh, ow, k, sn, sh, sw, hp, wp, R, stride)

    def build():
        d = N.ConvDesc()
        srow, simg = wp * 8, hp * wp * 8
        if stride == 1:
            d.a_nviews = 1
            d.a_view[0] = _t4((64, ow, hp, n), (8, srow, simg))
        else:
            d.a_nviews = 2
            for a in range(2):
                d.a_view[a] = _t4((64, ow, (hp - a + 1) // 2, n), (16, 2 * srow, simg), a * srow)
        bw, bh, bnn = choose_box(ow, oh, n, rows=64, exact=True)
        d.box_w, d.box_h, d.box_n = bw, bh, bnn
        d.grid_w, d.grid_h, d.grid_n = ow, oh, n
        d.ntaps = R
        for r in range(R):
            if stride == 1:
                d.tap_view[r], d.tap_dh[r] = 0, r
            else:
                d.tap_view[r], d.tap_dh[r] = r % 2, r // 2
            d.tap_woff[r] = r * 64
        d.o_nviews = 1
        d.o_view[0] = _t4((k, ow, oh, n), (sw, sh, sn))
        d.k_out, d.c_total, d.w_ld = k, 64, R * 64
        d.bn = 64
        return d

    rc = N.cuda().dmpi_wgrad_run(_det_wgrad_plan(_get_plan(key, build), dwwin.device), dy.data_ptr(), xp.data_ptr(),
                                 dwwin.data_ptr(), N.stream_ptr())
    N.count_launch()
    _check(rc, "dmpi_wgrad_run")
