"""Flat parameter / gradient arenas.

Every parameter of a model becomes a view into one contiguous fp32 buffer, its gradient a view into a
second one laid out in *bucket order* (reverse registration order, i.e. roughly the order in which
backward produces gradients), so that

* a DDP gradient bucket is a contiguous byte range that the NVLink all-reduce kernel processes in place
  (no copy-in / copy-out as in c10d's Reducer — SURVEY.md N8);
* one fused optimizer launch updates the whole model;
* the bf16 mirror consumed by the tcgen05 kernels has identical offsets and is refreshed by that same launch.

Convolution weights keep PyTorch's logical ``[K, C, R, S]`` shape (checkpoint parity with
ref:pytorch/resnet/main.py:139 / ref:pytorch/unet/train.py:216) but are stored physically as
``[K, R, S, C]`` (``channels_last``), which is the K-major operand layout of the implicit-GEMM kernel.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn as nn

PAD = 64  # elements; keeps every tensor 256-byte (fp32) / 128-byte (bf16) aligned
_ARENAS: list = []  # weak references to every live ParamArena (lets the optimizers find the one owning their params)


def _round_up(n: int, m: int = PAD) -> int:
    return (n + m - 1) // m * m


class _RawCudaBuffer:
    """Wraps a raw device pointer (e.g. inside the peer-mapped symmetric arena) for torch.as_tensor."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def tensor_from_ptr(ptr: int, nbytes: int, device: torch.device) -> torch.Tensor:
    return torch.as_tensor(_RawCudaBuffer(ptr, nbytes), device=device)


class TransposeEntry(C.Structure):
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("A", C.c_int), ("T", C.c_int), ("B", C.c_int),
                ("mode", C.c_int), ("ld", C.c_int64), ("tile_begin", C.c_int), ("tiles_a", C.c_int),
                ("tiles_b", C.c_int), ("pad_", C.c_int)]


def build_transpose_table(entries: Sequence[tuple], device: torch.device):
    """entries: (src [A][T][B] bf16, dst, A, T, B, mode, ld).  Returns (device table, n_entries, total_tiles)."""
    arr = (TransposeEntry * len(entries))()
    tile = 0
    for i, (src, dst, A, T, B, mode, ld) in enumerate(entries):
        ta, tb = (A + 31) // 32, (B + 31) // 32
        arr[i] = TransposeEntry(src.data_ptr(), dst.data_ptr(), A, T, B, mode, ld, tile, ta, tb, 0)
        tile += ta * tb * T
    raw = np.frombuffer(bytes(arr), dtype=np.uint8).copy()
    table = torch.from_numpy(raw).to(device)
    return table, len(entries), tile


@dataclass
class ParamSlot:
    name: str
    param: nn.Parameter
    kind: str            # 'conv', 'convt', 'linear', 'stem', 'vector', 'other'
    shape: Tuple[int, ...]
    numel: int
    offset: int = 0      # element offset in the flat arenas
    index: int = 0       # position in bucket-fill order
    t_offset: int = -1   # element offset of the transposed bf16 shadow
    t_shape: Tuple[int, int] = (0, 0)
    ready_cb: Optional[Callable[[], None]] = None
    fused: bool = False  # gradient written by a fused kernel and published through grad_ready() only


def _physical(p: torch.Tensor, kind: str) -> torch.Tensor:
    """Logical parameter -> tensor in the physical (arena) element order."""
    if kind in ("conv", "stem", "convt"):
        return p.permute(0, 2, 3, 1)
    return p


def _logical_view(flat: torch.Tensor, slot: ParamSlot) -> torch.Tensor:
    v = flat[slot.offset: slot.offset + slot.numel]
    if slot.kind in ("conv", "stem", "convt"):
        a, b, r, s = slot.shape
        return v.view(a, r, s, b).permute(0, 3, 1, 2)
    return v.view(slot.shape)


class ParamArena:
    def __init__(self, module: nn.Module, device: torch.device,
                 alloc: Optional[Callable[[str, int], torch.Tensor]] = None):
        """``alloc(name, nbytes)`` returns a uint8 tensor (lets DDP place grads/params/buffers inside the
        peer-mapped symmetric arena); default is a private allocation."""
        self.module = module
        self.device = device
        self.is_cuda = device.type == "cuda"
        def _zeros(shape, dtype):
            """Zero-initialised device tensor without a library fill kernel (own zero-fill kernel on CUDA)."""
            if not self.is_cuda:
                return torch.zeros(shape, dtype=dtype, device=device)
            t = torch.empty(shape, dtype=dtype, device=device)
            self._zero(t)
            return t

        alloc = alloc or (lambda name, nbytes: _zeros(nbytes, torch.uint8))

        kinds: Dict[int, str] = {}
        for m in module.modules():
            for pn, p in m.named_parameters(recurse=False):
                if isinstance(m, nn.ConvTranspose2d) and pn == "weight":
                    kinds[id(p)] = "convt"
                elif isinstance(m, nn.Conv2d) and pn == "weight":
                    kinds[id(p)] = "stem" if p.shape[1] < 8 else "conv"
                elif isinstance(m, nn.Linear) and pn == "weight":
                    kinds[id(p)] = "linear"
                elif p.dim() == 1:
                    kinds.setdefault(id(p), "vector")
        named = [(n, p) for n, p in module.named_parameters()]
        # Trainable parameters first (bucket-fill order), frozen ones (requires_grad=False) behind them: the reducer,
        # the fused optimizers and clipping work on the prefix [0, total_trainable) only, so a frozen parameter is
        # never bucketed, waited for, decayed or updated (torch DDP / torch.optim semantics).
        ordered = [np_ for np_ in reversed(named) if np_[1].requires_grad] + \
                  [np_ for np_ in reversed(named) if not np_[1].requires_grad]
        self.slots: List[ParamSlot] = []
        self.by_param: Dict[int, ParamSlot] = {}
        off = 0
        self.n_trainable = 0
        self.total_trainable = 0
        for idx, (name, p) in enumerate(ordered):
            slot = ParamSlot(name, p, kinds.get(id(p), "other"), tuple(p.shape), p.numel(), off, idx)
            off += _round_up(p.numel())
            self.slots.append(slot)
            self.by_param[id(p)] = slot
            if p.requires_grad:
                self.n_trainable = idx + 1
                self.total_trainable = off
        self.total = max(off, PAD)
        self.total_trainable = max(self.total_trainable, PAD) if self.n_trainable else 0

        self.flat_p = alloc("params", self.total * 4).view(torch.float32)
        self.flat_g = alloc("grads", self.total * 4).view(torch.float32)
        self._zero(self.flat_g)
        self.flat_s = _zeros(self.total, torch.bfloat16) if self.is_cuda else None

        with torch.no_grad():
            for s in self.slots:
                dst = self.flat_p[s.offset: s.offset + s.numel]
                src = s.param.data
                if self.is_cuda and s.kind in ("conv", "stem", "convt"):
                    # the [K,C,R,S] -> [K,R,S,C] re-layout happens once, on the host (two memcpys instead of a strided
                    # library copy kernel per weight)
                    src = src.cpu()
                dst.copy_(_physical(src, s.kind).reshape(-1))
                s.param.data = _logical_view(self.flat_p, s)
                s.param.grad = _logical_view(self.flat_g, s) if s.param.requires_grad else None

        # ---- buffers (BatchNorm running statistics): one byte range, fp32 first then int64
        bufs = [(n, b) for n, b in module.named_buffers() if b is not None]
        f32 = [(n, b) for n, b in bufs if b.dtype == torch.float32]
        i64 = [(n, b) for n, b in bufs if b.dtype == torch.int64]
        other = [(n, b) for n, b in bufs if b.dtype not in (torch.float32, torch.int64)]
        assert not other, f"unsupported buffer dtypes: {[n for n, _ in other]}"
        nb_f32 = sum(_round_up(b.numel(), 4) for _, b in f32)
        nb_i64 = sum(b.numel() for _, b in i64)
        self.buffer_bytes = _round_up(nb_f32 * 4 + nb_i64 * 8, 256)
        self.buffer_list = [b for _, b in f32] + [b for _, b in i64]
        if not self.is_cuda:
            # Stock autograd saves BatchNorm buffers for backward and checks their version counters; views of one flat
            # storage would share a single counter, so on the CPU (stock-module) path buffers stay separate tensors and
            # DDP coalesces them on the fly.
            self.flat_b = None
            f32, i64 = [], []
        else:
            self.flat_b = alloc("buffers", max(self.buffer_bytes, 256))
        fb32 = self.flat_b[: nb_f32 * 4].view(torch.float32) if self.flat_b is not None else None
        fb64 = self.flat_b[nb_f32 * 4: nb_f32 * 4 + nb_i64 * 8].view(torch.int64) if self.flat_b is not None else None
        o = 0
        with torch.no_grad():
            for name, b in f32:
                v = fb32[o: o + b.numel()].view(b.shape)
                v.copy_(b)
                self._set_buffer(name, v)
                o += _round_up(b.numel(), 4)
            o = 0
            for name, b in i64:
                v = fb64[o: o + b.numel()].view(b.shape)
                v.copy_(b)
                self._set_buffer(name, v)
                o += b.numel()

        # ---- transposed bf16 shadows (dgrad / ConvTranspose / Linear-dgrad operands) and stem window weights
        self.flat_t = None
        self.t_table = None
        self.stem_w: Dict[int, torch.Tensor] = {}
        self.stem_dw: Dict[int, torch.Tensor] = {}
        if self.is_cuda:
            toff = 0
            for s in self.slots:
                if s.kind == "conv":
                    k, c, r, q = s.shape
                    s.t_offset, s.t_shape = toff, (c, r * q * k)
                    toff += _round_up(c * r * q * k)
                elif s.kind == "convt":
                    cin, cout, r, q = s.shape
                    s.t_offset, s.t_shape = toff, (r * q * cout, cin)
                    toff += _round_up(cin * cout * r * q)
                elif s.kind == "linear":
                    n_out, k = s.shape
                    npad = _round_up(n_out, 64)
                    s.t_offset, s.t_shape = toff, (k, npad)
                    toff += _round_up(k * npad)
                elif s.kind == "stem":
                    k, c, r, q = s.shape
                    self.stem_w[id(s.param)] = _zeros((k, r * 64), torch.bfloat16)
                    self.stem_dw[id(s.param)] = _zeros((k, r * 64), torch.float32)
            self.flat_t = _zeros(max(toff, PAD), torch.bfloat16)
            entries = []
            for s in self.slots:
                if s.t_offset < 0:
                    continue
                src = self.flat_s[s.offset: s.offset + s.numel]
                dst = self.flat_t[s.t_offset:]
                if s.kind == "conv":
                    k, c, r, q = s.shape
                    entries.append((src, dst, k, r * q, c, 0, r * q * k))
                elif s.kind == "convt":
                    cin, cout, r, q = s.shape
                    entries.append((src, dst, cin, r * q, cout, 1, 0))
                elif s.kind == "linear":
                    n_out, k = s.shape
                    entries.append((src, dst, n_out, 1, k, 0, s.t_shape[1]))
            if entries:
                self.t_table = build_transpose_table(entries, device)
        self._shadow_version = -1
        self.fused_flags: Optional[torch.Tensor] = None  # CPU uint8 per slot, shared with the C++ autograd hooks
        self.bn_scratch: Optional[torch.Tensor] = None
        self._bn_slices: Dict[int, Tuple[torch.Tensor, torch.Tensor, torch.Tensor]] = {}
        self._bn_dirty: Dict[int, bool] = {}
        if self.is_cuda:
            self._build_bn_scratch()
            self.refresh_shadows(force=True)
        import weakref

        _ARENAS.append(weakref.ref(self))

    # ------------------------------------------------------------------ helpers
    def _set_buffer(self, dotted: str, value: torch.Tensor) -> None:
        mod = self.module
        parts = dotted.split(".")
        for p in parts[:-1]:
            mod = getattr(mod, p)
        mod._buffers[parts[-1]] = value

    def _build_bn_scratch(self) -> None:
        bns = [m for m in self.module.modules() if isinstance(m, nn.BatchNorm2d)]
        total = sum(4 * m.num_features for m in bns)
        self.bn_scratch = torch.empty(max(total, 4), dtype=torch.float32, device=self.device)
        self._zero(self.bn_scratch)
        o = 0
        for m in bns:
            c = m.num_features
            self._bn_slices[id(m)] = (self.bn_scratch[o: o + c], self.bn_scratch[o + c: o + 2 * c],
                                      self.bn_scratch[o + 2 * c: o + 4 * c])
            self._bn_dirty[id(m)] = False
            o += 4 * c

    def bn_stats(self, bn: nn.BatchNorm2d):
        """(sum, sqsum) accumulators for the conv epilogue — zero since begin_forward()."""
        s = self._bn_slices[id(bn)]
        return s[0], s[1]

    def bn_acc(self, bn: nn.BatchNorm2d) -> torch.Tensor:
        """Zeroed scratch for the BatchNorm backward reductions."""
        acc = self._bn_slices[id(bn)][2]
        if self._bn_dirty[id(bn)]:
            self._zero(acc)
        self._bn_dirty[id(bn)] = True
        return acc

    def begin_forward(self) -> None:
        """Once per forward pass: clear every BatchNorm accumulator with a single memset and make sure the bf16
        shadows reflect the current master weights."""
        if not self.is_cuda:
            return
        self._zero(self.bn_scratch)
        for k in self._bn_dirty:
            self._bn_dirty[k] = False
        self.refresh_shadows()

    def _param_version(self) -> int:
        return sum(s.param._version for s in self.slots)

    def mark_shadows_fresh(self) -> None:
        self._shadow_version = self._param_version()

    def refresh_shadows(self, force: bool = False) -> None:
        """bf16 mirror + transposed forms.  The fused optimizers write the mirror themselves and call
        refresh_derived(); this full path only runs after out-of-band updates (load_state_dict, broadcast,
        a torch optimizer), detected through the tensors' version counters."""
        if not self.is_cuda:
            return
        ver = self._param_version()
        if not force and ver == self._shadow_version:
            return
        from ..ops import kernels as K

        K.cast_bf16(self.flat_p, self.flat_s, self.total)
        self.refresh_derived()
        self._shadow_version = ver

    def refresh_derived(self) -> None:
        from ..ops import kernels as K

        if self.t_table is not None:
            K.transpose_table(*self.t_table)
        for s in self.slots:
            if s.kind == "stem":
                k, c, r, q = s.shape
                K.stem_pack(self.flat_p[s.offset: s.offset + s.numel], self.stem_w[id(s.param)], k, r, q, c)

    # ------------------------------------------------------------------ operand accessors (CUDA path)
    def w_natural(self, p: nn.Parameter) -> torch.Tensor:
        s = self.by_param[id(p)]
        v = self.flat_s[s.offset: s.offset + s.numel]
        if s.kind == "conv":
            k, c, r, q = s.shape
            return v.view(k, r * q * c)
        if s.kind == "convt":
            cin, cout, r, q = s.shape
            return v.view(cin, r * q * cout)
        if s.kind == "linear":
            return v.view(s.shape)
        raise ValueError(s.kind)

    def w_transposed(self, p: nn.Parameter) -> torch.Tensor:
        s = self.by_param[id(p)]
        a, b = s.t_shape
        return self.flat_t[s.t_offset: s.t_offset + a * b].view(a, b)

    def grad_physical(self, p: nn.Parameter) -> torch.Tensor:
        """fp32 gradient slot in physical element order (what the wgrad kernels accumulate into)."""
        s = self.by_param[id(p)]
        self._ensure_grad(s)
        if not s.fused:
            s.fused = True  # this parameter's gradient is published by grad_ready(), never by an autograd hook
            if self.fused_flags is not None and s.index < self.fused_flags.numel():
                self.fused_flags[s.index] = 1
        v = self.flat_g[s.offset: s.offset + s.numel]
        if s.kind in ("conv", "stem"):
            k, c, r, q = s.shape
            return v.view(k, r, q, c)
        if s.kind == "convt":
            cin, cout, r, q = s.shape
            return v.view(cin, r * q * cout)
        return v.view(s.shape)

    def _ensure_grad(self, s: ParamSlot) -> None:
        if not s.param.requires_grad:
            return  # frozen: the fused kernels still get a (tail-of-arena) slot to write into; nothing reads it
        g = s.param.grad
        if g is None or g.data_ptr() != self.flat_g.data_ptr() + s.offset * 4:
            # e.g. torch.optim's zero_grad(set_to_none=True): start again from a clean slot
            self._zero(self.flat_g[s.offset: s.offset + _round_up(s.numel)])
            s.param.grad = _logical_view(self.flat_g, s)

    def grad_ready(self, p: nn.Parameter) -> None:
        """Called by the fused backward kernels' wrappers once the slot of ``p`` holds this iteration's gradient."""
        s = self.by_param[id(p)]
        if s.ready_cb is not None:
            s.ready_cb()

    @staticmethod
    def _zero(t: torch.Tensor) -> None:
        from ..ops import kernels as K

        K.zero_(t)

    def zero_grad(self) -> None:
        self._zero(self.flat_g)
        for s in self.slots[: self.n_trainable]:
            if s.param.grad is None or s.param.grad.data_ptr() != self.flat_g.data_ptr() + s.offset * 4:
                s.param.grad = _logical_view(self.flat_g, s)


def get_arena(module: nn.Module) -> Optional[ParamArena]:
    return getattr(module, "_dmpi_arena", None)


def ensure_arena(module: nn.Module, alloc=None) -> ParamArena:
    """Creates (once) the arena of a root module whose parameters already live on their final device."""
    arena = get_arena(module)
    if arena is not None and alloc is None:
        s = arena.slots[0]
        if s.param.device == arena.device and s.param.data_ptr() == arena.flat_p.data_ptr() + s.offset * 4:
            return arena
        # parameters were moved / replaced after the arena was built (e.g. model.to(device)): rebuild
    first = next(module.parameters())
    arena = ParamArena(module, first.device, alloc)
    if not getattr(module, "_dmpi_sd_hook", False):
        # Checkpoints must be ordinary tensors (contiguous, own storage, torch-native layouts) so that the
        # reference scripts can torch.load them — not strided views into the flat arena.
        def _plain_state_dict(mod, sd, prefix, local_metadata):
            for k in list(sd.keys()):
                if k.startswith(prefix) and torch.is_tensor(sd[k]):
                    sd[k] = sd[k].detach().clone(memory_format=torch.contiguous_format)
            return sd

        module._register_state_dict_hook(_plain_state_dict)
        object.__setattr__(module, "_dmpi_sd_hook", True)
    object.__setattr__(module, "_dmpi_arena", arena)
    for m in module.modules():
        object.__setattr__(m, "_dmpi_arena_ref", arena)
    return arena
