"""Input pipeline helpers.

* :class:`CudaPrefetcher` — pinned-memory, side-stream host->device staging with one batch of look-ahead.  This is the
  role ``pin_memory=True`` + ``.to(device)`` + DDP's side-stream input copy play for the reference
  (ref:pytorch/resnet/main.py:96-102,126), except that the copy of batch i+1 overlaps the compute
  of batch i instead of serialising with it.
* :class:`SyntheticImages` / :class:`SyntheticSegmentation` — datasets of the reference's shapes for machines without
  the CIFAR-10 / microscopy data (there is no network in the benchmark environment).
* :class:`SegmentationFolder` (aliases ``BasicDataset`` / ``CarvanaDataset``) — file-backed segmentation dataset with
  the item contract of ref:pytorch/unet/data_loading.py:52-134, a parallel / rank-0-shared palette scan and a
  vectorised value -> class mapping.
"""
from __future__ import annotations

import os
from pathlib import Path
from typing import Iterable, Iterator, Optional, Sequence

import numpy as np
import torch
from torch.utils.data import Dataset


class CudaPrefetcher:
    """Wraps an iterable of CPU batches (tensor, tuple/list of tensors, or dict of tensors)."""

    def __init__(self, loader: Iterable, device: torch.device):
        self.loader = loader
        self.device = device
        self.stream = torch.cuda.Stream(device=device)
        self.bytes_per_batch = 0

    def _to_device(self, obj):
        if torch.is_tensor(obj):
            src = obj if obj.is_pinned() else obj.pin_memory()
            self.bytes_per_batch += src.numel() * src.element_size()
            return src.to(self.device, non_blocking=True), src
        if isinstance(obj, dict):
            out, keep = {}, []
            for k, v in obj.items():
                out[k], kp = self._to_device(v)
                keep.append(kp)
            return out, keep
        if isinstance(obj, (list, tuple)):
            pairs = [self._to_device(v) for v in obj]
            return type(obj)(p[0] for p in pairs), [p[1] for p in pairs]
        return obj, None

    def _record(self, obj):
        if torch.is_tensor(obj):
            obj.record_stream(torch.cuda.current_stream(self.device))
        elif isinstance(obj, dict):
            for v in obj.values():
                self._record(v)
        elif isinstance(obj, (list, tuple)):
            for v in obj:
                self._record(v)

    def __iter__(self) -> Iterator:
        it = iter(self.loader)
        nxt = None
        keep = None

        def preload():
            nonlocal nxt, keep
            try:
                batch = next(it)
            except StopIteration:
                nxt = None
                return
            self.bytes_per_batch = 0
            with torch.cuda.stream(self.stream):
                nxt, keep = self._to_device(batch)

        preload()
        while nxt is not None:
            torch.cuda.current_stream(self.device).wait_stream(self.stream)
            cur = nxt
            self._record(cur)
            preload()
            yield cur

    def __len__(self) -> int:
        return len(self.loader)


class SyntheticImages(Dataset):
    """Random normalised images + integer labels (CIFAR-10 / ImageNet shaped)."""

    def __init__(self, n: int, shape: Sequence[int] = (3, 32, 32), num_classes: int = 10, seed: int = 0):
        g = torch.Generator().manual_seed(seed)
        self.x = torch.randn((n, *shape), generator=g)
        self.y = torch.randint(0, num_classes, (n,), generator=g)

    def __len__(self) -> int:
        return self.x.shape[0]

    def __getitem__(self, i):
        return self.x[i], self.y[i]


class SyntheticSegmentation(Dataset):
    """Random images with blob masks, returned as the reference's ``{'image': CHW float, 'mask': HW float}`` dicts."""

    def __init__(self, n: int, hw: Sequence[int] = (240, 320), seed: int = 0):
        g = torch.Generator().manual_seed(seed)
        h, w = hw
        self.x = torch.rand((n, 3, h, w), generator=g)
        yy, xx = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
        cy = torch.randint(0, h, (n,), generator=g)
        cx = torch.randint(0, w, (n,), generator=g)
        r = torch.randint(min(h, w) // 8, min(h, w) // 3, (n,), generator=g)
        self.m = (((yy[None] - cy[:, None, None]) ** 2 + (xx[None] - cx[:, None, None]) ** 2) < (r ** 2)[:, None, None]).float()

    def __len__(self) -> int:
        return self.x.shape[0]

    def __getitem__(self, i):
        return {"image": self.x[i], "mask": self.m[i]}


# ------------------------------------------------------------------------------------------------------------------
# File-backed segmentation data.  Output contract of ref:pytorch/unet/data_loading.py:52-134 — items are
# ``{'image': float32 CHW in [0, 1], 'mask': float32 HW in {0, 1}}``, images resized by ``scale`` with bicubic and
# masks with nearest-neighbour interpolation — built differently:
#   * both folders are listed ONCE into stem -> path tables (no glob per item);
#   * the palette (sorted set of distinct mask values, which fixes the value -> class-index mapping) is found by a pool
#     of workers, can be handed in ready-made, and `SegmentationFolder.shared` lets rank 0 alone scan the masks and
#     publish the result to the other ranks instead of every rank decoding every mask at start-up;
#   * the value -> index mapping is one vectorised table lookup per mask instead of a Python loop over values.
_IMAGE_EXTS_TORCH = (".pt", ".pth")


def _decode(path) -> np.ndarray:
    """File -> numpy array (H, W) or (H, W, C).  .npy / .pt / .pth hold arrays directly, anything else goes to PIL."""
    ext = os.path.splitext(str(path))[1].lower()
    if ext == ".npy":
        return np.load(path)
    if ext in _IMAGE_EXTS_TORCH:
        return torch.load(path).numpy()
    from PIL import Image

    with Image.open(path) as im:
        return np.asarray(im)


def _pack_values(a: np.ndarray) -> np.ndarray:
    """(H, W) stays as it is; (H, W, C<=4) colour masks are packed into one int64 per pixel so that a colour is a
    scalar that can be sorted and searched."""
    if a.ndim == 2:
        return a.astype(np.int64, copy=False)
    if a.ndim != 3 or a.shape[-1] > 4:
        raise ValueError(f"a mask must be a 2-D label image or an image with up to 4 colour planes, got shape {a.shape}")
    out = np.zeros(a.shape[:2], dtype=np.int64)
    for c in range(a.shape[-1]):
        out = (out << 8) | a[..., c].astype(np.int64)
    return out


def _distinct_values(path) -> np.ndarray:
    return np.unique(_pack_values(_decode(path)))


def scan_palette(mask_paths: Sequence, workers: Optional[int] = None) -> np.ndarray:
    """Sorted distinct (packed) pixel values over all mask files, decoded by ``workers`` processes in parallel
    (the reference uses a multiprocessing.Pool for this step, ref:pytorch/unet/data_loading.py:66-72)."""
    paths = [str(p) for p in mask_paths]
    if not paths:
        return np.zeros((0,), dtype=np.int64)
    if workers is None:
        workers = min(len(paths), os.cpu_count() or 1, 16)
    if workers <= 1 or len(paths) < 8:
        parts = [_distinct_values(p) for p in paths]
    else:
        import multiprocessing as mp

        with mp.get_context("spawn" if torch.cuda.is_initialized() else "fork").Pool(workers) as pool:
            parts = pool.map(_distinct_values, paths, chunksize=max(1, len(paths) // (4 * workers)))
    return np.unique(np.concatenate(parts))


def _list_by_stem(folder: Path) -> dict:
    table: dict = {}
    with os.scandir(folder) as it:
        for e in it:
            if e.is_file() and not e.name.startswith("."):
                table.setdefault(os.path.splitext(e.name)[0], []).append(e.path)
    return table


class SegmentationFolder(Dataset):
    def __init__(self, images_dir: str, mask_dir: str, scale: float = 1.0, mask_suffix: str = "",
                 palette: Optional[Sequence[int]] = None, workers: Optional[int] = None):
        if not 0.0 < scale <= 1.0:
            raise ValueError(f"scale must lie in (0, 1], got {scale}")
        self.images_dir, self.mask_dir = Path(images_dir), Path(mask_dir)
        self.scale, self.mask_suffix = float(scale), mask_suffix
        images, masks = _list_by_stem(self.images_dir), _list_by_stem(self.mask_dir)
        if not images:
            raise RuntimeError(f"{images_dir} holds no image files")
        self.ids = sorted(images)
        self._image_path, self._mask_path = {}, {}
        for stem in self.ids:
            if len(images[stem]) != 1:
                raise RuntimeError(f"image id '{stem}' is ambiguous: {sorted(images[stem])}")
            m = masks.get(stem + mask_suffix, [])
            if len(m) != 1:
                what = "has no mask" if not m else f"has several masks {sorted(m)}"
                raise FileNotFoundError(f"image id '{stem}' {what} in {mask_dir} (suffix '{mask_suffix}')")
            self._image_path[stem], self._mask_path[stem] = images[stem][0], m[0]
        if palette is None:
            palette = scan_palette([self._mask_path[s] for s in self.ids], workers)
        self.palette = np.asarray(sorted(int(v) for v in palette), dtype=np.int64)
        self.mask_values = self.palette.tolist()  # name used by the reference's training script for logging

    @classmethod
    def shared(cls, images_dir: str, mask_dir: str, scale: float = 1.0, mask_suffix: str = "", src: int = 0,
               workers: Optional[int] = None) -> "SegmentationFolder":
        """Collective constructor: only rank ``src`` decodes the masks; the palette travels to the other ranks through
        the process group (``dist.broadcast_object``).  Without an initialised group it is the plain constructor."""
        from ..parallel import dist

        if not dist.is_initialized() or dist.get_world_size() == 1:
            return cls(images_dir, mask_dir, scale, mask_suffix, workers=workers)
        palette = None
        if dist.get_rank() == src:
            palette = cls(images_dir, mask_dir, scale, mask_suffix, workers=workers).palette.tolist()
        palette = dist.broadcast_object(palette, src)
        return cls(images_dir, mask_dir, scale, mask_suffix, palette=palette)

    def __len__(self) -> int:
        return len(self.ids)

    def _resized(self, arr: np.ndarray, nearest: bool) -> np.ndarray:
        from PIL import Image

        h, w = arr.shape[:2]
        nw, nh = int(self.scale * w), int(self.scale * h)
        if nw < 1 or nh < 1:
            raise ValueError(f"scale {self.scale} leaves no pixels of a {w}x{h} image")
        if (nw, nh) == (w, h):
            return arr
        return np.asarray(Image.fromarray(arr).resize((nw, nh), resample=Image.NEAREST if nearest else Image.BICUBIC))

    def class_index_mask(self, mask: np.ndarray) -> np.ndarray:
        """Pixel values -> position in the palette (one table lookup for the whole mask)."""
        packed = _pack_values(mask)
        idx = np.searchsorted(self.palette, packed)
        idx = np.clip(idx, 0, max(len(self.palette) - 1, 0))
        if len(self.palette) and not np.array_equal(self.palette[idx], packed):
            raise ValueError("mask contains a value that is not in the palette (was it scanned from another folder?)")
        return idx

    def __getitem__(self, i: int):
        stem = self.ids[i]
        image, mask = _decode(self._image_path[stem]), _decode(self._mask_path[stem])
        if image.shape[:2] != mask.shape[:2]:
            raise ValueError(f"'{stem}': image is {image.shape[1]}x{image.shape[0]} but its mask is "
                             f"{mask.shape[1]}x{mask.shape[0]}")
        image = self._resized(image, nearest=False)
        classes = self.class_index_mask(self._resized(mask, nearest=True))
        chw = image[None] if image.ndim == 2 else np.moveaxis(image, -1, 0)
        x = torch.from_numpy(np.array(chw, dtype=np.float32, order="C"))
        if float(x.max()) > 1.0:   # 8-bit data -> [0, 1]; data already in [0, 1] stays
            x = x / 255.0
        return {"image": x, "mask": torch.from_numpy((classes > 0).astype(np.float32))}


# Names under which the reference's scripts construct the dataset (ref:pytorch/unet/train.py:85: CarvanaDataset(images,
# masks, scale) — no mask suffix).
BasicDataset = SegmentationFolder


class CarvanaDataset(SegmentationFolder):
    def __init__(self, images_dir, mask_dir, scale=1, mask_suffix="", **kw):
        super().__init__(images_dir, mask_dir, scale, mask_suffix=mask_suffix, **kw)
