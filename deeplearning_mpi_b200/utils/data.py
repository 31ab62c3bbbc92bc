"""Input pipeline helpers.

* :class:`CudaPrefetcher` — pinned-memory, side-stream host->device staging with one batch of look-ahead.  This is the
  role ``pin_memory=True`` + ``.to(device)`` + DDP's side-stream input copy play for the reference
  (ref:pytorch/resnet/main.py:96-102,126), except that the copy of batch i+1 overlaps the compute
  of batch i instead of serialising with it.
* :class:`SyntheticImages` / :class:`SyntheticSegmentation` — datasets of the reference's shapes for machines without
  the CIFAR-10 / microscopy data (there is no network in the benchmark environment).
* :class:`BasicDataset` / :class:`CarvanaDataset` — file-backed segmentation dataset with the reference's
  preprocessing contract (ref:pytorch/unet/data_loading.py:52-134).
"""
from __future__ import annotations

from os import listdir
from os.path import isfile, join, splitext
from pathlib import Path
from typing import Iterable, Iterator, Sequence

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


def _load_image(path):
    from PIL import Image

    ext = splitext(str(path))[1]
    if ext == ".npy":
        return Image.fromarray(np.load(path))
    if ext in (".pt", ".pth"):
        return Image.fromarray(torch.load(path).numpy())
    return Image.open(path)


class BasicDataset(Dataset):
    """Image/mask folders -> ``{'image': float CHW in [0,1], 'mask': float HW in {0,1}}``.

    Same contract as ref:pytorch/unet/data_loading.py:52-129: ids from the image folder, masks matched by stem
    (+ ``mask_suffix``), both resized by ``scale`` (bicubic for images, nearest for masks), images scaled to [0, 1]
    when they contain values above 1, masks mapped through the sorted set of values found in the mask folder and then
    binarised (``> 0``)."""

    def __init__(self, images_dir: str, mask_dir: str, scale: float = 1.0, mask_suffix: str = ""):
        assert 0 < scale <= 1, "Scale must be between 0 and 1"
        self.images_dir, self.mask_dir = Path(images_dir), Path(mask_dir)
        self.scale, self.mask_suffix = scale, mask_suffix
        self.ids = sorted(splitext(f)[0] for f in listdir(images_dir) if isfile(join(images_dir, f)) and not f.startswith("."))
        if not self.ids:
            raise RuntimeError(f"No input file found in {images_dir}, make sure you put your images there")
        values = set()
        for idx in self.ids:
            files = list(self.mask_dir.glob(idx + self.mask_suffix + ".*"))
            if not files:
                raise FileNotFoundError(f"No mask file found for index '{idx}' with suffix '{mask_suffix}' in '{mask_dir}'")
            m = np.asarray(_load_image(files[0]))
            if m.ndim == 2:
                values.update(np.unique(m).tolist())
            elif m.ndim == 3:
                values.update(map(tuple, np.unique(m.reshape(-1, m.shape[-1]), axis=0).tolist()))
            else:
                raise ValueError(f"Loaded masks should have 2 or 3 dimensions, found {m.ndim}")
        self.mask_values = sorted(values)

    def __len__(self) -> int:
        return len(self.ids)

    @staticmethod
    def preprocess(mask_values, pil_img, scale: float, is_mask: bool):
        from PIL import Image

        w, h = pil_img.size
        nw, nh = int(scale * w), int(scale * h)
        assert nw > 0 and nh > 0, "Scale is too small, resized images would have no pixel"
        img = np.asarray(pil_img.resize((nw, nh), resample=Image.NEAREST if is_mask else Image.BICUBIC))
        if is_mask:
            mask = np.zeros((nh, nw), dtype=np.int64)
            for i, v in enumerate(mask_values):
                if img.ndim == 2:
                    mask[img == v] = i
                else:
                    mask[(img == np.asarray(v)).all(-1)] = i
            return mask
        img = img[np.newaxis, ...] if img.ndim == 2 else img.transpose((2, 0, 1))
        if (img > 1).any():
            img = img / 255.0
        return img

    def __getitem__(self, idx):
        name = self.ids[idx]
        mask_file = list(self.mask_dir.glob(name + self.mask_suffix + ".*"))
        img_file = list(self.images_dir.glob(name + ".*"))
        assert len(img_file) == 1, f"Either no image or multiple images found for the ID {name}: {img_file}"
        assert len(mask_file) == 1, f"Either no mask or multiple masks found for the ID {name}: {mask_file}"
        mask, img = _load_image(mask_file[0]), _load_image(img_file[0])
        assert img.size == mask.size, f"Image and mask {name} should be the same size, but are {img.size} and {mask.size}"
        img = self.preprocess(self.mask_values, img, self.scale, is_mask=False)
        mask = self.preprocess(self.mask_values, mask, self.scale, is_mask=True)
        return {"image": torch.as_tensor(np.ascontiguousarray(img)).float().contiguous(),
                "mask": torch.as_tensor((mask > 0).astype(np.float32)).contiguous()}


class CarvanaDataset(BasicDataset):
    def __init__(self, images_dir, mask_dir, scale=1):
        super().__init__(images_dir, mask_dir, scale, mask_suffix="")
