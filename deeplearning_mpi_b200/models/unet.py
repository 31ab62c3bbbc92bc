"""UNet for binary / multi-class segmentation with the reference's module tree and parameter names
(ref:pytorch/unet/model.py:5-81 — DoubleConv / DownBlock / UpBlock / UNet; 136 state_dict keys for
``out_classes=1``), so ``torch.save(ddp_model.state_dict())`` checkpoints are interchangeable with
ref:pytorch/unet/train.py:72-74,216.

CUDA forward (sm_100a kernels, NHWC bf16):
  * every Conv3x3+bias -> BatchNorm -> ReLU is one implicit-GEMM launch (bias + BN statistics in the epilogue)
    plus one fused normalise+ReLU pass;
  * the skip "concat" is free: the encoder's DoubleConv writes its output into the upper channel slice of a
    pre-allocated buffer and the decoder's ConvTranspose2d (a GEMM with a pixel-shuffle scatter epilogue)
    writes the lower slice (the reference materialises ``torch.cat`` — model.py:47);
  * the 1-channel head is a bandwidth-bound dot-product kernel.
``up_sample_mode='bilinear'`` (nn.Upsample, align_corners=True, ref:pytorch/unet/model.py:39-40) runs a native gather
kernel pair (csrc/kernels/misc/upsample.cu) that also writes straight into the concat buffer; ``conv_transpose`` is the
benchmarked mode.

``python -m deeplearning_mpi_b200.models.unet`` prints the module tree and the output shape for a 512x512 input, the
smoke block of ref:pytorch/unet/model.py:84-89.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ..ops import functional as FN
from ..utils.arena import ensure_arena
from .resnet import _cba


class DoubleConv(nn.Module):
    def __init__(self, in_channels: int, out_channels: int):
        super().__init__()
        self.double_conv = nn.Sequential(
            nn.Conv2d(in_channels, out_channels, kernel_size=3, padding=1),
            nn.BatchNorm2d(out_channels),
            nn.ReLU(inplace=True),
            nn.Conv2d(out_channels, out_channels, kernel_size=3, padding=1),
            nn.BatchNorm2d(out_channels),
            nn.ReLU(inplace=True),
        )

    def forward(self, x):
        return self.double_conv(x)

    def forward_sm100(self, x, arena, out_holder=None, stem_hw=None):
        dc = self.double_conv
        a = _cba(x, dc[0], dc[1], arena, relu=True, stem_hw=stem_hw)
        return _cba(a, dc[3], dc[4], arena, relu=True, out_holder=out_holder)


class DownBlock(nn.Module):
    def __init__(self, in_channels: int, out_channels: int):
        super().__init__()
        self.double_conv = DoubleConv(in_channels, out_channels)
        self.down_sample = nn.MaxPool2d(2)

    def forward(self, x):
        skip_out = self.double_conv(x)
        return self.down_sample(skip_out), skip_out

    def forward_sm100(self, x, arena, skip_holder, stem_hw=None):
        skip = self.double_conv.forward_sm100(x, arena, out_holder=skip_holder, stem_hw=stem_hw)
        # the skip tensor has two consumers (pool, decoder concat): the concat's gradient slice is added inside the
        # pooling backward kernel (GradLink) instead of by autograd
        link = FN.GradLink() if (torch.is_grad_enabled() and skip.requires_grad) else None
        return FN.MaxPool.apply(skip, 2, 2, 0, link), skip, link


class UpBlock(nn.Module):
    def __init__(self, in_channels: int, out_channels: int, up_sample_mode: str):
        super().__init__()
        if up_sample_mode == "conv_transpose":
            self.up_sample = nn.ConvTranspose2d(in_channels - out_channels, in_channels - out_channels, kernel_size=2,
                                                stride=2)
        elif up_sample_mode == "bilinear":
            self.up_sample = nn.Upsample(scale_factor=2, mode="bilinear", align_corners=True)
        else:
            raise ValueError("Unsupported `up_sample_mode` (can take one of `conv_transpose` or `bilinear`)")
        self.double_conv = DoubleConv(in_channels, out_channels)

    def forward(self, down_input, skip_input):
        x = self.up_sample(down_input)
        x = torch.cat([x, skip_input], dim=1)
        return self.double_conv(x)

    def forward_sm100(self, down, skip, cat_buf, arena, link=None):
        c_up = down.shape[-1]
        if isinstance(self.up_sample, nn.ConvTranspose2d):
            up = FN.ConvTranspose2x2.apply(down, self.up_sample.weight, self.up_sample.bias, [cat_buf[..., :c_up]], arena)
        else:
            up = FN.UpsampleBilinear2x.apply(down, [cat_buf[..., :c_up]])
        x = FN.ConcatView.apply(up, skip, [cat_buf], link)
        return self.double_conv.forward_sm100(x, arena)


class UNet(nn.Module):
    def __init__(self, out_classes: int = 2, up_sample_mode: str = "conv_transpose"):
        super().__init__()
        self.up_sample_mode = up_sample_mode
        self.down_conv1 = DownBlock(3, 64)
        self.down_conv2 = DownBlock(64, 128)
        self.down_conv3 = DownBlock(128, 256)
        self.down_conv4 = DownBlock(256, 512)
        self.double_conv = DoubleConv(512, 1024)
        self.up_conv4 = UpBlock(512 + 1024, 512, self.up_sample_mode)
        self.up_conv3 = UpBlock(256 + 512, 256, self.up_sample_mode)
        self.up_conv2 = UpBlock(128 + 256, 128, self.up_sample_mode)
        self.up_conv1 = UpBlock(128 + 64, 64, self.up_sample_mode)
        self.conv_last = nn.Conv2d(64, out_classes, kernel_size=1)

    def forward(self, x):
        if x.is_cuda:
            return self._forward_sm100(x)
        x, skip1_out = self.down_conv1(x)
        x, skip2_out = self.down_conv2(x)
        x, skip3_out = self.down_conv3(x)
        x, skip4_out = self.down_conv4(x)
        x = self.double_conv(x)
        x = self.up_conv4(x, skip4_out)
        x = self.up_conv3(x, skip3_out)
        x = self.up_conv2(x, skip2_out)
        x = self.up_conv1(x, skip1_out)
        return self.conv_last(x)

    def _forward_sm100(self, x):
        arena = ensure_arena(self)
        arena.begin_forward()
        n, c, h, w = x.shape
        assert h % 16 == 0 and w % 16 == 0, "UNet needs H and W divisible by 16"
        from ..ops.conv import smallc_geometry

        _, _, hp, wp = smallc_geometry(h, w, 3, 3, 1, 1)
        xp = FN.StageInput.apply(x, 1, hp, wp)
        dev = x.device
        # concat buffers: [upsampled (C_up) | skip (C_skip)] per decoder level
        cat1 = torch.empty((n, h, w, 128 + 64), dtype=torch.bfloat16, device=dev)
        cat2 = torch.empty((n, h // 2, w // 2, 256 + 128), dtype=torch.bfloat16, device=dev)
        cat3 = torch.empty((n, h // 4, w // 4, 512 + 256), dtype=torch.bfloat16, device=dev)
        cat4 = torch.empty((n, h // 8, w // 8, 1024 + 512), dtype=torch.bfloat16, device=dev)
        a, s1, l1 = self.down_conv1.forward_sm100(xp, arena, [cat1[..., 128:]], stem_hw=(h, w))
        a, s2, l2 = self.down_conv2.forward_sm100(a, arena, [cat2[..., 256:]])
        a, s3, l3 = self.down_conv3.forward_sm100(a, arena, [cat3[..., 512:]])
        a, s4, l4 = self.down_conv4.forward_sm100(a, arena, [cat4[..., 1024:]])
        a = self.double_conv.forward_sm100(a, arena)
        a = self.up_conv4.forward_sm100(a, s4, cat4, arena, l4)
        a = self.up_conv3.forward_sm100(a, s3, cat3, arena, l3)
        a = self.up_conv2.forward_sm100(a, s2, cat2, arena, l2)
        a = self.up_conv1.forward_sm100(a, s1, cat1, arena, l1)
        return FN.Head1x1.apply(a, self.conv_last.weight, self.conv_last.bias, arena)


def _smoke(size: int = 512) -> None:
    """Module tree + output shape for one ``size`` x ``size`` RGB image (on the GPU through the sm_100a kernels when
    one is present, else the CPU oracle path)."""
    model = UNet(out_classes=1)
    print(model)
    dev = torch.device("cuda" if torch.cuda.is_available() else "cpu")
    model = model.to(dev).eval()
    with torch.no_grad():
        out = model(torch.randn(1, 3, size, size, device=dev))
    n_params = sum(p.numel() for p in model.parameters())
    print(f"input (1, 3, {size}, {size}) -> output {tuple(out.shape)} on {dev.type}; {n_params:,} parameters, "
          f"{len(model.state_dict())} state_dict keys")


if __name__ == "__main__":
    import sys

    _smoke(int(sys.argv[1]) if len(sys.argv) > 1 else 512)
