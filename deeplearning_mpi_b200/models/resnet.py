"""ResNet family (18/34/50/101/152) with torchvision-identical module and parameter names, so checkpoints
round-trip with the reference trainer (``torchvision.models.resnet18`` + ``fc -> Linear(512, 10)``,
ref:pytorch/resnet/main.py:40-41,48-52,139; state_dict has the same 122 keys for ResNet-18/10 classes).

The ``nn.Conv2d`` / ``nn.BatchNorm2d`` / ``nn.Linear`` children are *parameter holders*.  On a CUDA device
the forward pass runs entirely on this framework's sm_100a kernels (NHWC bf16 activations, tcgen05
implicit-GEMM convolutions, fused BatchNorm+ReLU(+residual), fp32 master weights); on CPU the very same
modules are executed by stock PyTorch, which is the fp32 oracle used by the tests and the path used by the
gloo/CPU configuration.
"""
from __future__ import annotations

from typing import List, Optional, Type, Union

import torch
import torch.nn as nn

from ..ops import functional as FN
from ..utils.arena import ensure_arena


def _cba(x, conv, bn, arena, relu=True, residual=None, out_holder=None, stem_hw=None, link=None, link_role=None):
    return FN.ConvBnAct.apply(x, conv.weight, conv.bias, bn.weight, bn.bias, residual, conv, bn, arena, relu,
                              out_holder, stem_hw, link, link_role)


def _fork_link(x, first_conv):
    """A GradLink for the fork at a block input, or None when the sum has to stay with autograd (x needs no gradient,
    or the main branch starts with a strided convolution whose dgrad is parity-decomposed)."""
    if not (torch.is_grad_enabled() and x.requires_grad) or first_conv.stride[0] != 1:
        return None
    return FN.GradLink()


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes: int, planes: int, stride: int = 1, downsample: Optional[nn.Module] = None):
        super().__init__()
        self.conv1 = nn.Conv2d(inplanes, planes, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = nn.Conv2d(planes, planes, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):  # stock PyTorch path (CPU oracle)
        identity = x
        out = self.relu(self.bn1(self.conv1(x)))
        out = self.bn2(self.conv2(out))
        if self.downsample is not None:
            identity = self.downsample(x)
        return self.relu(out + identity)

    def forward_sm100(self, x, arena):
        # the identity-branch gradient is added inside conv1's dgrad epilogue (GradLink) instead of by autograd
        link = _fork_link(x, self.conv1)
        take = "take" if link is not None else None
        out = _cba(x, self.conv1, self.bn1, arena, relu=True, link=link, link_role=take)
        if self.downsample is not None:
            identity = _cba(x, self.downsample[0], self.downsample[1], arena, relu=False, link=link,
                            link_role="give_dx" if link is not None else None)
            return _cba(out, self.conv2, self.bn2, arena, relu=True, residual=identity)
        return _cba(out, self.conv2, self.bn2, arena, relu=True, residual=x, link=link,
                    link_role="give_dres" if link is not None else None)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes: int, planes: int, stride: int = 1, downsample: Optional[nn.Module] = None):
        super().__init__()
        self.conv1 = nn.Conv2d(inplanes, planes, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, stride, 1, bias=False)  # stride on the 3x3 (ResNet v1.5)
        self.bn2 = nn.BatchNorm2d(planes)
        self.conv3 = nn.Conv2d(planes, planes * 4, 1, bias=False)
        self.bn3 = nn.BatchNorm2d(planes * 4)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        identity = x
        out = self.relu(self.bn1(self.conv1(x)))
        out = self.relu(self.bn2(self.conv2(out)))
        out = self.bn3(self.conv3(out))
        if self.downsample is not None:
            identity = self.downsample(x)
        return self.relu(out + identity)

    def forward_sm100(self, x, arena):
        link = _fork_link(x, self.conv1)
        take = "take" if link is not None else None
        out = _cba(x, self.conv1, self.bn1, arena, relu=True, link=link, link_role=take)
        out = _cba(out, self.conv2, self.bn2, arena, relu=True)
        if self.downsample is not None:
            identity = _cba(x, self.downsample[0], self.downsample[1], arena, relu=False, link=link,
                            link_role="give_dx" if link is not None else None)
            return _cba(out, self.conv3, self.bn3, arena, relu=True, residual=identity)
        return _cba(out, self.conv3, self.bn3, arena, relu=True, residual=x, link=link,
                    link_role="give_dres" if link is not None else None)


class ResNet(nn.Module):
    def __init__(self, block: Type[Union[BasicBlock, Bottleneck]], layers: List[int], num_classes: int = 1000):
        super().__init__()
        self.inplanes = 64
        self.conv1 = nn.Conv2d(3, 64, 7, 2, 3, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = nn.MaxPool2d(3, 2, 1)
        self.layer1 = self._make_layer(block, 64, layers[0])
        self.layer2 = self._make_layer(block, 128, layers[1], stride=2)
        self.layer3 = self._make_layer(block, 256, layers[2], stride=2)
        self.layer4 = self._make_layer(block, 512, layers[3], stride=2)
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = nn.Linear(512 * block.expansion, num_classes)
        for m in self.modules():  # same initialisation as torchvision
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)

    def _make_layer(self, block, planes, blocks, stride=1):
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(nn.Conv2d(self.inplanes, planes * block.expansion, 1, stride, bias=False),
                                       nn.BatchNorm2d(planes * block.expansion))
        layers = [block(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * block.expansion
        for _ in range(1, blocks):
            layers.append(block(self.inplanes, planes))
        return nn.Sequential(*layers)

    def __setattr__(self, name, value):
        # Replacing the classifier after construction (model.fc = nn.Linear(512, 10), ref:pytorch/resnet/main.py:41)
        # must invalidate a previously built arena.
        if name == "fc" and "_dmpi_arena" in self.__dict__:
            del self.__dict__["_dmpi_arena"]
        super().__setattr__(name, value)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda:
            return self._forward_sm100(x)
        x = self.maxpool(self.relu(self.bn1(self.conv1(x))))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        return self.fc(torch.flatten(self.avgpool(x), 1))

    def _forward_sm100(self, x: torch.Tensor) -> torch.Tensor:
        arena = ensure_arena(self)
        arena.begin_forward()
        n, c, h, w = x.shape
        from ..ops.conv import smallc_geometry

        _, _, hp, wp = smallc_geometry(h, w, 7, 7, 2, 3)
        xp = FN.StageInput.apply(x, 3, hp, wp)
        a = _cba(xp, self.conv1, self.bn1, arena, relu=True, stem_hw=(h, w))
        a = FN.MaxPool.apply(a, 3, 2, 1)
        for layer in (self.layer1, self.layer2, self.layer3, self.layer4):
            for blk in layer:
                a = blk.forward_sm100(a, arena)
        a = FN.GlobalAvgPool.apply(a)
        return FN.LinearFn.apply(a, self.fc.weight, self.fc.bias, arena)


def resnet18(num_classes: int = 1000, **_ignored) -> ResNet:
    return ResNet(BasicBlock, [2, 2, 2, 2], num_classes)


def resnet34(num_classes: int = 1000, **_ignored) -> ResNet:
    return ResNet(BasicBlock, [3, 4, 6, 3], num_classes)


def resnet50(num_classes: int = 1000, **_ignored) -> ResNet:
    return ResNet(Bottleneck, [3, 4, 6, 3], num_classes)


def resnet101(num_classes: int = 1000, **_ignored) -> ResNet:
    return ResNet(Bottleneck, [3, 4, 23, 3], num_classes)


def resnet152(num_classes: int = 1000, **_ignored) -> ResNet:
    return ResNet(Bottleneck, [3, 8, 36, 3], num_classes)
