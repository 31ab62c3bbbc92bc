"""Loss modules and functional helpers with the names the reference scripts use
(``nn.CrossEntropyLoss`` ref:pytorch/resnet/main.py:113, ``nn.BCEWithLogitsLoss`` ref:pytorch/unet/train.py:162).
On CUDA fp32 logits they run the fused forward+backward sm_100a kernels; anywhere else they defer to PyTorch."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .ops import functional as FN
from .optim import clip_grad_norm_  # noqa: F401  (re-export: dmpi.nn.clip_grad_norm_)


def cross_entropy(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    if logits.is_cuda and logits.dtype == torch.float32 and logits.dim() == 2 and labels.dtype == torch.int64:
        return FN.SoftmaxCrossEntropy.apply(logits, labels)
    return F.cross_entropy(logits, labels)


def binary_cross_entropy_with_logits(logits: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
    if logits.is_cuda and logits.dtype == torch.float32 and logits.shape == targets.shape:
        return FN.BCEWithLogits.apply(logits, targets)
    return F.binary_cross_entropy_with_logits(logits, targets)


class CrossEntropyLoss(nn.Module):
    def forward(self, logits, labels):
        return cross_entropy(logits, labels)


class BCEWithLogitsLoss(nn.Module):
    def forward(self, logits, targets):
        return binary_cross_entropy_with_logits(logits, targets)
