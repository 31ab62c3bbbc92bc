"""Evaluation metrics of the reference trainers."""
from __future__ import annotations

from typing import Iterable, List

import torch


@torch.no_grad()
def top1_accuracy(model, device, loader: Iterable) -> float:
    """correct / total over a loader of (images, labels)  (ref:pytorch/resnet/main.py:57-73)."""
    model.eval()
    correct = total = 0
    for data in loader:
        images, labels = data[0].to(device), data[1].to(device)
        outputs = model(images)
        predicted = outputs.argmax(dim=1)
        total += labels.size(0)
        correct += int((predicted == labels).sum().item())
    return correct / max(total, 1)


def dice_scores(pred_mask: torch.Tensor, true_mask: torch.Tensor, eps: float = 1e-8) -> List[float]:
    """Per-sample Dice of binary masks [B, H, W]; an empty prediction of an empty mask scores 1.0
    (ref:pytorch/unet/train.py:125-137)."""
    b = pred_mask.shape[0]
    p = pred_mask.reshape(b, -1).float()
    t = true_mask.reshape(b, -1).float()
    inter = (p * t).sum(1)
    union = p.sum(1) + t.sum(1)
    dice = torch.where(union > 0, (2 * inter + eps) / (union + eps), torch.ones_like(union))
    return dice.tolist()


@torch.no_grad()
def evaluate_dice(model, device, loader: Iterable) -> float:
    """Mean per-sample Dice at threshold 0.5 over a loader of {'image','mask'} batches (ref:pytorch/unet/train.py:104-140)."""
    model.eval()
    scores: List[float] = []
    for batch in loader:
        images = batch["image"].to(device, dtype=torch.float32)
        masks = batch["mask"].to(device, dtype=torch.float32)
        pred = model(images)
        if pred.dim() == 4:
            pred = pred.squeeze(1)
        pred = (torch.sigmoid(pred) > 0.5).float()
        scores.extend(dice_scores(pred, masks))
    return sum(scores) / len(scores) if scores else 0.0
