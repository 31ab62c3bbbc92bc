"""Runs the ResNet stem max-pool forward/backward a few times (for ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeplearning_mpi_b200.ops import kernels as K
dev = torch.device("cuda:0")
x = torch.randn(256, 112, 112, 64, device=dev).bfloat16()
for _ in range(3):
    y, idx = K.maxpool_forward(x, 3, 2, 1, want_idx=True)
    dy = torch.randn_like(y)
    dx = K.maxpool_backward(dy, idx, tuple(x.shape), 3, 2, 1)
torch.cuda.synchronize()
print("ok")
