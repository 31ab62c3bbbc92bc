"""Run one conv shape a few times (for ncu): python bench/one_conv.py H C K R STRIDE PAD [fprop|dgrad|wgrad]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeplearning_mpi_b200.ops import conv as C

h, c, k, R, st, pad = [int(v) for v in sys.argv[1:7]]
what = sys.argv[7] if len(sys.argv) > 7 else "fprop"
N = int(os.environ.get("BS", "256"))
dev = torch.device("cuda:0")
x = torch.randn(N, h, h, c, device=dev).bfloat16()
w = (torch.randn(k, c, R, R, device=dev) * (c * R * R) ** -0.5).bfloat16()
w2d = w.permute(0, 2, 3, 1).contiguous().view(k, -1)
wt2d = w.permute(1, 2, 3, 0).contiguous().view(c, -1)
oh = (h + 2 * pad - R) // st + 1
dy = torch.randn(N, oh, oh, k, device=dev).bfloat16()
dw = torch.zeros(k, R, R, c, device=dev)
ssum = torch.zeros(k, device=dev); ssq = torch.zeros(k, device=dev)
for _ in range(3):
    if what == "fprop":
        C.conv2d_fprop(x, w2d, R, R, st, pad, stats=(ssum, ssq))
    elif what == "dgrad":
        C.conv2d_dgrad(dy, wt2d, R, R, st, pad, (h, h))
    else:
        C.conv2d_wgrad(dy, x, dw, R, R, st, pad)
torch.cuda.synchronize()
print("ok")
