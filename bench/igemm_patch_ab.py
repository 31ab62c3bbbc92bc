"""A/B of patch-mode vs per-tap implicit GEMM (3x3 stride-1 fprop): python bench/igemm_patch_ab.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeplearning_mpi_b200.ops import conv as C

dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def timeit(fn, iters=4):
    fn(); ts = []
    for _ in range(iters):
        flush.zero_(); a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return min(ts) * 1e3

shapes = [(16, 512, 64, 64), (16, 512, 128, 64), (16, 512, 192, 64), (16, 256, 64, 128), (16, 256, 128, 128),
          (16, 256, 384, 128), (16, 128, 128, 256), (16, 128, 256, 256), (256, 56, 64, 64), (256, 28, 128, 128)]
for (n, h, c, k) in shapes:
    x = torch.randn(n, h, h, c, device=dev).bfloat16()
    w2d = (torch.randn(k, 9 * c, device=dev) * (9 * c) ** -0.5).bfloat16()
    ssum = torch.zeros(k, device=dev); ssq = torch.zeros(k, device=dev)
    res = []
    for mode in (0, 1):  # 0: per-tap loads; 1: the planner's patch rules (resident filter for N <= 64, CTA pairs above)
        C._PATCH = mode
        res.append(timeit(lambda: C.conv2d_fprop(x, w2d, 3, 3, 1, 1, stats=(ssum, ssq))))
    fl = 2.0 * n * h * h * k * c * 9
    print(f"n{n} h{h} c{c} k{k}: taps {res[0]:.0f} us ({fl/res[0]/1e6:.0f} TF)  patch {res[1]:.0f} us ({fl/res[1]/1e6:.0f} TF)", flush=True)
