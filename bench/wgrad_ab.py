"""A/B of the two 3x3 weight-gradient kernels on given shapes: python bench/wgrad_ab.py  (prints us per call)."""
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

shapes = [(16, 512, 192, 64), (16, 512, 128, 64), (16, 256, 64, 128), (16, 256, 128, 128), (16, 256, 384, 128),
          (256, 28, 128, 128), (256, 56, 64, 128), (256, 56, 128, 64), (16, 128, 128, 256)]
for (n, h, c, k) in shapes:
    x = torch.randn(n, h, h, c, device=dev).bfloat16()
    dy = torch.randn(n, h, h, k, device=dev).bfloat16()
    dw = torch.zeros(k, 3, 3, c, device=dev)
    res = []
    for mode in (0, 2):
        C._WGRAD_PATCH = mode
        res.append(timeit(lambda: C.conv2d_wgrad(dy, x, dw, 3, 3, 1, 1)))
    print(f"n{n} h{h} c{c} k{k}: taps {res[0]:.0f} us  patch {res[1]:.0f} us", flush=True)
