"""Fixed cost per kernel launch inside a CUDA graph: the same tiny op captured REPS times back to back, time per
replayed launch.  A ResNet-18 32x32 step is ~140 such launches, so this floor IS that benchmark; for ResNet-50 it is
~350 launches per 18 ms step.  Prints one line per op."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deeplearning_mpi_b200.ops import conv as C  # noqa: E402
from deeplearning_mpi_b200.ops import kernels as K  # noqa: E402

dev = torch.device("cuda:0")
REPS = int(os.environ.get("REPS", "200"))


def per_launch(fn, reps=REPS):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            for _ in range(reps):
                fn()
    for _ in range(2):
        g.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        g.replay()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b) * 1e3 / reps)
    return best


def main():
    torch.manual_seed(0)
    n, h, c, k = 8, 8, 64, 64
    x = torch.randn(n, h, h, c, device=dev).bfloat16()
    y = torch.empty(n, h, h, k, device=dev, dtype=torch.bfloat16)
    out = torch.empty_like(y)
    w1 = torch.randn(k, c, device=dev).bfloat16()
    w3 = torch.randn(k, 9 * c, device=dev).bfloat16()
    w256 = torch.randn(256, 256, device=dev).bfloat16()
    x256 = torch.randn(n, h, h, 256, device=dev).bfloat16()
    y256 = torch.empty(n, h, h, 256, device=dev, dtype=torch.bfloat16)
    dw = torch.zeros(k, 3, 3, c, device=dev)
    dw1 = torch.zeros(k, 1, 1, c, device=dev)
    ssum, ssq = torch.zeros(k, device=dev), torch.zeros(k, device=dev)
    gamma, beta = torch.ones(k, device=dev), torch.zeros(k, device=dev)
    mean, invstd = torch.zeros(k, device=dev), torch.ones(k, device=dev)
    rm, rv = torch.zeros(k, device=dev), torch.ones(k, device=dev)
    nbt = torch.zeros((), dtype=torch.int64, device=dev)
    acc, dg, db = torch.zeros(2 * k, device=dev), torch.zeros(k, device=dev), torch.zeros(k, device=dev)
    dy = torch.empty_like(y)
    small = torch.zeros(256, device=dev)
    ops = {
        "aten fill (256 floats)": lambda: small.fill_(1.0),
        "memset (our zero_)": lambda: K.zero_(small),
        "igemm 1x1 64->64 bn64 (4 tiles)": lambda: C.conv2d_fprop(x, w1, 1, 1, out=y),
        "igemm 1x1 + stats (det)": lambda: C.conv2d_fprop(x, w1, 1, 1, out=y, stats=(ssum, ssq)),
        "igemm 3x3 64->64 patch": lambda: C.conv2d_fprop(x, w3, 3, 3, 1, 1, out=y),
        "igemm 1x1 256->256 bn256 pair": lambda: C.conv2d_fprop(x256, w256, 1, 1, out=y256),
        "wgrad 3x3 64->64": lambda: C.conv2d_wgrad(y, x, dw, 3, 3, 1, 1),
        "wgrad 1x1 64->64": lambda: C.conv2d_wgrad(y, x, dw1, 1, 1, 1, 0),
        "bn_forward": lambda: K.bn_forward(y, None, out, ssum, ssq, gamma, beta, mean, invstd, rm, rv, nbt, 1e-5, 0.1,
                                           True, True),
        "bn_backward (2 launches)": lambda: K.bn_backward(out, y, None, dy, None, mean, invstd, gamma, acc, dg, db,
                                                          beta=beta, relu_from_y=True),
    }
    for name, fn in ops.items():
        print(f"{per_launch(fn):8.2f} us/launch  {name}", flush=True)


if __name__ == "__main__":
    main()
