"""Per-layer timing of the tcgen05 conv kernels vs cuDNN (bf16, channels_last, cudnn.benchmark = True as the reference
sets it, ref:pytorch/resnet/main.py:29) on every distinct ResNet-50 (MODEL=resnet50, batch 256) or UNet (MODEL=unet,
batch 16, 512x512) convolution shape.  Writes gpurun_out/conv_bench_<model>.txt.  CUDA-event timed, best of 5, L2
flushed between iterations."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from deeplearning_mpi_b200.ops import conv as C

torch.backends.cudnn.benchmark = True
dev = torch.device("cuda:0")
MODEL = os.environ.get("MODEL", "resnet50")
# 1 GiB: larger than L2, and long enough (~170 us) that the host has queued the timed launch before the flush
# retires — with a short flush the Python launch latency of the wrapper leaked into the event interval
flush = torch.empty(1 << 30, dtype=torch.uint8, device=dev)

def timeit(fn, iters=5):
    for _ in range(2): fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return min(ts)

N = int(os.environ.get("BS", "256" if MODEL == "resnet50" else "16"))
shapes = [  # h, c, k, R, stride, pad
    (56, 64, 64, 1, 1, 0), (56, 64, 64, 3, 1, 1), (56, 64, 256, 1, 1, 0), (56, 256, 64, 1, 1, 0),
    (56, 256, 128, 1, 1, 0), (56, 128, 128, 3, 2, 1), (28, 128, 512, 1, 1, 0), (28, 512, 128, 1, 1, 0),
    (28, 128, 128, 3, 1, 1), (56, 256, 512, 1, 2, 0), (28, 256, 256, 3, 2, 1), (14, 256, 1024, 1, 1, 0),
    (14, 1024, 256, 1, 1, 0), (14, 256, 256, 3, 1, 1), (14, 512, 512, 3, 2, 1), (7, 512, 2048, 1, 1, 0),
    (7, 2048, 512, 1, 1, 0), (7, 512, 512, 3, 1, 1),
]
if MODEL == "unet":  # ref:pytorch/unet/model.py:51-68 at 512x512: every 3x3 DoubleConv shape of encoder and decoder
    shapes = [(512, 64, 64, 3, 1, 1), (256, 64, 128, 3, 1, 1), (256, 128, 128, 3, 1, 1), (128, 128, 256, 3, 1, 1),
              (128, 256, 256, 3, 1, 1), (64, 256, 512, 3, 1, 1), (64, 512, 512, 3, 1, 1), (32, 512, 1024, 3, 1, 1),
              (32, 1024, 1024, 3, 1, 1),
              # decoder: the first conv of an UpBlock reads the [up | skip] concat buffer
              (64, 1536, 512, 3, 1, 1), (128, 768, 256, 3, 1, 1), (256, 384, 128, 3, 1, 1), (512, 192, 64, 3, 1, 1)]
if os.environ.get("SHAPES"):
    shapes = [shapes[int(i)] for i in os.environ["SHAPES"].split(",")]
skip_ref = os.environ.get("SKIP_CUDNN") == "1"
lines = []
for (h, c, k, R, st, pad) in shapes:
    x = torch.randn(N, h, h, c, device=dev).bfloat16()
    w = (torch.randn(k, c, R, R, device=dev) * (c*R*R) ** -0.5).bfloat16()
    w2d = w.permute(0, 2, 3, 1).contiguous().view(k, -1)
    wt2d = w.permute(1, 2, 3, 0).contiguous().view(c, -1)
    oh = (h + 2*pad - R)//st + 1
    dy = torch.randn(N, oh, oh, k, device=dev).bfloat16()
    dw = torch.zeros(k, R, R, c, device=dev)
    ssum = torch.zeros(k, device=dev); ssq = torch.zeros(k, device=dev)
    flops = 2.0 * N * oh * oh * k * c * R * R
    xc = x.permute(0, 3, 1, 2); wc = w.contiguous(memory_format=torch.channels_last); dyc = dy.permute(0, 3, 1, 2)
    res = {}
    res["fprop"] = timeit(lambda: C.conv2d_fprop(x, w2d, R, R, st, pad, stats=(ssum, ssq)))
    if os.environ.get("FPROP_NOSTATS") == "1":  # the same GEMM without the BatchNorm statistics in the epilogue
        res["fprop_nostat"] = timeit(lambda: C.conv2d_fprop(x, w2d, R, R, st, pad))
    res["dgrad"] = timeit(lambda: C.conv2d_dgrad(dy, wt2d, R, R, st, pad, (h, h)))
    res["wgrad"] = timeit(lambda: C.conv2d_wgrad(dy, x, dw, R, R, st, pad))
    if st == 1:
        add = torch.randn_like(x)
        res["dgrad_add"] = timeit(lambda: C.conv2d_dgrad(dy, wt2d, R, R, st, pad, (h, h), addend=add))
    if not skip_ref:
      res["cudnn_f"] = timeit(lambda: F.conv2d(xc, wc, stride=st, padding=pad))
      res["cudnn_d"] = timeit(lambda: torch.ops.aten.convolution_backward(dyc, xc, wc, None, (st, st), (pad, pad), (1, 1), False, (0, 0), 1, (True, False, False)))
      res["cudnn_w"] = timeit(lambda: torch.ops.aten.convolution_backward(dyc, xc, wc, None, (st, st), (pad, pad), (1, 1), False, (0, 0), 1, (False, True, False)))
    line = f"h{h} c{c} k{k} R{R} s{st}: " + " ".join(f"{n}={v*1e3:.0f}us({flops/v/1e9:.0f}TF)" for n, v in res.items())
    print(line, flush=True); lines.append(line)
os.makedirs("gpurun_out", exist_ok=True)
open(f"gpurun_out/conv_bench_{MODEL}.txt", "w").write("\n".join(lines) + "\n")
