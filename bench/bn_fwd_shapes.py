"""BatchNorm forward on the ResNet-50 shapes (batch 256), with and without residual, L2 flushed in between; run under
`ncu --metrics gpu__time_duration.sum -k regex:bn_fwd --csv` and summarise with --summarize."""
import os, sys, csv
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
SHAPES = [(802816, 64), (802816, 256), (200704, 128), (200704, 512), (50176, 256), (50176, 1024), (12544, 512), (12544, 2048)]


def run():
    import torch
    from deeplearning_mpi_b200.ops import kernels as K
    dev = "cuda:0"
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for (M, C) in SHAPES:
        y = torch.randn(M, C, device=dev).bfloat16(); res = torch.randn(M, C, device=dev).bfloat16(); out = torch.empty_like(y)
        g, b = torch.rand(C, device=dev) + .5, torch.randn(C, device=dev)
        ssum, ssq = torch.zeros(C, device=dev), torch.zeros(C, device=dev); K.channel_stats(y, ssum, ssq)
        mean, istd = torch.empty(C, device=dev), torch.empty(C, device=dev)
        rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
        mask = torch.empty(M, C // 8, dtype=torch.uint8, device=dev)
        for _ in range(3):
            flush.zero_()
            K.bn_forward(y, None, out, ssum, ssq, g, b, mean, istd, rm, rv, None, 1e-5, 0.1, True, True)
            flush.zero_()
            K.bn_forward(y, res, out, ssum, ssq, g, b, mean, istd, rm, rv, None, 1e-5, 0.1, True, True, relu_mask=mask)
        torch.cuda.synchronize()


def summarize(path):
    rows = list(csv.reader(open(path, errors="ignore")))
    h = next(r for r in rows if "Kernel Name" in r)
    ki, vi = h.index("Kernel Name"), h.index("Metric Value")
    d = [float(r[vi].replace(",", "")) / 1e3 for r in rows[rows.index(h) + 2:] if len(r) > vi and "bn_fwd" in r[ki]]
    if len(d) % 2: d.insert(0, float("nan"))  # ncu drops the first matching launch of a process
    tot = 0.0
    for si, (M, C) in enumerate(SHAPES):
        t0, t1 = d[si * 6 + 4], d[si * 6 + 5]  # third repetition
        nb = M * C * 2 / 1e6
        print(f"M={M:7d} C={C:5d}: plain {t0:5.0f}us {2 * nb / t0:5.2f}TB/s   +res+mask {t1:5.0f}us {(3 + 1 / 16) * nb / t1:5.2f}TB/s")
        tot += t0 + t1
    print(f"total {tot:.0f} us")


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--summarize":
        summarize(sys.argv[2])
    else:
        run()
