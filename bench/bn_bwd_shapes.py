"""BatchNorm backward on the ResNet-50 shapes (batch 256), one call per (shape, gate mode), L2 flushed in between.
Meant to run under `ncu --metrics gpu__time_duration.sum -k regex:bn_bwd --csv` so that the reduce and the apply
pass are timed separately; `python bench/bn_bwd_shapes.py --summarize file.csv` prints GB/s per launch."""
import os, sys, csv
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
SHAPES = [(802816, 64), (802816, 256), (200704, 128), (200704, 512), (50176, 256), (50176, 1024), (12544, 512), (12544, 2048)]
MODES = ["from_y", "bits+dres", "none"]
# tensor passes (reads + writes of M*C*2 bytes) of the reduce / apply pass per mode (bit mask = 1/16 pass)
PASSES = {"from_y": (2, 3), "bits+dres": (3 + 1 / 16, 3), "none": (2, 3)}


def run():
    import torch
    from deeplearning_mpi_b200.ops import kernels as K
    dev = "cuda:0"
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for (M, C) in SHAPES:
        y = torch.randn(M, C, device=dev).bfloat16(); res = torch.randn(M, C, device=dev).bfloat16()
        out = torch.empty_like(y); dout = torch.randn(M, C, device=dev).bfloat16(); dy = torch.empty_like(y); dres = torch.empty_like(y)
        g, b = torch.rand(C, device=dev) + .5, torch.randn(C, device=dev)
        ssum, ssq = torch.zeros(C, device=dev), torch.zeros(C, device=dev); K.channel_stats(y, ssum, ssq)
        mean, istd = torch.empty(C, device=dev), torch.empty(C, device=dev)
        rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev); acc = torch.zeros(2 * C, device=dev)
        dg, db = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
        mask = torch.empty(M, C // 8, dtype=torch.uint8, device=dev)
        K.bn_forward(y, res, out, ssum, ssq, g, b, mean, istd, rm, rv, None, 1e-5, 0.1, True, True, relu_mask=mask)
        for _ in range(2):
            flush.zero_(); acc.zero_()
            K.bn_backward(dout, y, None, dy, None, mean, istd, g, acc, dg, db, beta=b, relu_from_y=True)
            flush.zero_(); acc.zero_()
            K.bn_backward(dout, y, out, dy, dres, mean, istd, g, acc, dg, db, relu_mask=mask)
            flush.zero_(); acc.zero_()
            K.bn_backward(dout, y, None, dy, None, mean, istd, g, acc, dg, db)
        torch.cuda.synchronize()


def summarize(path):
    rows = list(csv.reader(open(path, errors="ignore")))
    h = next(r for r in rows if "Kernel Name" in r)
    ki, vi = h.index("Kernel Name"), h.index("Metric Value")
    d = [(r[ki], float(r[vi].replace(",", "")) / 1e3) for r in rows[rows.index(h) + 2:] if len(r) > vi]
    d = [x for x in d if "bn_bwd" in x[0]]
    if d and "apply" in d[0][0]:
        d.insert(0, ("bn_bwd_reduce (not captured)", float("nan")))  # ncu drops the first matching launch of a process
    per_shape = len(MODES) * 2 * 2
    tot_r = tot_a = 0.0
    for si, (M, C) in enumerate(SHAPES):
        chunk = d[si * per_shape + per_shape // 2:(si + 1) * per_shape]  # second repetition
        line = f"M={M:7d} C={C:5d}:"
        for mi, mode in enumerate(MODES):
            (_, tr), (_, ta) = chunk[2 * mi], chunk[2 * mi + 1]
            pr, pa = PASSES[mode]
            nb = M * C * 2 / 1e3  # KB -> us * GB/s
            line += f"  {mode}: red {tr:5.0f}us {pr * nb / tr / 1e3:5.2f}TB/s  app {ta:5.0f}us {pa * nb / ta / 1e3:5.2f}TB/s |"
            tot_r += tr; tot_a += ta
        print(line)
    print(f"total reduce {tot_r:.0f} us, apply {tot_a:.0f} us")


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--summarize":
        summarize(sys.argv[2])
    else:
        run()
