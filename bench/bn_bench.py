"""BatchNorm kernel bandwidth on ResNet-50 shapes (bs256): achieved GB/s vs the measured HBM copy peak."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeplearning_mpi_b200.ops import kernels as K
dev = "cuda:0"
peak = 6551.0
try: peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception: pass
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timeit(fn, iters=5):
    fn(); ts = []
    for _ in range(iters):
        flush.zero_(); a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return min(ts) * 1e-3
for (M, C) in [(802816, 64), (802816, 256), (200704, 128), (200704, 512), (50176, 1024), (12544, 2048)]:
    y = torch.randn(M, C, device=dev).bfloat16(); res = torch.randn(M, C, device=dev).bfloat16()
    out = torch.empty_like(y); dout = torch.randn(M, C, device=dev).bfloat16(); dy = torch.empty_like(y); dres = torch.empty_like(y)
    g, b = torch.rand(C, device=dev) + .5, torch.randn(C, device=dev)
    ssum, ssq = torch.zeros(C, device=dev), torch.zeros(C, device=dev); K.channel_stats(y, ssum, ssq)
    mean, istd = torch.empty(C, device=dev), torch.empty(C, device=dev)
    rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev); acc = torch.zeros(2 * C, device=dev)
    dg, db = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
    nb = M * C * 2
    t_f = timeit(lambda: K.bn_forward(y, None, out, ssum, ssq, g, b, mean, istd, rm, rv, None, 1e-5, 0.1, True, True))
    t_fr = timeit(lambda: K.bn_forward(y, res, out, ssum, ssq, g, b, mean, istd, rm, rv, None, 1e-5, 0.1, True, True))
    t_b = timeit(lambda: (acc.zero_(), K.bn_backward(dout, y, out, dy, None, mean, istd, g, acc, dg, db)))
    t_by = timeit(lambda: (acc.zero_(), K.bn_backward(dout, y, None, dy, None, mean, istd, g, acc, dg, db, beta=b, relu_from_y=True)))
    t_br = timeit(lambda: (acc.zero_(), K.bn_backward(dout, y, out, dy, dres, mean, istd, g, acc, dg, db)))
    def gbs(passes, t): return passes * nb / t / 1e9
    print(f"M={M:7d} C={C:5d}: fwd {gbs(2,t_f):5.0f} GB/s ({gbs(2,t_f)/peak:.2f})  fwd+res {gbs(3,t_fr):5.0f} ({gbs(3,t_fr)/peak:.2f})  "
          f"bwd(act) {gbs(7,t_b):5.0f} ({gbs(7,t_b)/peak:.2f})  bwd(from_y) {gbs(5,t_by):5.0f} ({gbs(5,t_by)/peak:.2f}) [{t_by*1e6:.0f}us vs {t_b*1e6:.0f}us]  "
          f"bwd+dres {gbs(7,t_br):5.0f} ({gbs(7,t_br)/peak:.2f}) [{t_br*1e6:.0f}us]", flush=True)
