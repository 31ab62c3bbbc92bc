"""Probe: tcgen05 descriptors shifted by whole 128-byte rows inside a swizzle atom (see csrc/kernels/conv/umma_probe.cu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeplearning_mpi_b200 import _native as N

dev = torch.device("cuda:0")
torch.manual_seed(0)
rows = 192
x = torch.randn(rows, 64, device=dev).bfloat16()
w = torch.randn(64, 64, device=dev).bfloat16()
y = torch.randn(64, 128, device=dev).bfloat16()
lib = N.cuda()
for mode in (0, 1, 2, 3):
    res = []
    for shift in (0, 1, 2, 3, 5, 7, 8, 9, 17, 58, 59, 60):
        out = torch.zeros(128, 64, device=dev)
        src = y if (mode & 2) else w
        rc = lib.dmpi_umma_probe(x.data_ptr(), src.data_ptr(), out.data_ptr(), rows, shift, mode, N.stream_ptr())
        torch.cuda.synchronize()
        if mode & 2:
            ref = y.float().t() @ x[shift:shift + 64].float()
        else:
            ref = x[shift:shift + 128].float() @ w.float().t()
        err = (out - ref).abs().max().item() / (ref.abs().max().item() + 1e-6)
        res.append(f"s{shift}:{'ok' if err < 1e-2 else 'BAD(%.2f)' % err}")
    print(f"mode={mode} ({'MN' if mode & 2 else 'K'}-major, base_offset={'phase' if mode & 1 else '0'}): " + " ".join(res), flush=True)
