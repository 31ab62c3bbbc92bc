"""Peak-rate probe of the igemm kernel on plain GEMM shapes (1x1 conv, flat mode) vs cuBLAS."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeplearning_mpi_b200.ops import conv as C
dev = "cuda:0"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timeit(fn, iters=5):
    fn(); fn(); ts = []
    for _ in range(iters):
        flush.zero_(); a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return min(ts) * 1e-3
for (M, K, N) in [(16384, 4096, 4096), (16384, 4096, 256), (65536, 1024, 128), (65536, 1024, 64), (262144, 256, 256), (262144, 64, 64), (32768, 8192, 512)]:
    x = torch.randn(1, 1, M, K, device=dev).bfloat16(); w = (torch.randn(N, K, device=dev) * K ** -0.5).bfloat16()
    t = timeit(lambda: C.conv2d_fprop(x, w, 1, 1))
    x2 = x.view(M, K); t2 = timeit(lambda: x2 @ w.t())
    fl = 2.0 * M * K * N; by = 2.0 * (M * K + N * K + M * N)
    print(f"M={M:6d} K={K:5d} N={N:5d}: ours {t*1e6:8.1f} us {fl/t/1e12:7.1f} TF {by/t/1e9:6.0f} GB/s | cublas {t2*1e6:8.1f} us {fl/t2/1e12:7.1f} TF", flush=True)
