"""BatchNorm backward without caller scratch (``dmpi_bn_backward_noalloc``).

Who needs it: programs that cannot obtain device memory at the moment they call us — the whole GPU was handed to the
host application's own pool at start-up, or the call sits on a latency-critical path where cudaMalloc/cudaFree would
stall every stream, or the code runs as a plug-in that may not touch the host framework's allocator.  For them the
2*C floats of reduction scratch live in a fixed 16 KiB array inside the library image (limit: C <= 2048).

This test drives both entry points through the plain C ABI (ctypes on libdmpi_cuda.so, raw pointers, no wrapper) and
checks that the no-allocation variant reproduces ``dmpi_bn_backward`` (which gets zeroed scratch from the caller):
two calls in a row with different shapes — the second one reuses the shared array the first call left dirty — plus the
size-limit rejection.  The file name sorts last so a failure here cannot hide results of the other tests.
"""
import ctypes as C
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu

VP, I, I64 = C.c_void_p, C.c_int, C.c_int64
LIB = Path(__file__).resolve().parent.parent / "deeplearning_mpi_b200" / "_lib" / "libdmpi_cuda.so"


def _load():
    lib = C.CDLL(str(LIB))
    lib.dmpi_bn_backward.restype = I
    lib.dmpi_bn_backward.argtypes = [VP, VP, VP, VP, VP, I64, I, I64, I64, I64, I64, I64, VP, VP, VP, VP, VP, VP, VP]
    lib.dmpi_bn_backward_noalloc.restype = I
    lib.dmpi_bn_backward_noalloc.argtypes = [VP, VP, VP, VP, VP, I64, I, I64, I64, I64, I64, I64, VP, VP, VP, VP, VP, VP]
    return lib


def _case(lib, dev, m, c, with_res, seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    y = (torch.randn(m, c, device=dev, generator=g) * 1.5 + 0.3).bfloat16()
    act = torch.relu(torch.randn(m, c, device=dev, generator=g)).bfloat16()
    dout = torch.randn(m, c, device=dev, generator=g).bfloat16()
    yf = y.float()
    mean = yf.mean(0).contiguous()
    invstd = (yf.var(0, unbiased=False) + 1e-5).rsqrt().contiguous()
    gamma = (torch.rand(c, device=dev, generator=g) + 0.5).contiguous()
    stream = torch.cuda.current_stream().cuda_stream

    def run(noalloc):
        dy = torch.empty_like(y)
        dres = torch.empty_like(y) if with_res else None
        dgamma = torch.zeros(c, device=dev)
        dbeta = torch.zeros(c, device=dev)
        p = lambda t: t.data_ptr() if t is not None else None  # noqa: E731
        if noalloc:
            rc = lib.dmpi_bn_backward_noalloc(p(dout), p(y), p(act), p(dy), p(dres), m, c, c, c, c, c, c, p(mean),
                                              p(invstd), p(gamma), p(dgamma), p(dbeta), stream)
        else:
            acc = torch.zeros(2 * c, device=dev)
            rc = lib.dmpi_bn_backward(p(dout), p(y), p(act), p(dy), p(dres), m, c, c, c, c, c, c, p(mean), p(invstd),
                                      p(gamma), p(acc), p(dgamma), p(dbeta), stream)
        assert rc == 0, f"entry point returned {rc}"
        torch.cuda.synchronize()
        return dy, dres, dgamma, dbeta

    ref = run(False)
    got = run(True)
    for name, a, b in zip(("dy", "dres", "dgamma", "dbeta"), got, ref):
        if a is None:
            continue
        a, b = a.float(), b.float()
        err = (a - b).abs().max().item()
        den = b.abs().max().item() + 1e-12
        print(f"[noalloc m={m} c={c}] {name}: max_abs_err={err:.3e} rel={err / den:.3e}")
        # same tolerance as tests/test_kernels_gpu.py uses for BatchNorm backward (atomics reorder the sums)
        assert err <= 1e-2 * den, name


def test_bn_backward_noalloc_matches_scratch_variant_twice():
    lib = _load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    _case(lib, dev, 4096, 256, True, 1)    # first call on the device: resolves the static array's address
    _case(lib, dev, 1568, 2048, False, 2)  # second call: other shape, reuses the (dirty) shared array, C at the limit


def test_bn_backward_noalloc_rejects_oversize():
    lib = _load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    c = 2048 + 64
    t = torch.zeros(8, c, device=dev, dtype=torch.bfloat16)
    v = torch.ones(c, device=dev)
    rc = lib.dmpi_bn_backward_noalloc(t.data_ptr(), t.data_ptr(), None, t.data_ptr(), None, 8, c, c, c, c, c, c,
                                      v.data_ptr(), v.data_ptr(), v.data_ptr(), v.data_ptr(), v.data_ptr(),
                                      torch.cuda.current_stream().cuda_stream)
    assert rc == -4
