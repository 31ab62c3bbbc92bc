"""In-tree native build: nvcc (sm_100a) for the CUDA kernels, g++ for the host runtime.

Produces two shared libraries next to the Python package (they travel to the GPU box with the
source snapshot; they are git-ignored):

  deeplearning_mpi_b200/_lib/libdmpi_host.so   TCP store + reducer core       (no CUDA dependency)
  deeplearning_mpi_b200/_lib/libdmpi_cuda.so   all sm_100a kernels + P2P arena + CUDA reducer glue
  deeplearning_mpi_b200/_lib/dmpi_torch_hooks.so   C++ autograd hooks of the DDP wrapper (pybind11 module, links libtorch)

Objects are rebuilt only when the content hash of the translation unit, its local headers and the
flags changes.  Usage:  python build.py [--force] [--verbose]
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent
LIB_DIR = ROOT / "deeplearning_mpi_b200" / "_lib"
OBJ_DIR = ROOT / "build" / "obj"
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
# The system g++ links libstdc++ dynamically.  (Do NOT honour $CXX blindly: a toolchain that links libstdc++ statically
# makes our .so export a private copy of std:: symbols, which other libraries in the process — cuDNN — then bind to.)
CXX = os.environ.get("DMPI_CXX") or ("/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++")

NVCC_FLAGS = [
    "-ccbin", CXX, "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC", "-Xptxas", "-v",
]
NVCC_FLAGS += os.environ.get("DMPI_NVCC_EXTRA", "").split()  # e.g. -DDMPI_ABLATE (bench builds only)
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-Wall", "-pthread"]

CUDA_SOURCES = sorted(str(p.relative_to(ROOT)) for p in (ROOT / "csrc").rglob("*.cu"))
HOST_SOURCES = sorted(str(p.relative_to(ROOT)) for p in (ROOT / "csrc").rglob("*.cpp") if "hooks" not in p.parts)
TORCH_EXT_SOURCE = ROOT / "csrc" / "hooks" / "autograd_hooks.cpp"   # links against libtorch: built separately
TORCH_EXT_NAME = "dmpi_torch_hooks"


def _headers_digest() -> str:
    h = hashlib.sha256()
    for pat in ("csrc/**/*.cuh", "csrc/**/*.h", "include/*.h"):
        for p in sorted(ROOT.glob(pat)):
            h.update(p.read_bytes())
    return h.hexdigest()


def _digest(src: Path, flags: list[str], hdr: str) -> str:
    h = hashlib.sha256()
    h.update(src.read_bytes())
    h.update(" ".join(flags).encode())
    h.update(hdr.encode())
    return h.hexdigest()


def _compile(src_rel: str, force: bool, hdr: str, verbose: bool) -> tuple[str, str]:
    src = ROOT / src_rel
    is_cu = src.suffix == ".cu"
    flags = NVCC_FLAGS if is_cu else CXX_FLAGS
    obj = OBJ_DIR / (src_rel.replace("/", "__") + ".o")
    stamp = obj.with_suffix(".hash")
    dig = _digest(src, flags, hdr)
    if not force and obj.exists() and stamp.exists() and stamp.read_text() == dig:
        return str(obj), ""
    cmd = ([NVCC] + flags if is_cu else [CXX] + flags) + ["-I", str(ROOT / "include"), "-c", str(src), "-o", str(obj)]
    if not is_cu:
        cmd += ["-I", "/usr/local/cuda/include"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"compile failed: {' '.join(cmd)}\n{r.stdout}\n{r.stderr}")
    stamp.write_text(dig)
    log = r.stderr if is_cu else ""
    if verbose and log:
        print(f"--- {src_rel}\n{log}")
    return str(obj), log


def build(force: bool = False, verbose: bool = False) -> dict[str, str]:
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    hdr = _headers_digest()
    logs: dict[str, str] = {}
    objs: dict[str, str] = {}
    with cf.ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
        futs = {ex.submit(_compile, s, force, hdr, verbose): s for s in CUDA_SOURCES + HOST_SOURCES}
        for f in cf.as_completed(futs):
            s = futs[f]
            objs[s], logs[s] = f.result()
    (ROOT / "build" / "ptxas.log").write_text("\n".join(f"--- {k}\n{v}" for k, v in sorted(logs.items()) if v))

    host_objs = [objs[s] for s in HOST_SOURCES if "/cudaglue/" not in s]
    cuda_objs = [objs[s] for s in CUDA_SOURCES]
    host_lib = LIB_DIR / "libdmpi_host.so"
    cuda_lib = LIB_DIR / "libdmpi_cuda.so"

    def link(out: Path, inputs: list[str], extra: list[str]) -> None:
        if not inputs:
            return
        newest = max(os.path.getmtime(i) for i in inputs)
        if not force and out.exists() and os.path.getmtime(out) >= newest:
            return
        cmd = [CXX, "-shared", "-Wl,--exclude-libs,ALL", "-o", str(out)] + inputs + extra
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed: {' '.join(cmd)}\n{r.stderr}")

    link(host_lib, host_objs, ["-pthread"])
    link(cuda_lib, cuda_objs, ["-L/usr/local/cuda/lib64", "-lcudart", "-pthread", "-ldl",
                               "-Wl,-rpath,/usr/local/cuda/lib64"])
    out = {"host": str(host_lib), "cuda": str(cuda_lib)}
    ext = build_torch_ext(force)
    if ext:
        out["torch_hooks"] = ext
    return out


def build_torch_ext(force: bool = False) -> str:
    """The pybind11 / libtorch module with the C++ autograd hooks.  Optional: when it cannot be built (no torch headers)
    the DDP wrapper falls back to its Python hooks, so a failure here is reported and swallowed."""
    out = LIB_DIR / f"{TORCH_EXT_NAME}.so"
    stamp = OBJ_DIR / f"{TORCH_EXT_NAME}.hash"
    try:
        import sysconfig

        import torch
        from torch.utils import cpp_extension as ce

        flags = ["-O2", "-std=c++17", "-fPIC", "-shared", "-pthread", f"-DTORCH_EXTENSION_NAME={TORCH_EXT_NAME}",
                 "-DTORCH_API_INCLUDE_EXTENSION_H", f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"]
        incs = [f"-I{p}" for p in ce.include_paths()] + [f"-I{sysconfig.get_paths()['include']}"]
        libdirs = ce.library_paths()
        dig = hashlib.sha256(TORCH_EXT_SOURCE.read_bytes() + " ".join(flags).encode() + torch.__version__.encode()).hexdigest()
        if not force and out.exists() and stamp.exists() and stamp.read_text() == dig:
            return str(out)
        cmd = ([CXX] + flags + incs + [str(TORCH_EXT_SOURCE), "-o", str(out)] + [f"-L{d}" for d in libdirs] +
               ["-lc10", "-ltorch_cpu", "-ltorch", "-ltorch_python"] + [f"-Wl,-rpath,{d}" for d in libdirs])
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            print(f"[build] {TORCH_EXT_NAME}: compile failed, Python hooks will be used\n{r.stderr[-1500:]}")
            return ""
        stamp.write_text(dig)
        return str(out)
    except Exception as e:  # pragma: no cover
        print(f"[build] {TORCH_EXT_NAME}: skipped ({type(e).__name__}: {e})")
        return ""


if __name__ == "__main__":
    out = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(out)
