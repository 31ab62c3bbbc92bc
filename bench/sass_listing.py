"""SASS evidence for every kernel of libdmpi_cuda.so (CPU only, cuobjdump):
  docs/sass/libdmpi_cuda.sass.xz       the complete listing (xz -d to read; ~25 MB of text)
  docs/sass/excerpts/<kernel>.txt      per kernel: instruction count and the lines that prove the Blackwell / NVLink
                                       path (UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG / UTMASTG = TMA,
                                       UTCBAR = tcgen05.commit, LDGMC = multimem.ld_reduce, *.SYS = peer accesses, REDG)
Run after a build:  python bench/sass_listing.py   (bench/sass_summary.py writes the mnemonic-count table)."""
import lzma
import os
import re
import subprocess

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(root, "deeplearning_mpi_b200", "_lib", "libdmpi_cuda.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
out_dir = os.path.join(root, "docs", "sass")
ex_dir = os.path.join(out_dir, "excerpts")
os.makedirs(ex_dir, exist_ok=True)
for f in os.listdir(ex_dir):
    os.remove(os.path.join(ex_dir, f))
with lzma.open(os.path.join(out_dir, "libdmpi_cuda.sass.xz"), "wt", preset=9) as f:
    f.write(sass)

key = re.compile(r"\b(UTC\w*MMA|UTCBAR|LDTM|STTM|UTMALDG|UTMASTG|UTMAPF|UBLKCP|LDGMC|STGMC|REDG?|ATOMG|UCGABAR_\w+|"
                 r"\w+\.SYS\b|SYNCS\.\w+|ELECT)")
demangle = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True,
                          text=True).stdout.splitlines()
names = dict(zip(re.findall(r"Function : (\S+)", sass), demangle))
blocks = re.split(r"(?m)^\s*Function : ", sass)[1:]
index = []
for blk in blocks:
    mangled = blk.split("\n", 1)[0].strip()
    pretty = names.get(mangled, mangled)
    short = re.sub(r"\(.*", "", pretty.replace("void ", "").replace("dmpi::", ""))
    fname = re.sub(r"[^A-Za-z0-9_]+", "_", short).strip("_")[:80] + ".txt"
    lines = [l for l in blk.splitlines() if re.match(r"\s*/\*[0-9a-f]{4,}\*/", l)]
    seen, picked = {}, []
    for l in lines:
        m = key.search(l)
        if not m:
            continue
        op = re.search(r"\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", l)
        k = op.group(1) if op else m.group(1)
        seen[k] = seen.get(k, 0) + 1
        if seen[k] <= 3:  # first three occurrences of every distinct opcode
            picked.append(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l).rstrip())
    with open(os.path.join(ex_dir, fname), "w") as f:
        f.write(f"{pretty}\n{len(lines)} SASS instructions; opcode counts of the marked classes:\n  ")
        f.write("  ".join(f"{k} x{v}" for k, v in sorted(seen.items())) + "\n\n")
        f.write("\n".join(picked) + "\n")
    index.append((short, len(lines), fname))
with open(os.path.join(out_dir, "README.md"), "w") as f:
    f.write("# SASS evidence\n\n`libdmpi_cuda.sass.xz` is the complete `cuobjdump -sass` listing of the shipped library "
            "(`xz -dc docs/sass/libdmpi_cuda.sass.xz | less`); `summary.txt` counts the Blackwell-specific mnemonics per "
            "kernel; `excerpts/` holds, for every kernel, the first occurrences of each tcgen05 / TMA / mbarrier / "
            "multimem / peer-memory instruction with its address.  Regenerate with `python bench/sass_listing.py && "
            "python bench/sass_summary.py` after a build.\n\n| kernel | SASS instructions | excerpt |\n|---|---|---|\n")
    for short, n, fname in sorted(index):
        f.write(f"| `{short}` | {n} | [excerpts/{fname}](excerpts/{fname}) |\n")
print(len(index), "kernels")
