"""Regenerates docs/sass/summary.txt: per-kernel counts of the Blackwell-specific SASS mnemonics in libdmpi_cuda.so.
Runs on CPU (cuobjdump only):  python bench/sass_summary.py"""
import collections, os, re, subprocess

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(root, "deeplearning_mpi_b200", "_lib", "libdmpi_cuda.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
want = re.compile(r"^(UTC\w*MMA|UTCBAR|UTCATOMSWS|LDTM|UTMALDG|UTMASTG|UTMACCTL|UBLKCP|SYNCS|REDG|RED|LDGMC|STGMC|"
                  r"MEMBAR|LDG|STG|ELECT|UCGABAR_ARV|UCGABAR_WAIT|MAPA|LDS|STS)$")
out = ["# SASS evidence (cuobjdump -sass deeplearning_mpi_b200/_lib/libdmpi_cuda.so, sm_100a); regenerate with",
       "# python bench/sass_summary.py.  Per kernel: instruction count and Blackwell-specific mnemonics",
       "# (UTC*MMA = tcgen05.mma [.2CTA = cta_group::2], LDTM = tcgen05.ld, UTMALDG/UTMASTG = TMA load/store,",
       "# UTCBAR = tcgen05.commit [.MULTICAST = to both CTAs of a pair], SYNCS = mbarrier, REDG = red.global,",
       "# LDGMC / STGMC = multimem.ld_reduce / multimem.st (NVSwitch), LD/ST .SYS = peer (NVLink) accesses,",
       "# UCGABAR = cluster barrier, ELECT = elect.sync)", ""]
name, counts, n = None, collections.Counter(), 0


def flush():
    if name and n:
        keys = sorted(counts)
        out.append(name)
        out.append(f"    instructions={n}  " + "  ".join(f"{k}={counts[k]}" for k in keys))


for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        flush()
        name, counts, n = m.group(1), collections.Counter(), 0
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if m:
        n += 1
        op = m.group(1)
        base = op.split(".")[0]
        if want.match(base):
            key = base
            for suf in ("2CTA", "MULTICAST", "SYS"):
                if "." + suf in op:
                    key += "." + suf
            counts[key] += 1
flush()
os.makedirs(os.path.join(root, "docs", "sass"), exist_ok=True)
open(os.path.join(root, "docs", "sass", "summary.txt"), "w").write("\n".join(out) + "\n")
print("\n".join(l for l in out if "igemm" in l or "wgrad" in l or "nvls" in l)[:1500])
