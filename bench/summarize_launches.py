"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import collections, csv, re, sys
path = sys.argv[1]
rows = list(csv.reader(open(path)))
h = next(i for i, r in enumerate(rows) if 'Kernel Name' in r)
hdr = rows[h]; ki, vi, ui = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
agg = collections.defaultdict(lambda: [0, 0.0]); tot = 0.0
for r in rows[h + 1:]:
    if len(r) <= vi: continue
    name = re.sub(r'void |dmpi::|at::native::', '', re.sub(r'\(.*', '', r[ki]))[:64]
    v = float(r[vi].replace(',', '')) * {'ns': 1e-3, 'us': 1.0, 'ms': 1e3}.get(r[ui], 1e-3)
    agg[name][0] += 1; agg[name][1] += v; tot += v
print(f"# {path}: total {tot/1000:.2f} ms in {sum(a[0] for a in agg.values())} launches (ncu-serialised, cold cache: compare shares)")
for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 20]:
    print(f"{t/1000:8.3f} ms {100*t/tot:5.1f}%  x{c:4d}  {n}")
