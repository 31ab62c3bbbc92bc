"""Per-kernel roofline table from bench/ncu_step_metrics.sh output: time, DRAM traffic, achieved DRAM GB/s as a
fraction of the measured copy rate (MEASURED_PEAKS.json hbm_gbs, fallback 6551), tensor-pipe activity."""
import collections, csv, json, os, re, sys

path = sys.argv[1]
peak = 6551.0
try:
    peak = float(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
rows = list(csv.reader(open(path)))
h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
hdr = rows[h]
ki, mi, vi, ui, idi = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit"), hdr.index("ID")
per = collections.OrderedDict()
for r in rows[h + 1:]:
    if len(r) <= vi:
        continue
    d = per.setdefault(r[idi], {"name": re.sub(r"void |dmpi::|at::native::", "", re.sub(r"\(.*", "", r[ki]))[:48]})
    v = float(r[vi].replace(",", "") or 0)
    unit = r[ui]
    if "byte" in unit.lower():
        v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
    if r[mi] == "gpu__time_duration.sum":
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(unit, 1e-3)
    d[r[mi]] = v
agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0, 0.0])
for d in per.values():
    a = agg[d["name"]]
    t = d.get("gpu__time_duration.sum", 0.0)
    a[0] += 1
    a[1] += t
    a[2] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
    a[3] += d.get("sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_active", 0.0) * t
    a[4] += d.get("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 0.0) * t
tot = sum(a[1] for a in agg.values())
print(f"# {path}: {sum(a[0] for a in agg.values())} launches, {tot/1e3:.2f} ms (ncu-serialised, cold caches); DRAM fraction = "
      f"(dram read + write bytes) / time / {peak:.0f} GB/s (measured copy rate)")
print(f"{'kernel':48s} {'n':>4s} {'ms':>8s} {'share':>6s} {'DRAM GB':>8s} {'GB/s':>7s} {'of copy':>7s} {'tensor%':>7s}")
for n, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    gbs = a[2] / (a[1] * 1e-6) / 1e9 if a[1] else 0
    print(f"{n:48s} {a[0]:4d} {a[1]/1e3:8.3f} {100*a[1]/tot:5.1f}% {a[2]/1e9:8.3f} {gbs:7.0f} {gbs/peak:7.2f} {a[3]/a[1] if a[1] else 0:7.1f}")
