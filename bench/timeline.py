"""In-situ kernel timeline of one training step (CUPTI activity records through torch.profiler; kernels run as in the
benchmark: concurrently, warm caches, inside the replayed CUDA graph).  Unlike the ncu launch list nothing is
serialised or cache-flushed, so durations AND the idle gaps between kernels are the real ones.

    python bench/timeline.py --model resnet50 [--batch 256 --size 224] [--no-graph]
writes gpurun_out/timeline_<model>.csv (one row per launch: start_us, dur_us, gap_before_us, stream, name) and prints
totals by kernel name, the busy/idle split of the step and the largest gaps.
"""
import argparse
import collections
import csv
import os
import re
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("RANK", "0"); os.environ.setdefault("LOCAL_RANK", "0"); os.environ.setdefault("WORLD_SIZE", "1")
os.environ.setdefault("LOCAL_WORLD_SIZE", "1"); os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29533")

import bench as B  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet50")
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--size", type=int, default=0)
    ap.add_argument("--no-graph", dest="graph", action="store_false")
    ap.add_argument("--top", type=int, default=24)
    a = ap.parse_args()
    args = argparse.Namespace(model=a.model, batch=a.batch, size=a.size, impl="ours", graph=a.graph, steps=2, warmup=3,
                              no_e2e=True, gpus=1)
    batch, size = B.model_defaults(args)
    arm = B.build_ours(args, batch, size)
    step, device = arm["step"], arm["device"]
    host = B.make_host_batches(args, batch, size, n_batches=1)
    x, y = host[0][0].to(device), host[0][1].to(device)
    if a.graph:
        from deeplearning_mpi_b200.utils.graph import GraphedTrainStep

        step = GraphedTrainStep(step, (x, y))
    for _ in range(4):
        step(x, y)
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile

    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        step(x, y)
        step(x, y)
        torch.cuda.synchronize()
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    evs.sort(key=lambda e: e.time_range.start)
    # keep the second step only: split at the largest index where the kernel sequence repeats
    names = [e.name for e in evs]
    half = len(evs) // 2
    if names[:half] == names[half:2 * half]:
        evs = evs[half:2 * half]
    t0 = evs[0].time_range.start
    rows, agg = [], collections.defaultdict(lambda: [0, 0.0])
    end_prev = None
    busy_end = t0
    idle = 0.0
    for e in evs:
        s, d = e.time_range.start - t0, e.time_range.end - e.time_range.start
        gap = max(0.0, e.time_range.start - busy_end) if end_prev is not None else 0.0
        idle += gap
        busy_end = max(busy_end, e.time_range.end)
        end_prev = e.time_range.end
        short = re.sub(r"void |dmpi::|at::native::", "", re.sub(r"\(.*", "", e.name))[:60]
        rows.append((round(s, 2), round(d, 2), round(gap, 2), short))
        agg[short][0] += 1
        agg[short][1] += d
    span = busy_end - t0
    os.makedirs("gpurun_out", exist_ok=True)
    with open(f"gpurun_out/timeline_{a.model}.csv", "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["start_us", "dur_us", "gap_before_us", "name"])
        w.writerows(rows)
    tot = sum(v[1] for v in agg.values())
    print(f"# {a.model} batch {batch} size {size} graph={a.graph}: {len(rows)} device activities, span {span/1e3:.3f} ms, "
          f"sum of durations {tot/1e3:.3f} ms, idle (no kernel running) {idle/1e3:.3f} ms = {100*idle/span:.1f}% of the span")
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[: a.top]:
        print(f"{t/1e3:8.3f} ms {100*t/span:5.1f}%  x{c:4d}  avg {t/c:7.1f} us  {n}")
    gaps = sorted(rows, key=lambda r: -r[2])[:8]
    print("largest gaps (us before kernel):", [(g[2], g[3][:28]) for g in gaps])
    arm["finish"]()


if __name__ == "__main__":
    main()
