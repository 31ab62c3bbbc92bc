"""Which kernels of one eager training step are NOT ours?  Runs one warmed-up step of each model family under
torch.profiler and prints every device kernel with its launch count; for library (ATen) kernels it also prints the
Python frames that launched them, so the residue can be traced to a line.

    python bench/aten_audit.py [resnet50|resnet18|unet ...]  > gpurun_out/aten_audit.txt
"""
import collections
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import deeplearning_mpi_b200.nn as dnn  # noqa: E402
from deeplearning_mpi_b200 import optim  # noqa: E402
from deeplearning_mpi_b200.models import UNet, resnet18, resnet50  # noqa: E402


def build(name, dev):
    torch.manual_seed(0)
    if name == "unet":
        m = UNet(out_classes=1).to(dev)
        opt = optim.Adam(m.parameters(), lr=1e-4)
        x = torch.randn(2, 3, 128, 128, device=dev)
        y = (torch.rand(2, 128, 128, device=dev) > 0.5).float()

        def step():
            loss = dnn.binary_cross_entropy_with_logits(m(x).squeeze(1), y)
            opt.zero_grad()
            loss.backward()
            dnn.clip_grad_norm_(m.parameters(), 1.0)
            opt.step()
            return loss
        return step
    m = (resnet18(num_classes=10) if name == "resnet18" else resnet50(num_classes=1000)).to(dev)
    opt = optim.SGD(m.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-5)
    nc = 10 if name == "resnet18" else 1000
    x = torch.randn(8, 3, 64, 64, device=dev)
    y = torch.randint(0, nc, (8,), device=dev)

    def step():
        opt.zero_grad()
        loss = dnn.cross_entropy(m(x), y)
        loss.backward()
        opt.step()
        return loss
    return step


def main():
    dev = torch.device("cuda:0")
    for name in (sys.argv[1:] or ["resnet50", "unet"]):
        step = build(name, dev)
        for _ in range(3):
            step()
        torch.cuda.synchronize()
        from torch.profiler import ProfilerActivity, profile

        with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True) as prof:
            step()
            torch.cuda.synchronize()
        ours, lib = collections.Counter(), collections.Counter()
        stacks = collections.defaultdict(collections.Counter)
        for ev in prof.events():
            if ev.device_type != torch.autograd.DeviceType.CUDA:
                continue
            kname = ev.name
            if "dmpi::" in kname:
                ours[kname.split("(")[0][:80]] += 1
            else:
                lib[kname[:100]] += 1
        # CPU ops that launched library kernels, with their Python stacks
        for ev in prof.events():
            if ev.device_type == torch.autograd.DeviceType.CPU and ev.name.startswith("aten::") and ev.stack:
                ks = [k for k in ev.kernels] if hasattr(ev, "kernels") else []
                if ks:
                    frame = next((f for f in ev.stack if "deeplearning_mpi_b200" in f or "bench/" in f), ev.stack[0])
                    stacks[ev.name][frame.strip()[:150]] += len(ks)
        print(f"== {name}: {sum(ours.values())} own launches, {sum(lib.values())} library launches / memory ops")
        for k, v in lib.most_common():
            print(f"   LIB {v:4d}  {k}")
        for op, fr in stacks.items():
            for f, v in fr.most_common():
                print(f"   via {op:28s} x{v:3d}  {f}")
        for k, v in ours.most_common():
            print(f"   own {v:4d}  {k}")


if __name__ == "__main__":
    main()
