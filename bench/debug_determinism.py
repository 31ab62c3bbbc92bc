import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import deeplearning_mpi_b200.nn as dnn
from deeplearning_mpi_b200.models import resnet18
from deeplearning_mpi_b200.models.resnet import _cba
from deeplearning_mpi_b200.ops import functional as FN
from deeplearning_mpi_b200.ops.conv import smallc_geometry
from deeplearning_mpi_b200.utils.arena import ensure_arena
torch.manual_seed(0)
m = resnet18(num_classes=10).cuda()
x = torch.randn(16, 3, 64, 64, device="cuda"); y = torch.randint(0, 10, (16,), device="cuda")
arena = ensure_arena(m)
def fwd_stages():
    outs = []
    with torch.no_grad():
        arena.begin_forward()
        _, _, hp, wp = smallc_geometry(64, 64, 7, 7, 2, 3)
        xp = FN.StageInput.apply(x, 3, hp, wp); outs.append(("xp", xp.clone()))
        a = _cba(xp, m.conv1, m.bn1, arena, relu=True, stem_hw=(64, 64)); outs.append(("stem", a.clone()))
        a = FN.MaxPool.apply(a, 3, 2, 1); outs.append(("pool", a.clone()))
        for li, layer in enumerate((m.layer1, m.layer2, m.layer3, m.layer4)):
            for bi, blk in enumerate(layer):
                a = blk.forward_sm100(a, arena); outs.append((f"l{li+1}.{bi}", a.clone()))
    return outs
A = fwd_stages(); B = fwd_stages()
for (n, a), (_, b) in zip(A, B):
    d = (a.float() - b.float()); print(f"{n:8s} max|diff|={d.abs().max().item():.3e} n_diff={int((d != 0).sum())} of {d.numel()}")
def grads():
    for p in m.parameters():
        if p.grad is not None: p.grad.zero_()
    arena.flat_g.zero_()
    dnn.cross_entropy(m(x), y).backward(); torch.cuda.synchronize(); return arena.flat_g.clone()
G1 = grads(); G2 = grads()
print("grad rel diff run-to-run:", ((G1 - G2).norm() / G2.norm()).item())
for s in arena.slots[:8] + arena.slots[-4:]:
    a, b = G1[s.offset:s.offset+s.numel], G2[s.offset:s.offset+s.numel]
    print(f"  {s.name:30s} {((a-b).norm()/(b.norm()+1e-12)).item():.3e}")
