"""Per-stage forward error of the sm_100a ResNet path vs the CPU fp32 oracle (debug aid)."""
import copy, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeplearning_mpi_b200.models import resnet50, resnet18
from deeplearning_mpi_b200.models.resnet import _cba
from deeplearning_mpi_b200.ops import functional as FN
from deeplearning_mpi_b200.ops.conv import smallc_geometry
from deeplearning_mpi_b200.utils.arena import ensure_arena

def rel(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()

torch.manual_seed(0)
arch = sys.argv[1] if len(sys.argv) > 1 else "resnet50"
ref = (resnet50 if arch == "resnet50" else resnet18)(num_classes=1000)
ref.train()
fast = copy.deepcopy(ref).cuda()
x = torch.randn(32, 3, 96, 96)
with torch.no_grad():
    r = ref.relu(ref.bn1(ref.conv1(x)))
    arena = ensure_arena(fast); arena.begin_forward()
    n, c, h, w = x.shape
    _, _, hp, wp = smallc_geometry(h, w, 7, 7, 2, 3)
    xp = FN.StageInput.apply(x.cuda(), 3, hp, wp)
    a = _cba(xp, fast.conv1, fast.bn1, arena, relu=True, stem_hw=(h, w))
    print("stem", rel(a.permute(0, 3, 1, 2), r))
    r = ref.maxpool(r); a = FN.MaxPool.apply(a, 3, 2, 1)
    print("maxpool", rel(a.permute(0, 3, 1, 2), r))
    for li, (lr, lf) in enumerate(zip((ref.layer1, ref.layer2, ref.layer3, ref.layer4), (fast.layer1, fast.layer2, fast.layer3, fast.layer4))):
        for bi, (br, bf) in enumerate(zip(lr, lf)):
            r_in, a_in = r, a
            r = br(r); a = bf.forward_sm100(a, arena)
            # error of this block alone, given the oracle's input
            a_iso = bf.forward_sm100(r_in.permute(0, 2, 3, 1).contiguous().cuda().bfloat16(), arena)
            print(f"layer{li+1}.{bi}: cumulative {rel(a.permute(0,3,1,2), r):.4e}  isolated {rel(a_iso.permute(0,3,1,2), r):.4e}  shape {tuple(r.shape)}")
    r = torch.flatten(ref.avgpool(r), 1); a = FN.GlobalAvgPool.apply(a)
    print("avgpool", rel(a, r))
    r = ref.fc(r); a = FN.LinearFn.apply(a, fast.fc.weight, fast.fc.bias, arena)
    print("fc", rel(a, r), "|logits|", r.norm().item())
