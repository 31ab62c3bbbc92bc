"""Loads a checkpoint written by THIS framework through the reference's own, unmodified `build_model` with resume on
(ref:pytorch/resnet/main.py:36-54 or ref:pytorch/unet/train.py:60-75, imported from baseline/_ref) and prints the
model's eval-mode output checksum on a fixed input.  argv: resnet|unet  checkpoint  input.pt  output.pt"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = os.path.join(ROOT, "baseline", "_ref")
kind, ckpt, xin, yout = sys.argv[1:5]
for k, v in dict(RANK="0", LOCAL_RANK="0", WORLD_SIZE="1", MASTER_ADDR="127.0.0.1").items():
    os.environ.setdefault(k, v)
sys.path.insert(0, REF)
sys.path.insert(0, os.path.join(REF, "pytorch", "unet"))
import torch  # noqa: E402
import torch.distributed as tdist  # noqa: E402

tdist.init_process_group("nccl")
if kind == "resnet":
    import pytorch.resnet.main as ref  # noqa: E402

    ref.resume, ref.model_filepath = True, ckpt
    ddp = ref.build_model()
else:
    import pytorch.unet.train as ref  # noqa: E402

    ddp = ref.build_model(True, ckpt)
ddp.eval()
x = torch.load(xin).cuda()
with torch.no_grad():
    y = ddp(x)
torch.save(y.float().cpu(), yout)
tdist.destroy_process_group()
print("REF_RESUME_OK")
