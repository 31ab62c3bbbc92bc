"""End-to-end forward/backward of the sm_100a model path against the stock-PyTorch fp32 oracle (the same
nn.Module executed on CPU), plus optimizer-step parity and checkpoint round trips."""
import copy

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _rel(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def _cos(a, b):
    a, b = a.float().cpu().flatten(), b.float().cpu().flatten()
    return (torch.dot(a, b) / (a.norm() * b.norm() + 1e-20)).item()


def _round_weights_to_bf16(model):
    """Make the fp32 master weights exactly representable in bf16 so that oracle (fp32 weights) and kernels (bf16
    shadow of the same weights) multiply identical numbers; otherwise the 0.2 % weight rounding flips ~0.3 % of the
    ReLU gates, which alone is a ~5 % L2 difference in the gradients."""
    with torch.no_grad():
        for m in model.modules():
            if isinstance(m, (torch.nn.Conv2d, torch.nn.ConvTranspose2d, torch.nn.Linear)):
                m.weight.copy_(m.weight.bfloat16().float())


def _grads_logical(model):
    return {n: p.grad.detach().float().cpu().clone() for n, p in model.named_parameters()}


def _run_pair(make, x, target, loss_kind, tol_out, tol_grad, min_cos=0.85):
    import deeplearning_mpi_b200.nn as dnn

    torch.manual_seed(0)
    ref = make()
    ref.train()
    _round_weights_to_bf16(ref)
    fast = copy.deepcopy(ref).to(DEV)
    fast.train()
    # oracle on CPU, fp32
    out_ref = ref(x)
    if loss_kind == "xent":
        loss_ref = F.cross_entropy(out_ref, target)
    else:
        loss_ref = F.binary_cross_entropy_with_logits(out_ref.squeeze(1), target)
    loss_ref.backward()
    # ours
    out = fast(x.to(DEV))
    if loss_kind == "xent":
        loss = dnn.cross_entropy(out, target.to(DEV))
    else:
        loss = dnn.binary_cross_entropy_with_logits(out.squeeze(1), target.to(DEV))
    loss.backward()
    torch.cuda.synchronize()
    r_out = _rel(out, out_ref)
    print(f"out rel={r_out:.4e} loss={loss.item():.5f} ref={loss_ref.item():.5f}")
    # Randomly initialised BatchNorm networks amplify perturbations by ~1.2x per layer in training mode (mean-field
    # "BN chaos"), so bf16-vs-fp32 differences of whole networks are only bounded coarsely here; the tight
    # comparisons are the eval-mode check below, test_blocks_strict and the per-kernel tests.
    assert r_out < tol_out
    assert abs(loss.item() - loss_ref.item()) < 0.05 * abs(loss_ref.item()) + 1e-3
    g_ref = {n: p.grad for n, p in ref.named_parameters()}
    g = _grads_logical(fast)
    # bf16 activations flip a few ReLU gates relative to the fp32 oracle, so deep-network gradients are compared by
    # direction (cosine) and norm; the strict element-wise checks live in test_blocks_strict / test_conv_gpu.
    worst, worst_cos = 0.0, 1.0
    for n in g_ref:
        if n.endswith(".bias") and ("double_conv" in n) and g_ref[n].abs().max() < 1e-4:
            continue  # conv bias behind BatchNorm: analytically zero gradient
        r, c = _rel(g[n], g_ref[n]), _cos(g[n], g_ref[n])
        worst, worst_cos = max(worst, r), min(worst_cos, c)
        if r > tol_grad:
            print(f"  grad {n}: rel={r:.3e} cos={c:.4f} |ref|={g_ref[n].norm():.3e}")
    print(f"worst grad rel={worst:.4e} worst cos={worst_cos:.4f}")
    if min_cos is not None:
        assert worst < tol_grad and worst_cos > min_cos
    assert all(torch.isfinite(v).all() for v in g.values())
    # running statistics followed the same EMA
    for (n, b), (_, br) in zip(fast.named_buffers(), ref.named_buffers()):
        if b.dtype == torch.float32:
            if min_cos is not None:
                assert _rel(b, br) < 2e-2, n
        else:
            assert int(b.item()) == int(br.item()), n
    # eval mode (running statistics): no batch-statistics feedback, errors stay at bf16 rounding level
    ref.eval()
    fast.eval()
    fast.load_state_dict(ref.state_dict())
    with torch.no_grad():
        e_ref = ref(x)
        e_out = fast(x.to(DEV))
    r_eval = _rel(e_out, e_ref)
    print(f"eval-mode out rel={r_eval:.4e}")
    assert r_eval < 3e-2
    return ref, fast


def test_resnet18_cifar():
    from deeplearning_mpi_b200.models import resnet18

    x = torch.randn(64, 3, 64, 64)
    y = torch.randint(0, 10, (64,))
    _run_pair(lambda: resnet18(num_classes=10), x, y, "xent", 0.1, 0.6)


def test_resnet50_small():
    from deeplearning_mpi_b200.models import resnet50

    x = torch.randn(32, 3, 96, 96)
    y = torch.randint(0, 1000, (32,))
    _run_pair(lambda: resnet50(num_classes=1000), x, y, "xent", 0.5, 1.5, min_cos=None)  # 50 layers of BN chaos: see _run_pair


def test_unet_small():
    from deeplearning_mpi_b200.models import UNet

    x = torch.randn(2, 3, 64, 64)
    t = (torch.rand(2, 64, 64) > 0.5).float()
    _run_pair(lambda: UNet(out_classes=1), x, t, "bce", 0.1, 0.6)


class _Block(torch.nn.Module):
    """One residual block as a root module, fed with NCHW fp32 on CPU and NHWC bf16 on CUDA."""

    def __init__(self, blk):
        super().__init__()
        self.blk = blk

    def forward(self, x):
        if not x.is_cuda:
            return self._oracle(x)
        from deeplearning_mpi_b200.utils.arena import ensure_arena

        arena = ensure_arena(self)
        arena.begin_forward()
        return self.blk.forward_sm100(x, arena)


def _q(t):
    """bf16 storage emulation: the oracle keeps fp32 math but rounds wherever our kernels store bf16 (forward
    and, through the cast's backward, the gradients too), so ReLU gates agree with the kernel path."""
    return t.bfloat16().float()


def _oracle_block(blk, x):
    import torch.nn.functional as F

    def cb(conv, bn, a, relu, res=None):
        z = bn(_q(conv(a)))
        if res is not None:
            z = z + res
        return _q(F.relu(z) if relu else z)

    ident = x
    if hasattr(blk, "conv3"):
        o = cb(blk.conv1, blk.bn1, x, True)
        o = cb(blk.conv2, blk.bn2, o, True)
        if blk.downsample is not None:
            ident = cb(blk.downsample[0], blk.downsample[1], x, False)
        return cb(blk.conv3, blk.bn3, o, True, ident)
    o = cb(blk.conv1, blk.bn1, x, True)
    if blk.downsample is not None:
        ident = cb(blk.downsample[0], blk.downsample[1], x, False)
    return cb(blk.conv2, blk.bn2, o, True, ident)


_Block._oracle = lambda self, x: _oracle_block(self.blk, x)


@pytest.mark.parametrize("kind", ["basic", "basic_down", "bottleneck", "bottleneck_down"])
def test_blocks_strict(kind):
    from deeplearning_mpi_b200.models.resnet import BasicBlock, Bottleneck
    import torch.nn as nn

    torch.manual_seed(3)
    if kind == "basic":
        blk, cin = BasicBlock(64, 64), 64
    elif kind == "basic_down":
        ds = nn.Sequential(nn.Conv2d(64, 128, 1, 2, bias=False), nn.BatchNorm2d(128))
        blk, cin = BasicBlock(64, 128, 2, ds), 64
    elif kind == "bottleneck":
        blk, cin = Bottleneck(256, 64), 256
    else:
        ds = nn.Sequential(nn.Conv2d(256, 512, 1, 2, bias=False), nn.BatchNorm2d(512))
        blk, cin = Bottleneck(256, 128, 2, ds), 256
    ref = _Block(blk)
    for m in ref.modules():
        if isinstance(m, nn.BatchNorm2d):
            nn.init.uniform_(m.weight, 0.5, 1.5)
            nn.init.normal_(m.bias, 0, 0.3)
    _round_weights_to_bf16(ref)
    fast = copy.deepcopy(ref).to(DEV)
    x = torch.randn(8, cin, 16, 16).bfloat16().float()
    xr = x.clone().requires_grad_(True)
    out_ref = ref(xr)
    g = torch.randn_like(out_ref).bfloat16().float()
    out_ref.backward(g)
    xf = x.permute(0, 2, 3, 1).contiguous().to(DEV).bfloat16().requires_grad_(True)
    out = fast(xf)
    out.backward(g.permute(0, 2, 3, 1).contiguous().to(DEV).bfloat16())
    torch.cuda.synchronize()
    r = _rel(out.permute(0, 3, 1, 2), out_ref)
    rdx = _rel(xf.grad.permute(0, 3, 1, 2), xr.grad)
    print(f"[{kind}] out rel={r:.3e} dx rel={rdx:.3e}")
    assert r < 1e-2 and rdx < 3e-2
    gr = {n: p.grad for n, p in ref.named_parameters()}
    for n, p in fast.named_parameters():
        rr = _rel(p.grad, gr[n])
        print(f"   {n}: rel={rr:.3e}")
        assert rr < 3e-2, n


def test_training_reduces_loss_and_checkpoint_roundtrip(tmp_path):
    import torchvision

    import deeplearning_mpi_b200.nn as dnn
    from deeplearning_mpi_b200 import optim
    from deeplearning_mpi_b200.models import resnet18

    torch.manual_seed(1)
    model = resnet18(num_classes=10).to(DEV)
    opt = optim.SGD(model.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-5)
    x = torch.randn(32, 3, 32, 32, device=DEV)
    y = torch.randint(0, 10, (32,), device=DEV)
    losses = []
    for _ in range(12):
        opt.zero_grad()
        loss = dnn.cross_entropy(model(x), y)
        loss.backward()
        opt.step()
        losses.append(loss.item())
    print(losses)
    assert losses[-1] < 0.5 * losses[0]
    # checkpoint written by us loads into the stock torchvision model used by the reference trainer
    path = tmp_path / "ckpt.pth"
    torch.save(model.state_dict(), path)
    stock = torchvision.models.resnet18(weights=None)
    stock.fc = torch.nn.Linear(512, 10)
    stock.load_state_dict(torch.load(path, map_location="cpu"))
    model.eval()
    stock.eval()
    with torch.no_grad():
        a = model(x).cpu()
        b = stock(x.cpu())
    print("eval rel", _rel(a, b))
    assert _rel(a, b) < 5e-2
    # and back: stock weights load into ours (after the arena exists)
    model.load_state_dict(stock.state_dict())
    with torch.no_grad():
        a2 = model(x).cpu()
    assert _rel(a2, b) < 5e-2


def test_adam_clip_unet_steps():
    import deeplearning_mpi_b200.nn as dnn
    from deeplearning_mpi_b200 import optim
    from deeplearning_mpi_b200.models import UNet

    torch.manual_seed(2)
    model = UNet(out_classes=1).to(DEV)
    opt = optim.Adam(model.parameters(), lr=1e-3)
    x = torch.randn(2, 3, 64, 64, device=DEV)
    t = (torch.rand(2, 64, 64, device=DEV) > 0.5).float()
    losses = []
    for _ in range(8):
        pred = model(x).squeeze(1)
        loss = dnn.binary_cross_entropy_with_logits(pred, t)
        opt.zero_grad()
        loss.backward()
        norm = dnn.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
        losses.append(loss.item())
    print(losses, float(norm))
    assert losses[-1] < losses[0]


def test_cuda_graph_training_step():
    """Whole-iteration CUDA-graph capture: the replayed step trains like the eager one (SGD and Adam+clip)."""
    import deeplearning_mpi_b200.nn as dnn
    from deeplearning_mpi_b200 import optim
    from deeplearning_mpi_b200.models import UNet, resnet18
    from deeplearning_mpi_b200.utils.graph import GraphedTrainStep

    torch.manual_seed(5)
    model = resnet18(num_classes=10).to(DEV)
    opt = optim.SGD(model.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-5)

    def step(x, y):
        opt.zero_grad()
        loss = dnn.cross_entropy(model(x), y)
        loss.backward()
        opt.step()
        return loss

    x = torch.randn(32, 3, 32, 32, device=DEV)
    y = torch.randint(0, 10, (32,), device=DEV)
    g = GraphedTrainStep(step, (x, y), warmup=2)
    losses = [float(g(x, y).item()) for _ in range(12)]
    print("graphed resnet18:", losses)
    assert losses[-1] < 0.6 * losses[0] and g.launches_per_replay > 50
    # new data through the static inputs
    x2 = torch.randn(32, 3, 32, 32, device=DEV)
    l2 = float(g(x2, y).item())
    assert l2 > losses[-1]

    unet = UNet(out_classes=1).to(DEV)
    adam = optim.Adam(unet.parameters(), lr=1e-3)

    def ustep(xi, ti):
        pred = unet(xi).squeeze(1)
        loss = dnn.binary_cross_entropy_with_logits(pred, ti)
        adam.zero_grad()
        loss.backward()
        dnn.clip_grad_norm_(unet.parameters(), 1.0)
        adam.step()
        return loss

    xi = torch.randn(2, 3, 64, 64, device=DEV)
    ti = (torch.rand(2, 64, 64, device=DEV) > 0.5).float()
    gu = GraphedTrainStep(ustep, (xi, ti), warmup=2)
    ul = [float(gu(xi, ti).item()) for _ in range(8)]
    print("graphed unet:", ul)
    assert ul[-1] < ul[0]
    # warm-up steps + replays advanced the device-side step count (the capture pass itself records but does not execute)
    assert float(adam.state["step_t"].item()) == 2 + 8


@pytest.mark.parametrize("kind", ["resnet18", "resnet50", "unet"])
def test_odd_input_shapes(kind):
    """Non-square / odd spatial sizes: every tile planner (pixel boxes, patch tiles, parity views) has ragged edges here.
    Eval-mode forward is compared strictly against the fp32 oracle; the training step must stay finite and agree with
    the oracle on the loss."""
    import deeplearning_mpi_b200.nn as dnn
    from deeplearning_mpi_b200.models import UNet, resnet18, resnet50

    torch.manual_seed(3)
    if kind == "unet":
        make, x = (lambda: UNet(out_classes=1)), torch.randn(3, 3, 48, 80)
        target = (torch.rand(3, 48, 80) > 0.5).float()
    else:
        make = (lambda: resnet18(num_classes=10)) if kind == "resnet18" else (lambda: resnet50(num_classes=10))
        x, target = torch.randn(5, 3, 75, 101), torch.randint(0, 10, (5,))
    ref = make()
    _round_weights_to_bf16(ref)
    fast = copy.deepcopy(ref).to(DEV)
    ref.eval()
    fast.eval()
    with torch.no_grad():
        r_eval = _rel(fast(x.to(DEV)), ref(x))
    print(f"{kind} eval rel={r_eval:.4e}")
    assert r_eval < 3e-2
    ref.train()
    fast.train()
    if kind == "unet":
        loss_ref = F.binary_cross_entropy_with_logits(ref(x).squeeze(1), target)
        loss = dnn.binary_cross_entropy_with_logits(fast(x.to(DEV)).squeeze(1), target.to(DEV))
    else:
        loss_ref = F.cross_entropy(ref(x), target)
        loss = dnn.cross_entropy(fast(x.to(DEV)), target.to(DEV))
    loss.backward()
    torch.cuda.synchronize()
    assert abs(loss.item() - loss_ref.item()) < 0.1 * abs(loss_ref.item()) + 1e-2, (loss.item(), loss_ref.item())
    assert all(torch.isfinite(p.grad).all() for p in fast.parameters() if p.grad is not None)


def test_wgrad_side_stream_matches_default():
    """The opt-in side stream for weight gradients must produce the same gradients as the in-order path, up to the
    run-to-run noise of the in-order path itself (fp32 atomics reorder the BatchNorm sums; random-init BN networks
    amplify that), which is measured by running the in-order path twice."""
    import deeplearning_mpi_b200.nn as dnn
    from deeplearning_mpi_b200.models import resnet18
    from deeplearning_mpi_b200.ops import functional as FN

    torch.manual_seed(5)
    base = resnet18(num_classes=10)
    x, y = torch.randn(32, 3, 64, 64).to(DEV), torch.randint(0, 10, (32,)).to(DEV)

    def run(flag):
        FN._WGRAD_STREAM = flag
        try:
            m = copy.deepcopy(base).to(DEV)
            m.train()
            dnn.cross_entropy(m(x), y).backward()
            torch.cuda.synchronize()
            assert not FN._pending_wgrads
            return _grads_logical(m)
        finally:
            FN._WGRAD_STREAM = False

    a, b, c = run(False), run(False), run(True)
    names = [n for n in a if a[n].norm() > 1e-6]
    noise = max(_rel(b[n], a[n]) for n in names)
    diff = max(_rel(c[n], a[n]) for n in names)
    print(f"in-order run-to-run noise = {noise:.3e}, side-stream vs in-order = {diff:.3e}")
    assert diff <= 3 * noise + 2e-2


def test_training_step_is_reproducible():
    """Same weights, same batch, twice.  Default (level 1): the forward pass is bit-identical and the gradients agree to
    fp32 summation order (only the split-K weight-gradient adds are unordered, and they do not feed back into the step).
    Level 2 (deterministic split-K as well): the whole gradient arena is bit-identical.  Before the BatchNorm reductions
    were made order-independent the same comparison gave ~20 % (ReLU gate flips downstream of 1-ulp differences in the
    batch statistics)."""
    import deeplearning_mpi_b200.nn as dnn
    from deeplearning_mpi_b200.models import resnet50
    from deeplearning_mpi_b200.ops import conv as CV
    from deeplearning_mpi_b200.utils.arena import ensure_arena

    torch.manual_seed(11)
    base = resnet50(num_classes=10)
    x, y = torch.randn(16, 3, 96, 96).to(DEV), torch.randint(0, 10, (16,)).to(DEV)

    def twice():
        outs, grads = [], []
        for _ in range(2):
            m = copy.deepcopy(base).to(DEV)
            m.train()
            out = m(x)
            dnn.cross_entropy(out, y).backward()
            torch.cuda.synchronize()
            outs.append(out.detach().clone())
            grads.append(ensure_arena(m).flat_g.clone())
        assert torch.equal(outs[0], outs[1]), "forward pass is not bit-reproducible"
        return grads

    saved = CV._DETERMINISTIC_WGRAD
    try:
        CV._DETERMINISTIC_WGRAD = False
        g = twice()
        rel = ((g[0] - g[1]).norm() / (g[1].norm() + 1e-20)).item()
        print(f"level 1: run-to-run gradient rel diff = {rel:.3e}")
        assert rel < 1e-5
        CV._DETERMINISTIC_WGRAD = True
        g = twice()
        assert torch.equal(g[0], g[1]), "level 2: gradients are not bit-reproducible"
    finally:
        CV._DETERMINISTIC_WGRAD = saved


def _oracle_resnet(model, x):
    """fp32 math with bf16 rounding wherever the kernels store bf16 (see _q): stem, every block, pooled features."""
    a = _q(F.relu(model.bn1(_q(model.conv1(x)))))
    a = model.maxpool(a)
    for layer in (model.layer1, model.layer2, model.layer3, model.layer4):
        for blk in layer:
            a = _oracle_block(blk, a)
    a = _q(model.avgpool(a).flatten(1))
    return model.fc(a)


@pytest.mark.parametrize("arch", ["resnet18", "resnet50"])
def test_resnet_gradients_vs_bf16_oracle(arch):
    """Whole-network gradients are ASSERTED (not only finite): against an oracle that rounds to bf16 exactly where the
    kernels store bf16, the ReLU gates agree and what remains is accumulation-order noise amplified by depth."""
    import deeplearning_mpi_b200.nn as dnn
    from deeplearning_mpi_b200 import models

    torch.manual_seed(0)
    ref = getattr(models, arch)(num_classes=100)
    ref.train()
    # A randomly initialised train-mode BatchNorm ResNet is chaotic: with residual-branch gammas of O(1) the ORACLE
    # itself answers a 0.2 % input perturbation with uncorrelated ResNet-50 gradients (global rel 1.26, cos -0.1,
    # measured on the CPU), so no implementation can be asserted against it.  With the last BatchNorm of every block
    # scaled by 0.1 (the usual "small / zero residual gamma" start) the same perturbation moves the oracle's gradients
    # by rel 0.15 / cos 0.95 (both depths) - that is the regime asserted here.
    for name, m in ref.named_modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            last = name.endswith("bn3") or (arch == "resnet18" and name.endswith("bn2"))
            k = 0.1 if last else 1.0
            torch.nn.init.uniform_(m.weight, 0.5 * k, 1.5 * k)
            torch.nn.init.normal_(m.bias, 0, 0.2)
    _round_weights_to_bf16(ref)
    fast = copy.deepcopy(ref).to(DEV)
    # (a larger batch / image keeps the per-channel BatchNorm sample counts of the last stage in the hundreds: with a
    # handful of samples per channel the train-mode statistics amplify rounding noise chaotically)
    x = torch.randn(32, 3, 96, 96).bfloat16().float()
    y = torch.randint(0, 100, (32,))
    loss_ref = F.cross_entropy(_oracle_resnet(ref, x), y)
    loss_ref.backward()
    loss = dnn.cross_entropy(fast(x.to(DEV)), y.to(DEV))
    loss.backward()
    torch.cuda.synchronize()
    g_ref = {n: p.grad for n, p in ref.named_parameters()}
    g = _grads_logical(fast)
    worst, worst_cos, tot_num, tot_den = 0.0, 1.0, 0.0, 0.0
    for n in g_ref:
        r, c = _rel(g[n], g_ref[n]), _cos(g[n], g_ref[n])
        worst, worst_cos = max(worst, r), min(worst_cos, c)
        tot_num += (g[n].float().cpu() - g_ref[n]).norm().item() ** 2
        tot_den += g_ref[n].norm().item() ** 2
    glob = (tot_num / tot_den) ** 0.5
    print(f"[{arch}] loss {loss.item():.4f} vs {loss_ref.item():.4f}; gradients: global rel {glob:.3e}, worst tensor rel "
          f"{worst:.3e}, worst cos {worst_cos:.4f}")
    assert abs(loss.item() - loss_ref.item()) < 2e-2 * abs(loss_ref.item())
    # the per-block test (test_blocks_strict) holds 3e-2; the rest is depth x BatchNorm amplification of
    # summation-order differences, bounded here by twice the oracle's own sensitivity quoted above
    assert glob < 0.3 and worst_cos > 0.85, (glob, worst, worst_cos)


def test_loss_curve_matches_stock_torch():
    """N optimizer steps of this framework against stock PyTorch (torchvision ResNet-18,
    cuDNN, fp32, torch.optim.SGD) from the same weights on the same batches: the two loss curves stay together."""
    import torchvision

    import deeplearning_mpi_b200.nn as dnn
    from deeplearning_mpi_b200 import optim
    from deeplearning_mpi_b200.models import resnet18

    torch.manual_seed(11)
    ours = resnet18(num_classes=10).to(DEV)
    stock = torchvision.models.resnet18(weights=None)
    stock.fc = torch.nn.Linear(512, 10)
    stock.load_state_dict(ours.state_dict())
    stock = stock.to(DEV).train()
    o1 = optim.SGD(ours.parameters(), lr=0.02, momentum=0.9, weight_decay=1e-5)
    o2 = torch.optim.SGD(stock.parameters(), lr=0.02, momentum=0.9, weight_decay=1e-5)
    g = torch.Generator().manual_seed(5)
    proto = torch.randn(10, 3, 32, 32, generator=g)
    batches = []
    for _ in range(4):   # learnable data: class prototype + noise
        y = torch.randint(0, 10, (128,), generator=g)
        batches.append(((proto[y] + 0.7 * torch.randn(128, 3, 32, 32, generator=g)).to(DEV), y.to(DEV)))
    la, lb = [], []
    for step in range(24):
        x, y = batches[step % 4]
        o1.zero_grad()
        l1 = dnn.cross_entropy(ours(x), y)
        l1.backward()
        o1.step()
        o2.zero_grad()
        l2 = F.cross_entropy(stock(x), y)
        l2.backward()
        o2.step()
        la.append(l1.item())
        lb.append(l2.item())
    print("ours ", [round(v, 3) for v in la])
    print("stock", [round(v, 3) for v in lb])
    diffs = [abs(a - b) for a, b in zip(la, lb)]
    assert la[-1] < 0.5 * la[0] and lb[-1] < 0.5 * lb[0]
    assert diffs[0] < 2e-2 and max(diffs[:6]) < 0.1
    assert sum(diffs) / len(diffs) < 0.1 and max(diffs) < 0.3, (max(diffs), sum(diffs) / len(diffs))


@pytest.mark.parametrize("kind", ["resnet", "unet"])
def test_reference_script_resumes_from_our_checkpoint(kind, tmp_path):
    """save_checkpoint() output goes through the REFERENCE's own build_model(resume) (baseline/_ref, unmodified,
    ref:pytorch/resnet/main.py:36-54 / ref:pytorch/unet/train.py:60-75) in a separate process: strict load_state_dict
    through stock DDP succeeds and its eval output equals ours."""
    import os
    import subprocess
    import sys

    from deeplearning_mpi_b200.models import UNet, resnet18
    from deeplearning_mpi_b200.parallel import dist
    from deeplearning_mpi_b200.parallel.ddp import DistributedDataParallel
    from deeplearning_mpi_b200.utils.checkpoint import save_checkpoint

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if not os.path.isdir(os.path.join(root, "baseline", "_ref", "pytorch")):
        pytest.skip("reference arm not installed (baseline/_ref)")
    for k, v in dict(RANK="0", LOCAL_RANK="0", WORLD_SIZE="1", LOCAL_WORLD_SIZE="1", MASTER_ADDR="127.0.0.1",
                     MASTER_PORT="29877").items():
        os.environ.setdefault(k, v)
    torch.manual_seed(2)
    created = not dist.is_initialized()
    if created:
        dist.init_process_group("nccl")
    try:
        model = (resnet18(num_classes=10) if kind == "resnet" else UNet(out_classes=1)).to(DEV)
        ddp = DistributedDataParallel(model, device_ids=[0], output_device=0)
        x = torch.randn(4, 3, 32, 32)
        ddp.train()
        ddp(x.to(DEV))              # one training forward so that the running statistics are not the initial ones
        path = str(tmp_path / "ckpt.pth")
        save_checkpoint(ddp, path)
        ddp.eval()
        with torch.no_grad():
            mine = ddp(x.to(DEV)).float().cpu()
    finally:
        if created:
            dist.destroy_process_group()
    torch.save(x, tmp_path / "x.pt")
    env = dict(os.environ, MASTER_PORT="29878")
    r = subprocess.run([sys.executable, os.path.join(root, "tests", "mp", "ref_resume_worker.py"), kind, path,
                        str(tmp_path / "x.pt"), str(tmp_path / "y.pt")], capture_output=True, text=True, timeout=600,
                       env=env, cwd=root)
    assert r.returncode == 0 and "REF_RESUME_OK" in r.stdout, (r.stdout + r.stderr)[-3000:]
    theirs = torch.load(tmp_path / "y.pt")
    rel = _rel(mine, theirs)
    print(f"[{kind}] ours vs reference-loaded eval output: rel {rel:.3e}")
    assert rel < 5e-2


@pytest.mark.parametrize("example", ["resnet", "unet"])
def test_example_trainers_graph_flag(example, tmp_path):
    """The shipped CLIs reach the benchmarked (CUDA-graph) path: `--graph` trains and the loss falls."""
    import os
    import re
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, RANK="0", LOCAL_RANK="0", WORLD_SIZE="1", LOCAL_WORLD_SIZE="1", MASTER_ADDR="127.0.0.1",
               MASTER_PORT="29881")
    for d in ("saved_models", "logs", "data"):
        (tmp_path / d).mkdir()
    if example == "resnet":
        cmd = [sys.executable, os.path.join(root, "examples", "resnet", "main.py"), "--synthetic", "--graph", "--num_epochs",
               "8", "--batch_size", "64", "--synthetic_size", "128", "--learning_rate", "0.02"]
        pat = r"Epoch: (\d+), Loss: ([0-9.]+)"
    else:
        cmd = [sys.executable, os.path.join(root, "examples", "unet", "train.py"), "--synthetic", "--graph", "--num_epochs",
               "8", "--batch_size", "4", "--synthetic_size", "20", "--image_size", "64", "--learning_rate", "0.001"]
        pat = r"Epoch (\d+) finished with loss: ([0-9.]+)"
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=str(tmp_path))
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    losses = [float(m[1]) for m in re.findall(pat, r.stdout)]
    print(example, losses)
    # (random labels / blobs: the set is memorised over the epochs; single epochs may go up)
    assert len(losses) == 8 and all(v == v for v in losses) and min(losses[4:]) < 0.9 * losses[0], losses
