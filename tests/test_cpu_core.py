"""CPU-only unit tests: store, bucket planner, reducer state machine, sampler parity, state_dict parity, conv lowering
helpers.  (No GPU needed; the native host library is built by build.py / __graft_entry__.build().)"""
import ctypes as C
import threading

import pytest
import torch


@pytest.fixture(scope="module", autouse=True)
def _built():
    import build

    build.build()


def test_tcp_store_set_get_add_wait():
    from deeplearning_mpi_b200.parallel.store import TCPStore

    master = TCPStore("127.0.0.1", 0, True, timeout_s=10)
    client = TCPStore("127.0.0.1", master.port, False, timeout_s=10)
    master.set("k", b"hello")
    assert client.get("k") == b"hello"
    assert client.add("ctr", 3) == 3 and master.add("ctr", 4) == 7
    assert not client.check("missing") and client.check("k")
    with pytest.raises(TimeoutError):
        client.get("missing", timeout_s=0.2)
    # a blocked get is released by a later set from another client
    got = {}
    t = threading.Thread(target=lambda: got.setdefault("v", client.get("late", timeout_s=5)))
    t.start()
    master.set("late", b"x" * 100000)
    t.join(5)
    assert got["v"] == b"x" * 100000
    # barrier + all_gather across two "ranks"
    res = [None, None]

    def rank(r, st):
        res[r] = st.all_gather("g", r, 2, f"r{r}".encode())
        st.barrier("b", r, 2)

    ts = [threading.Thread(target=rank, args=(0, master)), threading.Thread(target=rank, args=(1, client))]
    [x.start() for x in ts]
    [x.join(10) for x in ts]
    assert res[0] == res[1] == [b"r0", b"r1"]
    client.close()
    master.close()


def test_bucket_planner_matches_torch():
    import torch.distributed as tdist
    import torchvision

    from deeplearning_mpi_b200 import _native as N

    model = torchvision.models.resnet18(weights=None)
    params = list(reversed(list(model.parameters())))
    sizes = [p.numel() * 4 for p in params]
    arr = (C.c_int64 * len(sizes))(*sizes)
    out = (C.c_int * len(sizes))()
    nb = N.host().dmpi_plan_buckets(arr, len(sizes), 1 << 20, 25 << 20, out)
    ref, _ = tdist._compute_bucket_assignment_by_size(params, [1 << 20, 25 << 20])
    mine = [[i for i, b in enumerate(out) if b == k] for k in range(nb)]
    assert mine == [list(b) for b in ref]


def test_reducer_state_machine_in_order_launch():
    from deeplearning_mpi_b200 import _native as N

    host = N.host()
    bucket_of = (C.c_int * 6)(0, 0, 1, 1, 1, 2)
    r = host.dmpi_reducer_create(6, bucket_of, 3)
    out = (C.c_int * 3)()
    host.dmpi_reducer_reset(r)
    # bucket 2 completes first but must wait for buckets 0 and 1
    assert host.dmpi_reducer_mark_ready(r, 5, out, 3) == 0
    assert host.dmpi_reducer_mark_ready(r, 2, out, 3) == 0
    assert host.dmpi_reducer_mark_ready(r, 0, out, 3) == 0
    assert host.dmpi_reducer_mark_ready(r, 1, out, 3) == 1 and out[0] == 0
    assert host.dmpi_reducer_mark_ready(r, 3, out, 3) == 0
    assert host.dmpi_reducer_pending(r) == 2
    assert host.dmpi_reducer_mark_ready(r, 4, out, 3) == 2 and list(out)[:2] == [1, 2]
    assert host.dmpi_reducer_pending(r) == 0
    assert host.dmpi_reducer_mark_ready(r, 4, out, 3) == -1  # double mark
    order = (C.c_int * 6)()
    assert host.dmpi_reducer_ready_order(r, order, 6) == 6 and list(order) == [5, 2, 0, 1, 3, 4]
    host.dmpi_reducer_reset(r)
    assert host.dmpi_reducer_pending(r) == 3
    host.dmpi_reducer_destroy(r)


@pytest.mark.parametrize("n,world,drop_last,shuffle", [(103, 4, False, True), (103, 4, True, True), (10, 3, False, False),
                                                       (3, 8, False, True), (50000, 8, False, True)])
def test_sampler_matches_torch(n, world, drop_last, shuffle):
    from torch.utils.data.distributed import DistributedSampler as TorchSampler

    from deeplearning_mpi_b200.parallel.sampler import DistributedSampler

    ds = list(range(n))
    for rank in range(world):
        for epoch in (0, 3):
            a = DistributedSampler(ds, num_replicas=world, rank=rank, shuffle=shuffle, seed=7, drop_last=drop_last)
            b = TorchSampler(ds, num_replicas=world, rank=rank, shuffle=shuffle, seed=7, drop_last=drop_last)
            a.set_epoch(epoch)
            b.set_epoch(epoch)
            assert list(a) == list(b) and len(a) == len(b)


def test_state_dict_parity_with_reference_models():
    import os
    import sys

    import torchvision

    from deeplearning_mpi_b200.models import UNet, resnet18, resnet50

    t = torchvision.models.resnet18(weights=None)
    t.fc = torch.nn.Linear(512, 10)
    m = resnet18(num_classes=10)
    assert list(m.state_dict().keys()) == list(t.state_dict().keys()) and len(m.state_dict()) == 122
    assert [tuple(v.shape) for v in m.state_dict().values()] == [tuple(v.shape) for v in t.state_dict().values()]
    m.load_state_dict(t.state_dict())
    x = torch.randn(2, 3, 32, 32)
    m.eval(), t.eval()
    assert torch.allclose(m(x), t(x), atol=1e-5)
    assert list(resnet50().state_dict().keys()) == list(torchvision.models.resnet50(weights=None).state_dict().keys())
    ref_dir = "/root/reference/pytorch/unet"
    if os.path.isdir(ref_dir):
        sys.path.insert(0, ref_dir)
        from model import UNet as RefUNet

        r = RefUNet(out_classes=1)
        u = UNet(out_classes=1)
        assert list(u.state_dict().keys()) == list(r.state_dict().keys()) and len(u.state_dict()) == 136
        u.load_state_dict(r.state_dict())
        u.eval(), r.eval()
        xi = torch.randn(1, 3, 32, 32)
        assert torch.allclose(u(xi), r(xi), atol=1e-5)
    with pytest.raises(ValueError):
        UNet(up_sample_mode="nearest")
    assert UNet(out_classes=2, up_sample_mode="bilinear")(torch.randn(1, 3, 32, 32)).shape == (1, 2, 32, 32)


def test_cpu_arena_and_optimizers_match_torch():
    """On CPU the arena re-homes params/grads and the optimizers follow torch.optim step for step."""
    import copy

    import deeplearning_mpi_b200.nn as dnn
    from deeplearning_mpi_b200 import optim
    from deeplearning_mpi_b200.models import resnet18
    from deeplearning_mpi_b200.utils.arena import ensure_arena

    torch.manual_seed(0)
    ours = resnet18(num_classes=10)
    ref = copy.deepcopy(ours)
    # eval-mode BatchNorm keeps the comparison well conditioned (training-mode BN with 4 samples amplifies the 1e-7
    # differences between channels_last and contiguous CPU conv kernels); gradients still flow to every parameter
    ours.eval()
    ref.eval()
    arena = ensure_arena(ours)
    assert arena.flat_p.numel() >= sum(p.numel() for p in ours.parameters())
    o1 = optim.SGD(ours.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-5)
    o2 = torch.optim.SGD(ref.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-5)
    x, y = torch.randn(4, 3, 32, 32), torch.randint(0, 10, (4,))
    for _ in range(3):
        o1.zero_grad()
        o2.zero_grad()
        dnn.cross_entropy(ours(x), y).backward()
        torch.nn.functional.cross_entropy(ref(x), y).backward()
        o1.step()
        o2.step()
    for (n, a), (_, b) in zip(ours.named_parameters(), ref.named_parameters()):
        assert torch.allclose(a, b, atol=1e-4, rtol=1e-3), n
    # Adam + clip
    a1 = optim.Adam(ours.parameters(), lr=1e-3)
    a2 = torch.optim.Adam(ref.parameters(), lr=1e-3)
    for _ in range(2):
        a1.zero_grad()
        a2.zero_grad()
        dnn.cross_entropy(ours(x), y).backward()
        torch.nn.functional.cross_entropy(ref(x), y).backward()
        n1 = dnn.clip_grad_norm_(ours.parameters(), 1.0)
        n2 = torch.nn.utils.clip_grad_norm_(ref.parameters(), 1.0)
        assert torch.allclose(n1, n2, rtol=1e-4)
        a1.step()
        a2.step()
    for (n, a), (_, b) in zip(ours.named_parameters(), ref.named_parameters()):
        assert torch.allclose(a, b, atol=2e-3, rtol=1e-2), n


def test_choose_box_and_dice_and_dataset(tmp_path):
    import numpy as np
    from PIL import Image

    from deeplearning_mpi_b200.ops.conv import choose_box, smallc_geometry
    from deeplearning_mpi_b200.utils.data import CarvanaDataset
    from deeplearning_mpi_b200.utils.metrics import dice_scores

    for (w, h, n) in [(56, 56, 256), (28, 28, 256), (7, 7, 256), (512, 512, 4), (14, 14, 3)]:
        bw, bh, bn = choose_box(w, h, n)
        assert 1 <= bw * bh * bn <= 128
        ew, eh, en = choose_box(w, h, n, rows=64, exact=True)
        assert ew * eh * en == 64
    assert smallc_geometry(224, 224, 7, 7, 2, 3) == (112, 112, 230, 230)
    d = dice_scores(torch.tensor([[[1, 0], [0, 0]], [[0, 0], [0, 0]]]), torch.tensor([[[1, 1], [0, 0]], [[0, 0], [0, 0]]]))
    assert abs(d[0] - 2 / 3) < 1e-6 and d[1] == 1.0
    (tmp_path / "images").mkdir()
    (tmp_path / "masks").mkdir()
    rng = np.random.default_rng(0)
    for i in range(3):
        Image.fromarray(rng.integers(0, 255, (20, 30, 3), dtype=np.uint8)).save(tmp_path / "images" / f"im{i}.png")
        Image.fromarray(((rng.random((20, 30)) > 0.5) * 255).astype(np.uint8)).save(tmp_path / "masks" / f"im{i}.png")
    ds = CarvanaDataset(str(tmp_path / "images"), str(tmp_path / "masks"), scale=0.5)
    item = ds[0]
    assert item["image"].shape == (3, 10, 15) and item["mask"].shape == (10, 15)
    assert item["image"].max() <= 1.0 and set(item["mask"].unique().tolist()) <= {0.0, 1.0}
    assert ds.mask_values == [0, 255] and len(ds) == 3
    # colour masks (one packed value per colour), a suffix, the parallel palette scan and a ready-made palette
    from deeplearning_mpi_b200.utils.data import SegmentationFolder, scan_palette

    (tmp_path / "rgb_i").mkdir()
    (tmp_path / "rgb_m").mkdir()
    colours = np.array([[0, 0, 0], [255, 0, 0], [0, 0, 255]], dtype=np.uint8)
    want = {}
    for i in range(9):
        lab = rng.integers(0, 3, (8, 12))
        want[f"s{i}"] = lab
        Image.fromarray(rng.integers(0, 255, (8, 12), dtype=np.uint8)).save(tmp_path / "rgb_i" / f"s{i}.png")
        Image.fromarray(colours[lab]).save(tmp_path / "rgb_m" / f"s{i}_mask.png")
    pal = scan_palette(sorted((tmp_path / "rgb_m").glob("*.png")), workers=2)
    assert pal.tolist() == [0, 255, 255 << 16]
    ds2 = SegmentationFolder(str(tmp_path / "rgb_i"), str(tmp_path / "rgb_m"), 1.0, "_mask", palette=pal)
    it = ds2[4]
    assert it["image"].shape == (1, 8, 12) and it["image"].max() <= 1.0
    # packed order: black 0 < blue 255 < red 255 << 16  ->  class indices 0, 1, 2
    remap = np.array([0, 2, 1])
    assert np.array_equal(ds2.class_index_mask(colours[want["s4"]]), remap[want["s4"]])
    assert np.array_equal(it["mask"].numpy(), (want["s4"] > 0).astype(np.float32))
    with pytest.raises(FileNotFoundError):
        SegmentationFolder(str(tmp_path / "rgb_i"), str(tmp_path / "rgb_m"), 1.0, "")  # suffix missing: no masks match
    assert SegmentationFolder.shared(str(tmp_path / "images"), str(tmp_path / "masks"), 0.5).mask_values == [0, 255]


def test_conv_planner_invariants():
    """Host-side tile planning of the implicit GEMM (ops/conv.py) over every image size the models can produce: the
    patch-mode tile must fit the 128 accumulator rows and the kernel's operand ring, and the N-tile / CTA-pair choice
    must stay inside what igemm.cu instantiates (BN 64 / 128 / 256, pairs only for BN >= 128)."""
    from deeplearning_mpi_b200.ops import conv as CV

    for ow in list(range(1, 140)) + [160, 224, 256, 384, 512, 1000]:
        for oh in (1, 2, 7, 14, 28, 56, 57, 112, 512):
            t = CV.choose_patch_tile(ow, oh, 1)
            if t is None:
                assert ow < 8
                continue
            tw, th, eff = t
            pw = tw + 2
            assert 8 <= tw <= 128 and 1 <= th <= oh and (th - 1) * pw + tw <= 128
            tiles = -(-ow // tw) * -(-oh // th)
            assert abs(eff - ow * oh / (tiles * 128.0)) < 1e-9 and 0 < eff <= 1.0
    for n_out in (8, 64, 65, 128, 192, 256, 512, 1000, 2048):
        for ktot in (64, 128, 256, 576, 4608):
            bn = CV._pick_bn(n_out, ktot)
            assert bn in (64, 128, 256) and (bn == 64) == (n_out <= 64)
            # the 256-wide tile: exact multiples, or widths such as 192 that pad to the same columns with 128-wide tiles
            assert bn != 256 or n_out % 256 == 0 or -(-n_out // 256) * 256 <= -(-n_out // 128) * 128
            cg = CV._pick_cg(bn, ktot)
            assert cg in (1, 2) and (cg == 1 or bn >= 128)
    ring = {(64, 1): 192, (64, 2): 180, (128, 1): 160, (256, 1): 144, (128, 2): 144, (256, 2): 160}  # KiB, igemm.cu IgemmCfg
    for bn_, cg_ in ring:   # every candidate the autotuner may force fits its ring as well
        for c_in in (64, 192):
            t = CV._patch_geom(3, 3, 1, 1, c_in, 512, 512, bn_, cg_)
            if t is not None:
                pw = t[0] + 2
                rows = max(128 + 2 * pw + 2, pw * (t[1] + 2))
                assert 2 * (-(-rows * 128 // 1024) * 1024) + 2 * (bn_ * 64 * 2 // cg_) <= ring[(bn_, cg_)] * 1024
    for c_in in (64, 128, 256):
        for n_out in (64, 128, 256):
            for ow, oh in ((56, 56), (28, 28), (128, 128), (256, 256), (512, 512), (20, 12)):
                bn = CV._pick_bn(n_out, 9 * c_in)
                cg = CV._pick_cg(bn, 9 * c_in)
                t = CV._patch_ok(3, 3, 1, 1, c_in, ow, oh, n_out, cg)
                if t is None:
                    continue
                pw = t[0] + 2
                rows = max(128 + 2 * pw + 2, pw * (t[1] + 2))
                patch_bytes = -(-rows * 128 // 1024) * 1024
                tap_bytes = bn * 64 * 2 // cg
                assert 2 * patch_bytes + 2 * tap_bytes <= ring[(bn, cg)] * 1024
                assert t[2] >= CV._PATCH_MIN_EFF
    # everything that is not a stride-1 pad-1 3x3 on whole 64-channel blocks stays on the per-tap path
    assert CV._patch_ok(1, 1, 1, 0, 64, 56, 56) is None and CV._patch_ok(3, 3, 2, 1, 64, 56, 56) is None
    assert CV._patch_ok(3, 3, 1, 1, 32, 56, 56) is None


def test_fault_injector_spec():
    """DMPI_FAULT parsing (utils/runtime.py::FaultInjector): inactive by default, fires once on the named rank / step,
    rejects unknown fields."""
    from deeplearning_mpi_b200.utils.runtime import FaultInjector

    assert not FaultInjector("").active
    f = FaultInjector("rank=1,step=2,kind=raise")
    assert f.active and (f.rank, f.step, f.kind) == (1, 2, "raise")
    for _ in range(5):
        f.tick(0)  # another rank: counts steps, never fires
    f = FaultInjector("rank=1,step=2,kind=raise")
    f.tick(1), f.tick(1)
    with pytest.raises(RuntimeError, match="injected failure on rank 1 at step 2"):
        f.tick(1)
    f.tick(1)  # later steps run normally
    s = FaultInjector("rank=0,step=0,stall=0.01")
    assert s.kind == "stall" and s.stall_s == 0.01
    s.tick(0)
    with pytest.raises(ValueError):
        FaultInjector("rank=1,when=now")
    with pytest.raises(ValueError):
        FaultInjector("rank=1,step=1,kind=explode")


def test_reducer_launch_order_rebuild():
    """Bucket launch order re-derived from the observed gradient-ready order: a bucket holding a late
    gradient no longer blocks the buckets behind it; the order is a permutation and can be reset."""
    import ctypes as C

    from deeplearning_mpi_b200 import _native as N

    host = N.host()
    bucket_of = (C.c_int * 6)(0, 0, 1, 1, 2, 2)
    r = host.dmpi_reducer_create(6, bucket_of, 3)
    out = (C.c_int * 3)()
    launched = []

    def mark(i):
        n = host.dmpi_reducer_mark_ready(r, i, out, 3)
        assert n >= 0
        launched.extend(out[j] for j in range(n))

    # iteration 0: parameter 1 (bucket 0) arrives last -> with the initial order nothing can go out before it
    host.dmpi_reducer_reset(r)
    for i in (0, 2, 3, 4, 5):
        mark(i)
    assert launched == []
    mark(1)
    assert launched == [0, 1, 2]
    assert host.dmpi_reducer_pending(r) == 0
    assert host.dmpi_reducer_rebuild(r) == 3
    host.dmpi_reducer_launch_order(r, out, 3)
    assert [out[i] for i in range(3)] == [1, 2, 0]
    # iteration 1, same arrival order: buckets 1 and 2 leave as soon as they are complete
    launched.clear()
    host.dmpi_reducer_reset(r)
    mark(0); mark(2); mark(3)
    assert launched == [1]
    mark(4); mark(5)
    assert launched == [1, 2]
    mark(1)
    assert launched == [1, 2, 0]
    # incomplete iteration: rebuild refuses; bad permutations are rejected; identity can be restored
    host.dmpi_reducer_reset(r)
    mark(0)
    assert host.dmpi_reducer_rebuild(r) == -1
    assert host.dmpi_reducer_set_launch_order(r, (C.c_int * 3)(0, 0, 1), 3) == -2
    assert host.dmpi_reducer_set_launch_order(r, (C.c_int * 3)(0, 1, 2), 3) == 0
    host.dmpi_reducer_destroy(r)


def test_frozen_parameters_stay_out_of_buckets_and_updates():
    """requires_grad=False parameters sit behind the trainable ones in the arena: not bucketed, not updated."""
    from deeplearning_mpi_b200 import optim
    from deeplearning_mpi_b200.utils.arena import ensure_arena

    torch.manual_seed(0)
    m = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.ReLU(), torch.nn.Linear(16, 4))
    m[0].weight.requires_grad_(False)
    frozen_before = m[0].weight.detach().clone()
    arena = ensure_arena(m)
    assert arena.n_trainable == 3 and len(arena.slots) == 4 and arena.slots[-1].param is m[0].weight
    assert arena.total_trainable <= arena.slots[-1].offset
    opt = optim.SGD([p for p in m.parameters() if p.requires_grad], lr=0.1, momentum=0.9, weight_decay=1e-2)
    for _ in range(2):
        opt.zero_grad()
        m(torch.randn(5, 8)).square().mean().backward()
        opt.step()
    assert torch.equal(m[0].weight, frozen_before) and m[0].weight.grad is None
    assert not torch.equal(m[2].weight, torch.zeros_like(m[2].weight))


def test_cpp_autograd_hooks_drive_the_reducer():
    """csrc/hooks/autograd_hooks.cpp on the CPU: C++ post-accumulate-grad hooks re-home fresh gradient tensors into
    their bucket views, mark the reducer core ready and run the finalize callback at the end of backward."""
    import ctypes as C

    from deeplearning_mpi_b200 import _native as N

    ext = N.torch_hooks()
    if ext is None:
        pytest.skip("dmpi_torch_hooks.so not built")
    host = N.host()
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(4, 8), torch.nn.Tanh(), torch.nn.Linear(8, 2))
    params = list(reversed(list(net.parameters())))          # bucket-fill order: reverse registration
    names = [f"p{i}" for i in range(len(params))]
    bucket_of = (C.c_int * 4)(0, 0, 1, 1)
    red = host.dmpi_reducer_create(4, bucket_of, 2)
    views = [torch.zeros_like(p) for p in params]
    ctx = ext.create(red, 0, N.fn_addr(host, "dmpi_reducer_mark_ready"), N.fn_addr(host, "dmpi_reducer_pending"), 0, 2)
    flags = torch.zeros(4, dtype=torch.uint8)
    ext.install(ctx, params, views, names, flags)
    x = torch.randn(6, 4)
    # not armed (no_sync): gradients are re-homed, nothing is marked
    net(x).square().sum().backward()
    assert ext.completed(ctx) == 0 and ext.take_launched(ctx) == 0
    assert all(p.grad.data_ptr() == v.data_ptr() for p, v in zip(params, views))
    # armed: both buckets launch, finalize runs once
    for p in params:
        p.grad = None
    host.dmpi_reducer_reset(red)
    ext.arm(ctx, True)
    net(x).square().sum().backward()
    assert ext.completed(ctx) == 1 and ext.take_launched(ctx) == 2 and host.dmpi_reducer_pending(red) == 0
    ref = torch.nn.Sequential(torch.nn.Linear(4, 8), torch.nn.Tanh(), torch.nn.Linear(8, 2))
    ref.load_state_dict(net.state_dict())
    ref(x).square().sum().backward()
    for v, q in zip(views, reversed(list(ref.parameters()))):
        assert torch.allclose(v, q.grad)
    # a parameter owned by a fused kernel is never published by its hook: the pass ends with a bucket incomplete
    flags[3] = 1
    host.dmpi_reducer_reset(red)
    ext.arm(ctx, True)
    with pytest.raises(RuntimeError, match="incomplete"):
        net(x).square().sum().backward()
    host.dmpi_reducer_destroy(red)


def test_tile_candidates_and_splitk_planner():
    """Pure-Python planning helpers of ops/conv.py: the candidate list the autotuner / force_config draw from, and the
    split-K chooser (every split must own at least one K block, the grid must fit the device)."""
    from deeplearning_mpi_b200.ops import conv as CV

    c64 = CV.tile_candidates(64, 3, 3, 1, 1, 64, 56, 56)
    assert {(bn, cg) for bn, cg, _ in c64} == {(64, 1), (64, 2)} and any(pt is not None for _, _, pt in c64)
    wide = CV.tile_candidates(512, 1, 1, 1, 0, 256, 14, 14)
    assert {(bn, cg) for bn, cg, _ in wide} == {(64, 1), (64, 2), (128, 1), (128, 2), (256, 1), (256, 2)}
    assert all(pt is None for _, _, pt in wide)                      # no patch mode for 1x1
    assert all(bn <= 128 for bn, _, _ in CV.tile_candidates(128, 3, 3, 1, 1, 128, 28, 28))
    with CV.force_config(128, 2, False):
        assert CV._forced == (128, 2, False) and CV._pick_splitk(256, 36, 4) is None   # forcing disables split-K
    assert CV._forced is None
    old = CV._SPLITK
    CV._SPLITK = True
    try:
        for n_out, num_kb, m_tiles in [(256, 36, 4), (512, 72, 1), (64, 16, 2), (128, 18, 16), (256, 17, 3)]:
            sk = CV._pick_splitk(n_out, num_kb, m_tiles)
            assert sk is not None
            bn, splits = sk
            units = m_tiles * -(-n_out // bn)
            per = -(-num_kb // splits)
            assert 2 <= splits <= 8 and units * splits <= 148 and per * (splits - 1) < num_kb
        assert CV._pick_splitk(512, 72, 98) is None      # enough tiles already
        assert CV._pick_splitk(64, 9, 2) is None         # reduction too short
    finally:
        CV._SPLITK = old
    assert CV._m_tiles(2, 2, 128, False) == 4 and CV._m_tiles(1, 1, 128, True) == 1
