"""ResNet data-parallel trainer — the capability of ref:pytorch/resnet/main.py (and, with ``--variant legacy``,
of ref:pytorch/resnet/resnet.py: evaluate/save *before* the epoch's training, batch 32, no per-step loss read).

Same CLI: --num_epochs --batch_size --learning_rate --random_seed --model_dir --model_filename --resume.
ResNet-18 with a 10-class head on CIFAR-10 (32x32), one process per GPU, DistributedSampler sharding,
SGD(momentum 0.9, weight decay 1e-5), cross-entropy, LOCAL_RANK-0 evaluation + checkpoint every 10 epochs.
Everything under the nn.Module surface runs on this framework (NVLink all-reduce, tcgen05 convolutions, ...).

    torchrun --nproc_per_node 8 --nnodes 1 --master_addr 127.0.0.1 --master_port 29500 examples/resnet/main.py --synthetic

``--graph`` replays the whole iteration as one CUDA graph (the speed bench.py reports; 3-4x on this launch-bound
configuration).  ``--synthetic`` replaces CIFAR-10 (which needs a download, examples/resnet/download.py) with random data of its shape;
``--arch resnet50 --image_size 224`` selects the BASELINE.json configuration.
"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _bootstrap  # noqa: F401,E402

import numpy as np  # noqa: E402
import torch  # noqa: E402
from torch.utils.data import DataLoader  # noqa: E402

import deeplearning_mpi_b200.nn as dnn  # noqa: E402
from deeplearning_mpi_b200 import models, optim  # noqa: E402
from deeplearning_mpi_b200.parallel import dist  # noqa: E402
from deeplearning_mpi_b200.parallel.ddp import DistributedDataParallel  # noqa: E402
from deeplearning_mpi_b200.parallel.sampler import DistributedSampler  # noqa: E402
from deeplearning_mpi_b200.utils.checkpoint import load_checkpoint, save_checkpoint  # noqa: E402
from deeplearning_mpi_b200.utils.data import CudaPrefetcher, SyntheticImages  # noqa: E402
from deeplearning_mpi_b200.utils.metrics import top1_accuracy  # noqa: E402
from deeplearning_mpi_b200.utils.runtime import TrainConfig, set_random_seeds, torchrun_env  # noqa: E402

LOCAL_RANK, WORLD_SIZE, WORLD_RANK = torchrun_env()


def build_model(cfg, arch: str, device):
    model = getattr(models, arch)(num_classes=10 if arch in ("resnet18", "resnet34") else 1000)
    model = model.to(device)
    ddp_model = DistributedDataParallel(model, device_ids=[LOCAL_RANK], output_device=LOCAL_RANK)
    if cfg.resume:
        load_checkpoint(ddp_model, cfg.model_filepath, LOCAL_RANK)
    return ddp_model


def build_loaders(cfg, args):
    if cfg.synthetic:
        shape = (3, args.image_size, args.image_size)
        classes = 10 if args.arch in ("resnet18", "resnet34") else 1000
        train_set = SyntheticImages(args.synthetic_size, shape, classes, seed=cfg.random_seed)
        test_set = SyntheticImages(max(args.synthetic_size // 5, cfg.batch_size), shape, classes, seed=cfg.random_seed + 1)
        workers = 0
    else:
        import torchvision
        import torchvision.transforms as transforms

        transform = transforms.Compose([
            transforms.RandomCrop(32, padding=4), transforms.RandomHorizontalFlip(), transforms.ToTensor(),
            transforms.Normalize((0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010)),
        ])
        # data must be prefetched (examples/resnet/download.py): downloading is not multi-process safe
        train_set = torchvision.datasets.CIFAR10(root="./data", train=True, download=False, transform=transform)
        test_set = torchvision.datasets.CIFAR10(root="./data", train=False, download=False, transform=transform)
        workers = 8 if args.variant == "legacy" else 15
    sampler = DistributedSampler(dataset=train_set)
    pin = torch.cuda.is_available() and args.variant != "legacy"
    train_loader = DataLoader(train_set, batch_size=cfg.batch_size, sampler=sampler, num_workers=workers, pin_memory=pin)
    test_bs = 128 if args.variant == "legacy" else cfg.batch_size
    test_loader = DataLoader(test_set, batch_size=test_bs, shuffle=False, num_workers=workers, pin_memory=pin)
    return train_loader, test_loader


def evaluate_and_save(ddp_model, device, test_loader, cfg, epoch):
    accuracy = top1_accuracy(ddp_model, device, test_loader)
    save_checkpoint(ddp_model, cfg.model_filepath)
    print("-" * 75)
    print("Epoch: {}, Accuracy: {}".format(epoch, accuracy))
    print("-" * 75)


def run(cfg, args):
    on_gpu = cfg.backend != "gloo"
    device = torch.device(f"cuda:{LOCAL_RANK}") if on_gpu else torch.device("cpu")
    ddp_model = build_model(cfg, args.arch, device)
    train_loader, test_loader = build_loaders(cfg, args)
    criterion = dnn.CrossEntropyLoss()
    optimizer = optim.SGD(ddp_model.parameters(), lr=cfg.learning_rate, momentum=0.9, weight_decay=1e-5)

    def train_step(inputs, labels):
        optimizer.zero_grad()
        loss = criterion(ddp_model(inputs), labels)
        loss.backward()
        optimizer.step()
        return loss

    if cfg.graph and on_gpu:
        from deeplearning_mpi_b200.utils.graph import LazyGraphedStep

        train_step = LazyGraphedStep(train_step)
    for epoch in range(cfg.num_epochs):
        if args.variant == "legacy" and epoch % 10 == 0 and LOCAL_RANK == 0:
            evaluate_and_save(ddp_model, device, test_loader, cfg, epoch)
        print("Local Rank: {}, Epoch: {}, Training ...".format(LOCAL_RANK, epoch))
        ddp_model.train()
        total_loss = []
        batches = CudaPrefetcher(train_loader, device) if on_gpu else train_loader
        for index, data in enumerate(batches):
            if cfg.steps_per_epoch and index >= cfg.steps_per_epoch:
                break
            if args.variant != "legacy":
                print(f"Local Rank: {LOCAL_RANK}, index: {index}", end="\r")
            inputs, labels = data[0].to(device), data[1].to(device)
            loss = train_step(inputs, labels)
            if args.variant != "legacy":
                total_loss.append(loss.item())
        if total_loss:
            print("Local Rank: {}, Epoch: {}, Loss: {}".format(LOCAL_RANK, epoch, np.mean(total_loss)))
        if args.variant != "legacy" and epoch % 10 == 0 and LOCAL_RANK == 0:
            evaluate_and_save(ddp_model, device, test_loader, cfg, epoch)
        print(f"Epoch {epoch} completed")


def init_processes(cfg, args):
    dist.init_process_group(cfg.backend)
    try:
        run(cfg, args)
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    defaults = TrainConfig(num_epochs=100, batch_size=128, learning_rate=0.1, random_seed=0, model_dir="saved_models",
                           model_filename="resnet_distributed.pth")
    parser = TrainConfig.parser(defaults)
    parser.add_argument("--arch", default="resnet18", choices=["resnet18", "resnet34", "resnet50", "resnet101", "resnet152"])
    parser.add_argument("--image_size", type=int, default=32)
    parser.add_argument("--synthetic_size", type=int, default=2048, help="samples in the synthetic training set")
    parser.add_argument("--variant", default="main", choices=["main", "legacy"],
                        help="'legacy' = behaviour of the reference's older resnet.py")
    ns = parser.parse_args()
    if ns.variant == "legacy" and "--batch_size" not in sys.argv:
        ns.batch_size = 32
    config = TrainConfig.from_args(ns)
    if not os.path.isdir(config.model_dir):
        raise OSError(f"The model directory '{config.model_dir}' does not exist. Please create it before running the script.")
    set_random_seeds(config.random_seed)
    init_processes(config, ns)
