"""Single-process CIFAR-10 prefetch so the trainers can use download=False (capability of
ref:pytorch/resnet/download.py).  Needs network access; the benchmark machines have none — use --synthetic there."""
import torchvision

if __name__ == "__main__":
    for train in (True, False):
        torchvision.datasets.CIFAR10(root="./data", train=train, download=True)
    print("CIFAR-10 is in ./data")
