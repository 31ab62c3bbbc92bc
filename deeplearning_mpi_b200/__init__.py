"""deeplearning_mpi_b200 — a Blackwell (sm_100a) native data-parallel training framework.

Capabilities mirror the PyTorch DDP examples of unlikeghost/DeepLearning-MPI (hello_world send/recv,
ResNet, UNet) but everything below the ``nn.Module`` surface is ours: TCP-store rendezvous, a C++
reducer, NVLink peer-to-peer collectives, tcgen05/TMA convolution kernels, fused BatchNorm/ReLU,
fused optimizers.  See DESIGN.md.
"""
__version__ = "0.1.0"

from . import _native  # noqa: F401
