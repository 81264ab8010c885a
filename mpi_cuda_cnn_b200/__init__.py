"""mpi_cuda_cnn_b200 — a Blackwell-native data-parallel CNN trainer.

Same capabilities as AnselObergfell/MPI-CUDA-CNN (serial CPU trainer, multi-process CPU trainer,
single-GPU trainer, multi-GPU data-parallel trainer for an MNIST-shaped CNN), rebuilt for B200:
hand-written sm_100a kernels behind a C ABI (``include/b200cnn.h``), a native training engine that
replays one CUDA graph per step, and a fused all-reduce + SGD kernel that reads peer GPUs'
gradients directly over NVLink (``include/b200cnn_comm.h``).  PyTorch is the tensor / process-group
framework and the numerical reference layer.

Layout::

    models/    network topologies, the PyTorch reference model (oracle + NCCL baseline)
    ops/       kernel wrappers (torch.autograd Functions), CPU engine binding
    parallel/  one-process-per-GPU launcher, symmetric memory, fused / NCCL data parallelism
    utils/     IDX files, synthetic data, glibc rand, timing, clocks
    engine.py  the native training engine (flagship path)
    cli.py     the four entrypoints: cnn, cnnmpi, cudacnn, cudampi
"""
from . import _native  # noqa: F401
from .engine import Engine, mnist_layers, pool_layers  # noqa: F401

__version__ = "0.1.0"
