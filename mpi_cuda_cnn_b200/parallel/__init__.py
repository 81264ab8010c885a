"""Data parallelism over NVLink: one process per GPU.

``launcher``  rank/device discovery, process group, NVLink topology map, a spawn helper
              (replaces mpirun + MPI_Init/Comm_rank/Comm_size, /root/reference/cnnmpi.c:419-422)
``symm``      peer-mapped "regions" (torch symmetric memory with NVLS multicast, or CUDA IPC)
``comm``      the fused all-reduce + SGD communicator (csrc/comm/allreduce_sgd.cu)
``trainer``   the data-parallel training loop (fused path and the NCCL + elementwise baseline)
"""
from .launcher import init_distributed, spawn, topology  # noqa: F401
from .comm import FusedComm  # noqa: F401
from .trainer import DataParallelTrainer  # noqa: F401
