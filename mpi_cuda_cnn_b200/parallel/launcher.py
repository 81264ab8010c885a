"""Process / device discovery for one-process-per-GPU runs on a single NVSwitch box.

Replaces the reference's start-up: ``mpirun -np 8`` (/root/reference/Makefile:44) plus
MPI_Init / MPI_Comm_rank / MPI_Comm_size (cnnmpi.c:419-422), which never selected a GPU
(SURVEY D18).  Here rank r owns GPU r; the rendezvous is torch.distributed over 127.0.0.1.
"""
import os
import socket

import torch
import torch.distributed as dist


def env_rank_world():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", os.environ.get("RANK", "0")))


def init_distributed(backend=None):
    """Returns (rank, world, device).  Uses RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* when present."""
    rank, world, local = env_rank_world()
    cuda = torch.cuda.is_available()
    device = torch.device(f"cuda:{local}") if cuda else torch.device("cpu")
    if cuda:
        torch.cuda.set_device(device)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29512")
        backend = backend or ("nccl" if cuda else "gloo")
        kwargs = {"device_id": device} if (cuda and backend == "nccl") else {}
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kwargs)
    return rank, world, device


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _entry(rank, world, port, fn, args):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    fn(rank, world, *args)
    if dist.is_initialized():
        dist.destroy_process_group()


def spawn(fn, world, args=()):
    """Start ``world`` ranks of fn(rank, world, *args) on this host (our own 'mpirun -np')."""
    import torch.multiprocessing as mp

    mp.spawn(_entry, args=(world, free_port(), fn, args), nprocs=world, join=True)


def topology():
    """NVLink / peer-access map of the box: {'gpus': n, 'peer': [[bool]], 'nvlinks': [[active links]]}."""
    n = torch.cuda.device_count()
    info = {"gpus": n, "peer": [[i == j or torch.cuda.can_device_access_peer(i, j) for j in range(n)] for i in range(n)], "nvlinks": None}
    try:
        import pynvml

        pynvml.nvmlInit()
        links = []
        for i in range(n):
            h = pynvml.nvmlDeviceGetHandleByIndex(i)
            active = 0
            for l in range(18):
                try:
                    if pynvml.nvmlDeviceGetNvLinkState(h, l) == pynvml.NVML_FEATURE_ENABLED:
                        active += 1
                except pynvml.NVMLError:
                    break
            links.append(active)
        info["nvlinks"] = links
        pynvml.nvmlShutdown()
    except Exception:
        pass
    return info
