"""Data-parallel training loop (configuration 4; replaces /root/reference/CUDAMPI.c:338-482).

Two update paths over the same engine:
  fused     per-layer fused all-reduce + 1/world + SGD kernels over NVLink, launched from inside the
            captured step so they overlap the rest of backward (the product)
  nccl      torch.distributed all_reduce on the flat gradient, then a scale+SGD elementwise kernel
            (the designated baseline: what BASELINE.md compares against)
Replicas start from rank 0's parameters (broadcast over NVLink), shard the data contiguously by rank
like cnnmpi.c:456-458, and rank 0 evaluates (cnnmpi.c:521).
"""
import torch
import torch.distributed as dist

from ..engine import Engine
from ..ops import functional as F200
from .comm import FusedComm


class DataParallelTrainer:
    def __init__(self, rank, world, device, layers=None, batch=256, mode="fused", seed=0, algo="auto", prefer_region=None):
        assert mode in ("fused", "nccl")
        self.rank, self.world, self.device, self.mode, self.batch = rank, world, torch.device(device), mode, batch
        self.engine = Engine(layers, max_batch=batch, device=self.device)
        self.comm = None
        if mode == "fused":
            self.comm = FusedComm(self.engine.flat_floats, rank, world, self.device, prefer=prefer_region)
            if algo == "nvls" and not self.comm.has_multicast:
                if rank == 0:
                    print("[b200cnn] no NVLS multicast mapping on this box: using the two-shot P2P kernel instead")
                algo = "two_shot"
            if algo != "auto":
                self.comm.set_algo(algo)
            self.engine.attach_comm(self.comm)
        # Same seed everywhere AND an explicit broadcast of rank 0's parameters.
        self.engine.init_params(seed=seed)
        self.sync_params(0)

    def sync_params(self, root=0):
        """Make every replica a copy of rank ``root``'s parameters (over NVLink on the fused path).  Called after the
        initialisation and again after ANY rank-local change of the parameters, e.g. a checkpoint loaded on rank 0
        (the reference never synchronises its replicas at all: SURVEY D7)."""
        if self.world <= 1:
            return
        torch.cuda.synchronize()
        dist.barrier()  # nobody is still training on the old parameters, root's new ones are in place
        if self.comm is not None:
            self.comm.broadcast_params(root)
        else:
            dist.broadcast(self.flat_params(), src=root)
        torch.cuda.synchronize()
        dist.barrier()

    def check_comm(self):
        """Raise if a kernel of the fused communicator ever gave up waiting for a peer (it then skipped its update)."""
        if self.comm is not None and self.comm.error():
            raise RuntimeError(f"rank {self.rank}: the fused communicator timed out waiting for a peer; the replicas are no longer in sync")

    def flat_params(self):
        from ..engine import _tensor_from_ptr

        return _tensor_from_ptr(self.engine.lib.b200cnn_engine_params(self.engine.handle), self.engine.flat_floats, self.device)

    def flat_grads(self):
        from ..engine import _tensor_from_ptr

        return _tensor_from_ptr(self.engine.lib.b200cnn_engine_grads(self.engine.handle), self.engine.flat_floats, self.device)

    def train_step(self, images, labels, rate, use_graph=True):
        """One mini-batch of ``batch`` samples per rank; rate is the learning rate (cnn.c:446)."""
        lr = rate / images.shape[0]
        if self.mode == "fused":
            # 1/world is applied inside the fused kernel.
            self.engine.train_step(images, labels, lr, use_graph=use_graph)
        else:
            # the baseline gets the same graph-captured forward + backward as the product; only the update differs
            self.engine.forward_backward(images, labels, use_graph=use_graph)
            g = self.flat_grads()
            if self.world > 1:
                dist.all_reduce(g)
            F200.sgd_update_(self.flat_params(), g, lr / self.world, zero_grad=False)

    def param_checksum(self):
        return float(self.flat_params().double().sum().item())

    def close(self):
        self.engine.close()
        if self.comm is not None:
            self.comm.close()
