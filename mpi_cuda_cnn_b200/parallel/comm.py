"""Python handle of the fused all-reduce + SGD communicator (include/b200cnn_comm.h)."""
import ctypes as C

import torch
import torch.distributed as dist

from .. import _native as N
from ..engine import _tensor_from_ptr
from .symm import Region

_ALGO = {"auto": N.ALGO_AUTO, "one_shot": N.ALGO_ONE_SHOT, "two_shot": N.ALGO_TWO_SHOT, "nvls": N.ALGO_NVLS,
         "ll_one_shot": N.ALGO_LL_ONE_SHOT, "ll_two_shot": N.ALGO_LL_TWO_SHOT}


class FusedComm:
    """One rank's end of the NVLink communicator over a flat buffer of ``flat_floats`` floats.

    The region [flags | params | grads0 | grads1] lives in peer-mapped memory; ``params`` /
    ``grads(parity)`` are zero-copy torch views of the local part.
    """

    def __init__(self, flat_floats, rank, world, device, group=None, prefer=None):
        self.lib = N.cuda()
        self.rank, self.world, self.device = rank, world, torch.device(device)
        self.flat_floats = int((flat_floats + 3) // 4 * 4)
        nbytes = self.lib.b200cnn_comm_region_bytes(self.flat_floats)
        self.region = Region(nbytes, rank, world, self.device, group=group, prefer=prefer)
        arr = (C.c_void_p * world)(*[C.c_void_p(p) for p in self.region.ptrs])
        mc = C.c_void_p(self.region.multicast_ptr) if self.region.multicast_ptr else C.c_void_p(0)
        with torch.cuda.device(self.device):
            self.handle = self.lib.b200cnn_comm_create(rank, world, arr, mc, self.flat_floats)
        if not self.handle:
            raise RuntimeError("comm_create: " + self.lib.b200cnn_last_error().decode())
        self.group = group
        if world > 1:
            dist.barrier(group=group)

    @property
    def has_multicast(self):
        return bool(self.lib.b200cnn_comm_has_multicast(self.handle))

    def params(self):
        return _tensor_from_ptr(self.lib.b200cnn_comm_params(self.handle), self.flat_floats, self.device)

    def grads(self, parity=0):
        return _tensor_from_ptr(self.lib.b200cnn_comm_grads(self.handle, parity), self.flat_floats, self.device)

    def set_algo(self, name):
        N.check(self.lib.b200cnn_comm_set_algo(self.handle, _ALGO[name]), "comm_set_algo")

    def _st(self, stream=None):
        return C.c_void_p((stream or torch.cuda.current_stream()).cuda_stream)

    def allreduce_sgd(self, lr, offset=0, count=None, algo="auto", parity=0, max_blocks=0, stream=None):
        count = self.flat_floats - offset if count is None else count
        N.check(self.lib.b200cnn_comm_allreduce_sgd(self.handle, offset, count, float(lr), _ALGO[algo], parity, max_blocks, self._st(stream)), "allreduce_sgd")

    def allreduce(self, offset=0, count=None, algo="auto", parity=0, max_blocks=0, stream=None):
        count = self.flat_floats - offset if count is None else count
        N.check(self.lib.b200cnn_comm_allreduce(self.handle, offset, count, _ALGO[algo], parity, max_blocks, self._st(stream)), "allreduce")

    def broadcast_params(self, root=0, stream=None):
        N.check(self.lib.b200cnn_comm_broadcast_params(self.handle, root, self._st(stream)), "broadcast_params")

    def barrier(self, stream=None):
        N.check(self.lib.b200cnn_comm_barrier(self.handle, self._st(stream)), "comm_barrier")

    def wait_weights(self, stream=None):
        """Enqueue the wait for the peers' shares of the last two-shot push updates (see include/b200cnn_comm.h)."""
        N.check(self.lib.b200cnn_comm_wait_weights(self.handle, self._st(stream)), "comm_wait_weights")

    def error(self):
        return self.lib.b200cnn_comm_error(self.handle)

    def close(self):
        if getattr(self, "handle", None):
            self.lib.b200cnn_comm_destroy(self.handle)
            self.handle = None
