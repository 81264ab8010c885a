"""Peer-mapped device memory ("regions") for the fused communicator.

Preferred: torch symmetric memory (CUDA VMM; gives an NVLS multicast address when the fabric has
one).  Fallback: cudaMalloc + CUDA IPC handles exchanged with all_gather_object.  Either way the
result is: this rank's region pointer, every peer's pointer as mapped here, an optional multicast
pointer — all the native communicator needs.
"""
import ctypes as C
import os

import torch
import torch.distributed as dist

from .. import _native as N


class Region:
    def __init__(self, nbytes, rank, world, device, group=None, prefer=None):
        self.nbytes, self.rank, self.world, self.device = int(nbytes), rank, world, device
        self.ptrs = None
        self.multicast_ptr = 0
        self.kind = None
        self._keep = []
        prefer = prefer or os.environ.get("B200CNN_REGION", "symm")
        if world == 1:
            self._alloc_local()
        elif prefer == "symm":
            try:
                self._alloc_symm(group)
            except Exception as exc:  # pragma: no cover - depends on the box
                if rank == 0:
                    print(f"[b200cnn] symmetric memory unavailable ({type(exc).__name__}: {exc}); using CUDA IPC")
                self._alloc_ipc(group)
        else:
            self._alloc_ipc(group)

    def _alloc_local(self):
        t = torch.zeros(self.nbytes, dtype=torch.uint8, device=self.device)
        self._keep.append(t)
        self.ptrs = [t.data_ptr()]
        self.kind = "local"

    def _alloc_symm(self, group):
        import torch.distributed._symmetric_memory as symm

        group = group or dist.group.WORLD
        t = symm.empty(self.nbytes, dtype=torch.uint8, device=self.device)
        hdl = symm.rendezvous(t, group)
        t.zero_()
        torch.cuda.synchronize()
        self._keep += [t, hdl]
        self.ptrs = [int(p) for p in hdl.buffer_ptrs]
        try:
            self.multicast_ptr = int(hdl.multicast_ptr) if hdl.has_multicast_support(torch.device(self.device).type, torch.device(self.device).index or 0) else 0
        except Exception:
            self.multicast_ptr = int(getattr(hdl, "multicast_ptr", 0) or 0)
        self.kind = "symm"
        dist.barrier(group=group)

    def _alloc_ipc(self, group):
        lib = N.cuda()
        group = group or dist.group.WORLD
        base = lib.b200cnn_ipc_alloc(self.nbytes)
        if not base:
            raise RuntimeError("ipc_alloc: " + lib.b200cnn_last_error().decode())
        buf = C.create_string_buffer(64)
        N.check(lib.b200cnn_ipc_get_handle(base, buf), "ipc_get_handle")
        handles = [None] * self.world
        dist.all_gather_object(handles, buf.raw, group=group)
        ptrs = []
        for r in range(self.world):
            if r == self.rank:
                ptrs.append(int(base))
            else:
                p = lib.b200cnn_ipc_open(handles[r])
                if not p:
                    raise RuntimeError("ipc_open: " + lib.b200cnn_last_error().decode())
                ptrs.append(int(p))
        self.ptrs = ptrs
        self.kind = "ipc"
        self._ipc_base = base
        dist.barrier(group=group)

    def local_ptr(self):
        return self.ptrs[self.rank]
