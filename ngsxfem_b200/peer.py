"""Scalar exchange between the GPUs of one node over peer memory (include/ngsx_abi.h, section (e); csrc/ngsx_peer.cu): the sums
`Integrate()` all-reduces in the reference (cutint/python_cutint.cpp:66-70) are published into the peers' mailboxes by a kernel that
is part of the rank's captured step -- no collective library call, no host work per step.  torch.distributed is used once, at
set-up, to hand the CUDA IPC handles round."""
import ctypes

import torch

from . import _lib
from ._lib import NgsxError, check
from .ngs import _ptr, _stream

MAXVALS = 4


class PeerExchange:
    def __init__(self, rank: int, world: int, device, group=None):
        import torch.distributed as dist
        L = _lib.load()
        self.rank, self.world, self.device = rank, world, torch.device(device)
        self._L = L
        self.mailbox = ctypes.c_void_p()
        self._opened = []
        handle = (ctypes.c_ubyte * 64)()
        # every rank runs the same collectives whatever fails locally (a rank that raised early would leave the others waiting)
        err = None
        try:
            with torch.cuda.device(self.device):
                check(L.ngsx_peer_alloc(world, ctypes.byref(self.mailbox), handle))
        except Exception as exc:                                                           # noqa: BLE001
            err = exc
        handles = [bytes(handle) if err is None else None]
        if world > 1:
            handles = [None] * world
            dist.all_gather_object(handles, bytes(handle) if err is None else None, group=group)
        ptrs = []
        if err is None and all(h is not None for h in handles):
            try:
                for r in range(world):
                    if r == rank:
                        ptrs.append(self.mailbox.value)
                        continue
                    p = ctypes.c_void_p()
                    buf = (ctypes.c_ubyte * 64).from_buffer_copy(handles[r])
                    with torch.cuda.device(self.device):
                        check(L.ngsx_peer_open(buf, ctypes.byref(p)))
                    self._opened.append(p)
                    ptrs.append(p.value)
            except Exception as exc:                                                       # noqa: BLE001
                err = exc
        elif err is None:
            err = NgsxError(-3, "a peer could not allocate / export its mailbox")
        if world > 1:                                                                      # also the barrier: every mailbox is mapped before anyone publishes
            ok = torch.tensor([0 if err is not None else 1], device=self.device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            if int(ok.item()) == 0 and err is None:
                err = NgsxError(-3, "a peer could not map the mailboxes (CUDA IPC)")
        if err is not None:
            self.close()
            raise err
        self.peers = torch.tensor(ptrs, dtype=torch.int64, device=self.device)            # device array of the world mailbox addresses
        self.step = torch.zeros(1, dtype=torch.int64, device=self.device)
        self.prev = torch.zeros(MAXVALS, dtype=torch.float64, device=self.device)          # totals of the previous step
        self.last = torch.zeros(MAXVALS, dtype=torch.float64, device=self.device)

    def exchange(self, local: torch.Tensor) -> torch.Tensor:
        """publish this step's local sums (<= 4 doubles, device) to every rank; returns the buffer that holds the totals of the PREVIOUS
        step once the kernel has run (stream-ordered; capturable)"""
        n = local.numel()
        if n > MAXVALS or local.dtype != torch.float64 or not local.is_cuda:
            raise NgsxError(-1, "PeerExchange.exchange: up to 4 float64 values on the device")
        check(self._L.ngsx_peer_exchange(_ptr(local), n, self.rank, self.world, _ptr(self.peers), self.mailbox, _ptr(self.step), _ptr(self.prev),
                                         _stream()))
        return self.prev

    def collect_last(self, n: int) -> torch.Tensor:
        """totals of the last published step (waits on the device for the peers' values of that step)"""
        check(self._L.ngsx_peer_collect(self.mailbox, n, self.world, _ptr(self.step), _ptr(self.last), _stream()))
        return self.last[:n]

    def close(self):
        L = self._L
        torch.cuda.synchronize(self.device)
        for p in self._opened:
            L.ngsx_peer_close(p)
        self._opened = []
        if self.mailbox:
            L.ngsx_peer_free(self.mailbox)
            self.mailbox = ctypes.c_void_p()
