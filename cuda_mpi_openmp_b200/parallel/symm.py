"""Peer-memory groups.

``DistPeerGroup``  — one rank per process (torch.distributed): buffers are cudaMalloc'ed
by libb200labs, their CUDA-IPC handles are all-gathered through the process group and
every peer's buffer is mapped into this process, so kernels dereference peer memory
directly over NVLink.
``LocalPeerGroup`` — ``world`` virtual ranks inside one process on one GPU; "peer"
pointers are simply the other virtual ranks' buffers.  Used to test the multi-rank
kernels on a single GPU (the kernels run phase by phase so none waits on another).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional

import numpy as np
import torch

from .. import _native as N

N.register("b200_dev_malloc", C.c_int, [C.POINTER(C.c_void_p), C.c_size_t])
N.register("b200_dev_free", C.c_int, [C.c_void_p])
N.register("b200_ipc_export", C.c_int, [C.c_void_p, C.c_void_p])
N.register("b200_ipc_open", C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)])
N.register("b200_ipc_close", C.c_int, [C.c_void_p])
N.register("b200_peer_can_access", C.c_int, [C.c_int, C.c_int])
N.register("b200_enable_peer_access", C.c_int, [C.c_int])
N.register("b200_symm_barrier", C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_void_p])
N.register("b200_symm_workspace_bytes", C.c_size_t, [])
N.register("b200_symm_offset", C.c_size_t, [C.c_char_p])


class _RawCuda:
    """Exposes a raw device allocation to torch through __cuda_array_interface__."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 3, "strides": None}


def _wrap(ptr: int, nbytes: int, device: torch.device) -> torch.Tensor:
    return torch.as_tensor(_RawCuda(ptr, nbytes), device=device)


class SymmetricBuffer:
    """``nbytes`` on every rank at peer-mapped addresses.  ``local`` is this rank's
    memory as a uint8 tensor; ``peer_ptrs[r]`` is rank r's buffer as seen from here;
    ``ptrs_dev`` is the same table on the device (what the kernels take)."""

    def __init__(self, group: "PeerGroup", nbytes: int, rank: int, local_ptr: int,
                 peer_ptrs: List[int], opened: List[int]):
        self.group, self.nbytes, self.rank = group, nbytes, rank
        self.local_ptr, self.peer_ptrs, self._opened = local_ptr, peer_ptrs, opened
        self.local = _wrap(local_ptr, nbytes, group.device)
        self.ptrs_dev = torch.tensor(peer_ptrs, dtype=torch.int64, device=group.device)

    def view(self, dtype: torch.dtype, *shape) -> torch.Tensor:
        n = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
        return self.local[:n].view(dtype).reshape(*shape)

    def peer_address(self, r: int, byte_offset: int = 0) -> int:
        return self.peer_ptrs[r] + byte_offset

    def peer_view(self, r: int, dtype: torch.dtype, n: int, byte_offset: int = 0) -> torch.Tensor:
        """``n`` elements of rank r's buffer as a tensor on THIS rank's device (a peer mapping:
        reading it moves the bytes over NVLink).  For checks and small copies."""
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        return _wrap(self.peer_ptrs[r] + byte_offset, nbytes, self.group.device).view(dtype)

    def free(self):
        """Collective over the group: every rank first closes its mappings of the peers' memory,
        then - after a barrier - releases its own.  (Freeing memory that a peer still has mapped,
        and then handing out the same address again, is how a re-allocation right after a free
        came back with `invalid argument`.)"""
        L = N.lib()
        had_peers = bool(self._opened)
        for p in self._opened:
            N.check(L.b200_ipc_close(C.c_void_p(p)), "b200_ipc_close")
        self._opened = []
        if had_peers or getattr(self.group, "world", 1) > 1:
            self.group.barrier()
        if self.local_ptr:
            N.check(L.b200_dev_free(C.c_void_p(self.local_ptr)), "b200_dev_free")
            self.local_ptr = 0


def _dev_malloc(nbytes: int) -> int:
    p = C.c_void_p()
    N.check(N.lib().b200_dev_malloc(C.byref(p), nbytes), "b200_dev_malloc")
    return int(p.value)


class PeerGroup:
    rank: int
    world: int
    device: torch.device

    def alloc(self, nbytes: int) -> SymmetricBuffer:  # pragma: no cover - interface
        raise NotImplementedError

    def barrier(self) -> None:
        raise NotImplementedError

    def workspace(self) -> SymmetricBuffer:
        """The SymmWorkspace (flags, lab3 slots, sort samples/cursor) of this group."""
        if getattr(self, "_ws", None) is None:
            self._ws = self.alloc(int(N.lib().b200_symm_workspace_bytes()))
            self._epoch = 0
            self.barrier()
        return self._ws

    def next_epoch(self) -> int:
        self._epoch += 1
        return self._epoch

    @staticmethod
    def ws_offset(field: str) -> int:
        off = int(N.lib().b200_symm_offset(field.encode()))
        if off == 2 ** 64 - 1:
            raise KeyError(field)
        return off

    def ws_error(self) -> int:
        ws = self.workspace()
        off = self.ws_offset("error")
        return int(ws.local[off: off + 8].view(torch.int64).item())


class DistPeerGroup(PeerGroup):
    """All ranks of a torch.distributed process group on one NVSwitch node."""

    def __init__(self, device: Optional[torch.device] = None, pg=None):
        import torch.distributed as dist

        self.dist, self.pg = dist, pg
        self.rank, self.world = dist.get_rank(pg), dist.get_world_size(pg)
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self._ws = None

    def alloc(self, nbytes: int) -> SymmetricBuffer:
        L = N.lib()
        nbytes = (int(nbytes) + 255) // 256 * 256
        with torch.cuda.device(self.device):
            local = _dev_malloc(nbytes)
            handle = (C.c_ubyte * 64)()
            N.check(L.b200_ipc_export(C.c_void_p(local), handle), "b200_ipc_export")
            mine = bytes(handle)
            gathered: List[Optional[bytes]] = [None] * self.world
            self.dist.all_gather_object(gathered, (mine, os.getpid()), group=self.pg)
            peers, opened = [], []
            for r, (h, pid) in enumerate(gathered):
                if r == self.rank:
                    peers.append(local)
                    continue
                p = C.c_void_p()
                buf = (C.c_ubyte * 64).from_buffer_copy(h)
                N.check(L.b200_ipc_open(buf, C.byref(p)), f"b200_ipc_open(rank {r})")
                peers.append(int(p.value))
                opened.append(int(p.value))
        return SymmetricBuffer(self, nbytes, self.rank, local, peers, opened)

    def barrier(self) -> None:
        torch.cuda.synchronize(self.device)
        self.dist.barrier(group=self.pg)

    def device_barrier(self) -> None:
        """Barrier enqueued on the current stream: signalling over peer memory only."""
        ws = self.workspace()
        N.check(N.lib().b200_symm_barrier(N.ptr(ws.ptrs_dev), self.rank, self.world, self.next_epoch(),
                                          N.current_stream_ptr()), "b200_symm_barrier")

    def all_reduce_max(self, value: float) -> float:
        t = torch.tensor([value], dtype=torch.float64, device=self.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.pg)
        return float(t.item())


class SoloPeerGroup(PeerGroup):
    """A single rank without torch.distributed (bench/tests at N = 1)."""

    def __init__(self, device: torch.device | str = "cuda:0"):
        self.rank, self.world, self.device = 0, 1, torch.device(device)
        self.dist, self.pg = None, None
        self._ws = None

    def alloc(self, nbytes: int) -> SymmetricBuffer:
        nbytes = (int(nbytes) + 255) // 256 * 256
        with torch.cuda.device(self.device):
            p = _dev_malloc(nbytes)
        return SymmetricBuffer(self, nbytes, 0, p, [p], [])

    def barrier(self) -> None:
        torch.cuda.synchronize(self.device)

    def all_reduce_max(self, value: float) -> float:
        return float(value)


class LocalPeerGroup:
    """``world`` virtual ranks in one process on one device (tests, single-GPU runs)."""

    def __init__(self, world: int, device: torch.device | str = "cuda:0"):
        self.world, self.device = world, torch.device(device)
        self.ranks = [_VirtualRank(self, r) for r in range(world)]

    def alloc(self, nbytes: int) -> List[SymmetricBuffer]:
        nbytes = (int(nbytes) + 255) // 256 * 256
        with torch.cuda.device(self.device):
            ptrs = [_dev_malloc(nbytes) for _ in range(self.world)]
        return [SymmetricBuffer(self.ranks[r], nbytes, r, ptrs[r], list(ptrs), []) for r in range(self.world)]

    def workspaces(self) -> List[SymmetricBuffer]:
        if getattr(self, "_ws", None) is None:
            self._ws = self.alloc(int(N.lib().b200_symm_workspace_bytes()))
            self._epoch = 0
        return self._ws

    def next_epoch(self) -> int:
        self._epoch += 1
        return self._epoch


class _VirtualRank(PeerGroup):
    def __init__(self, parent: LocalPeerGroup, rank: int):
        self.parent, self.rank, self.world, self.device = parent, rank, parent.world, parent.device

    def barrier(self) -> None:
        torch.cuda.synchronize(self.device)


def init_distributed(backend: Optional[str] = None) -> DistPeerGroup:
    """Rendezvous from the torchrun environment (RANK / LOCAL_RANK / WORLD_SIZE /
    MASTER_ADDR / MASTER_PORT) and return the peer group of all ranks."""
    import torch.distributed as dist

    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        dist.init_process_group(backend=backend or "nccl", device_id=torch.device("cuda", local_rank))
    return DistPeerGroup(torch.device("cuda", local_rank))
