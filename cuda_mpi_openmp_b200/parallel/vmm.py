"""VMM symmetric buffers with an optional NVSwitch multicast (NVLS) mapping.

The native side (csrc/vmm.cu) creates, exports, imports, maps and binds; this module
moves the POSIX file descriptors between the ranks (unix sockets, SCM_RIGHTS — a
descriptor cannot travel through ``all_gather_object``) and keeps the ranks in the
order a multicast object needs (every device added before any memory is bound).

    buf = vmm.alloc(group, nbytes)          # collective over the group
    buf.local            uint8 tensor over this rank's memory
    buf.peer_ptrs[r]     rank r's memory, mapped here (unicast, like the IPC buffers)
    buf.mc_ptr           multicast address (0 when the system has no NVLS)

(No counterpart in the reference: it has no multi-GPU code at all.)
"""
from __future__ import annotations

import ctypes as C
import os
import socket
import struct
import time
import uuid
from typing import Dict, List

import torch

from .. import _native as N
from .symm import PeerGroup, _wrap

N.register("b200_vmm_query", C.c_int, [C.POINTER(C.c_int), C.POINTER(C.c_int)])
N.register("b200_vmm_granularity", C.c_int, [C.c_int, C.POINTER(C.c_size_t)])
N.register("b200_vmm_create", C.c_int, [C.c_size_t, C.POINTER(C.c_uint64)])
N.register("b200_vmm_release", C.c_int, [C.c_uint64])
N.register("b200_vmm_export_fd", C.c_int, [C.c_uint64, C.POINTER(C.c_int)])
N.register("b200_vmm_import_fd", C.c_int, [C.c_int, C.POINTER(C.c_uint64)])
N.register("b200_vmm_map", C.c_int, [C.c_uint64, C.c_size_t, C.POINTER(C.c_void_p)])
N.register("b200_vmm_unmap", C.c_int, [C.c_void_p, C.c_size_t])
N.register("b200_mc_create", C.c_int, [C.c_int, C.c_size_t, C.POINTER(C.c_uint64)])
N.register("b200_mc_add_device", C.c_int, [C.c_uint64])
N.register("b200_mc_bind", C.c_int, [C.c_uint64, C.c_uint64, C.c_size_t])
N.register("b200_mc_unbind", C.c_int, [C.c_uint64, C.c_size_t])
N.register("b200_selftest_pingpong", C.c_int,
           [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p])
N.register("b200_selftest_multimem", C.c_int,
           [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_uint64, C.c_void_p])


def query() -> Dict[str, bool]:
    """{'vmm': fd-exportable cuMemCreate works, 'multicast': the device is behind NVLS}."""
    v, m = C.c_int(0), C.c_int(0)
    N.check(N.lib().b200_vmm_query(C.byref(v), C.byref(m)), "b200_vmm_query")
    return {"vmm": bool(v.value), "multicast": bool(m.value)}


def granularity(world: int) -> int:
    g = C.c_size_t(0)
    N.check(N.lib().b200_vmm_granularity(world, C.byref(g)), "b200_vmm_granularity")
    return int(g.value)


def _create(nbytes: int) -> int:
    h = C.c_uint64(0)
    N.check(N.lib().b200_vmm_create(nbytes, C.byref(h)), "b200_vmm_create")
    return int(h.value)


def _export(handle: int) -> int:
    fd = C.c_int(-1)
    N.check(N.lib().b200_vmm_export_fd(handle, C.byref(fd)), "b200_vmm_export_fd")
    return int(fd.value)


def _import(fd: int) -> int:
    h = C.c_uint64(0)
    N.check(N.lib().b200_vmm_import_fd(fd, C.byref(h)), "b200_vmm_import_fd")
    return int(h.value)


def _map(handle: int, nbytes: int) -> int:
    p = C.c_void_p()
    N.check(N.lib().b200_vmm_map(handle, nbytes, C.byref(p)), "b200_vmm_map")
    return int(p.value)


def exchange_fds(group: PeerGroup, outgoing: Dict[int, List[int]]) -> Dict[int, List[int]]:
    """Send ``outgoing[dest]`` (file descriptors) to each other rank; returns what the
    other ranks sent here, by sender.  Collective.  Every rank listens on an abstract
    unix socket, the names go round by all_gather_object, then each rank connects and
    sends — the kernel queues the messages, so nobody needs a second thread.

    Abstract sockets are reachable by any process in the network namespace, so a message
    is only accepted when it (a) comes from the pid the sender announced through the
    process group (SO_PEERCRED), (b) carries the random token this rank handed out in
    the same all_gather, (c) names a rank that has not delivered yet.  Anything else is
    dropped (its descriptors closed) and the rank keeps accepting until the deadline.
    Descriptors received before a failure are closed on the way out."""
    world, rank = group.world, group.rank
    if world == 1:
        return {}
    name = "\0b200vmm-%d-%s" % (os.getpid(), uuid.uuid4().hex)
    token = uuid.uuid4().bytes                       # 16 random bytes; senders must echo them
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    incoming: Dict[int, List[int]] = {}
    ok = False
    try:
        srv.bind(name)
        srv.listen(world)
        cards: list = [None] * world
        group.dist.all_gather_object(cards, (name, os.getpid(), token), group=group.pg)
        for dest in range(world):
            if dest == rank:
                continue
            fds = outgoing.get(dest, [])
            with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as c:
                c.connect(cards[dest][0])
                header = cards[dest][2] + b"%4d%4d" % (rank, len(fds))
                if fds:
                    socket.send_fds(c, [header], fds)
                else:
                    c.sendall(header)
        deadline = time.monotonic() + 60.0
        while len(incoming) < world - 1:
            left = deadline - time.monotonic()
            if left <= 0:
                raise RuntimeError(f"fd exchange: rank {rank} heard from {sorted(incoming)} only")
            srv.settimeout(left)
            conn, _addr = srv.accept()
            with conn:
                conn.settimeout(max(0.1, deadline - time.monotonic()))
                fds: List[int] = []
                try:
                    cred = conn.getsockopt(socket.SOL_SOCKET, socket.SO_PEERCRED, struct.calcsize("3i"))
                    pid, uid, _gid = struct.unpack("3i", cred)
                    msg, fds, _flags, _a = socket.recv_fds(conn, 24, 16)
                    fds = list(fds)
                    sender, count = int(msg[16:20]), int(msg[20:24])    # ValueError on a garbled header
                    good = (len(msg) == 24 and msg[:16] == token and 0 <= sender < world and sender != rank
                            and sender not in incoming and uid == os.getuid() and pid == cards[sender][1]
                            and count == len(fds))
                except (ValueError, OSError, struct.error):
                    good = False
                if not good:
                    for fd in fds:
                        os.close(fd)
                    continue                             # not one of ours: keep listening
                incoming[sender] = fds
        group.dist.barrier(group=group.pg)
        ok = True
        return incoming
    finally:
        srv.close()
        if not ok:
            for fds in incoming.values():
                for fd in fds:
                    try:
                        os.close(fd)
                    except OSError:
                        pass


class VmmBuffer:
    """``nbytes`` (padded to the granularity) on every rank: local mapping, peer
    mappings, and — when available — the multicast mapping of all copies."""

    def __init__(self, group: PeerGroup, nbytes: int, handle: int, local_ptr: int, peer_ptrs: List[int],
                 peer_handles: List[int], mc_handle: int, mc_ptr: int):
        self.group, self.nbytes = group, nbytes
        self.handle, self.local_ptr, self.peer_ptrs = handle, local_ptr, peer_ptrs
        self._peer_handles, self.mc_handle, self.mc_ptr = peer_handles, mc_handle, mc_ptr
        self.local = _wrap(local_ptr, nbytes, group.device)
        self.ptrs_dev = torch.tensor(peer_ptrs, dtype=torch.int64, device=group.device)

    def view(self, dtype: torch.dtype, n: int) -> torch.Tensor:
        return self.local[: n * torch.empty((), dtype=dtype).element_size()].view(dtype)

    def free(self) -> None:
        L = N.lib()
        torch.cuda.synchronize(self.group.device)
        self.group.barrier()
        if self.mc_ptr:
            L.b200_vmm_unmap(C.c_void_p(self.mc_ptr), self.nbytes)
            L.b200_mc_unbind(self.mc_handle, self.nbytes)
            self.mc_ptr = 0
        self.group.barrier()
        if self.mc_handle:
            L.b200_vmm_release(self.mc_handle)
            self.mc_handle = 0
        for r, p in enumerate(self.peer_ptrs):
            if r != self.group.rank and p:
                L.b200_vmm_unmap(C.c_void_p(p), self.nbytes)
        for h in self._peer_handles:
            L.b200_vmm_release(h)
        self.peer_ptrs, self._peer_handles = [], []
        if self.local_ptr:
            del self.local
            L.b200_vmm_unmap(C.c_void_p(self.local_ptr), self.nbytes)
            L.b200_vmm_release(self.handle)
            self.local_ptr = 0


def alloc(group: PeerGroup, nbytes: int, multicast: bool = True, map_peers: bool = True) -> VmmBuffer:
    """Collective: a zero-filled VMM buffer on every rank of ``group`` (a DistPeerGroup or
    SoloPeerGroup).  With ``multicast`` and NVLS hardware the copies are also bound to one
    multicast object (``mc_ptr``); without the hardware ``mc_ptr`` is 0 — callers that need
    it must check."""
    world, rank = group.world, group.rank
    with torch.cuda.device(group.device):
        caps = query()
        if not caps["vmm"]:
            raise RuntimeError("VMM allocations with POSIX fd export are not supported on this device")
        use_mc = bool(multicast and caps["multicast"] and world > 1)   # a multicast object needs >= 2 devices
        if world > 1:
            flags: List[bool] = [False] * world
            group.dist.all_gather_object(flags, use_mc, group=group.pg)
            use_mc = all(flags)
        gran = granularity(world if use_mc else 1)
        size = (max(int(nbytes), 1) + gran - 1) // gran * gran
        handle = _create(size)
        local_ptr = _map(handle, size)
        _wrap(local_ptr, size, group.device).zero_()
        torch.cuda.synchronize(group.device)

        mc_handle = 0
        my_fd = _export(handle) if (world > 1 and map_peers) else -1
        mc_fd = -1
        if use_mc and rank == 0:
            h = C.c_uint64(0)
            N.check(N.lib().b200_mc_create(world, size, C.byref(h)), "b200_mc_create")
            mc_handle = int(h.value)
            if world > 1:
                mc_fd = _export(mc_handle)
        outgoing: Dict[int, List[int]] = {}
        for dest in range(world):
            if dest == rank:
                continue
            fds = [my_fd] if my_fd >= 0 else []
            if mc_fd >= 0:
                fds.append(mc_fd)
            outgoing[dest] = fds
        incoming = exchange_fds(group, outgoing)
        for fd in (my_fd, mc_fd):
            if fd >= 0:
                os.close(fd)

        peer_ptrs, peer_handles = [0] * world, []
        peer_ptrs[rank] = local_ptr
        for sender, fds in sorted(incoming.items()):
            fds = list(fds)
            if map_peers:
                fd = fds.pop(0)
                h = _import(fd)
                os.close(fd)
                peer_handles.append(h)
                peer_ptrs[sender] = _map(h, size)
            if use_mc and sender == 0:
                fd = fds.pop(0)
                mc_handle = _import(fd)
                os.close(fd)
            for fd in fds:
                os.close(fd)

        mc_ptr = 0
        if use_mc:
            N.check(N.lib().b200_mc_add_device(mc_handle), "b200_mc_add_device")
            group.barrier()  # every device must be added before any rank binds
            N.check(N.lib().b200_mc_bind(mc_handle, handle, size), "b200_mc_bind")
            group.barrier()
            mc_ptr = _map(mc_handle, size)
        group.barrier()
    return VmmBuffer(group, size, handle, local_ptr, peer_ptrs, peer_handles, mc_handle, mc_ptr)


# ---- self-tests of the runtime (tools/symm_selftest.py, tests) --------------------
def pingpong(my_flag_ptr: int, peer_flag_ptr: int, initiator: bool, iters: int, base: int,
             device: torch.device, relaxed: bool = False) -> float:
    """Round-trip time (ns) of a flag bounced ``iters`` times between two ranks; both ranks
    call it at the same time, the initiator's number is the measurement."""
    out = torch.zeros(2, dtype=torch.int64, device=device)
    N.check(N.lib().b200_selftest_pingpong(C.c_void_p(my_flag_ptr), C.c_void_p(peer_flag_ptr), int(initiator),
                                           iters, base, int(relaxed), N.ptr(out), N.current_stream_ptr()),
            "b200_selftest_pingpong")
    ns, timed_out = out.tolist()
    if timed_out:
        raise RuntimeError("pingpong: peer did not answer within the timeout")
    return ns / iters


def multimem(mc_ptr: int, n: int, mode: str, value: int = 0, out: torch.Tensor | None = None) -> None:
    """One launch of the multimem self-test kernel over ``n`` uint64 elements at ``mc_ptr``:
    mode 'red' | 'ld_reduce' (needs ``out``, int64[n]) | 'st'."""
    code = {"red": 0, "ld_reduce": 1, "st": 2}[mode]
    N.check(N.lib().b200_selftest_multimem(C.c_void_p(mc_ptr), N.ptr(out) if out is not None else None, n, code,
                                           value, N.current_stream_ptr()), "b200_selftest_multimem")
