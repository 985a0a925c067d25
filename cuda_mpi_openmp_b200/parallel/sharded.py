"""The lab workloads sharded over the ranks of a PeerGroup.

======================  ===================================  ===========================
workload                fused (product)                      nccl (baseline)
======================  ===================================  ===========================
``ShardedRoberts``      stencil kernel loads the halo row    isend/irecv of the halo row,
(row bands)             from the next GPU's band in place    then the same kernel
``ShardedClassifier``   statistics kernel all-reduces the    moments kernel, NCCL
(row bands)             integer moments over peer memory     all_reduce, finalize kernel
``ShardedSort``         partition kernel stores keys into    local partition, counts +
(key shards)            the destination GPUs' buckets        keys by NCCL all_to_all
======================  ===================================  ===========================
(The reference has no multi-GPU code; both columns are new.  See csrc/mgpu_*.cu.)
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from .. import _native as N
from ..ops import classify as CL
from ..ops import sort as ST
from ..ops.roberts import roberts as roberts_op
from .symm import PeerGroup, SymmetricBuffer

N.register("b200_lab3_stats_allreduce", C.c_int,
           [C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
            C.c_int, C.c_int, C.c_uint64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p])
N.register("b200_lab3_stats_nvls", C.c_int,
           [C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
            C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p])
N.register("b200_lab3_lut_build_allgather", C.c_int,
           [C.POINTER(C.c_void_p), C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_int,
            C.c_void_p])
N.register("b200_lab3_apply_lut", C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p])
N.register("b200_lab3_coarse_bytes", C.c_size_t, [])
N.register("b200_lab3_lut2_bytes", C.c_size_t, [])
N.register("b200_lab3_lut_build2_allgather", C.c_int,
           [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_void_p])
N.register("b200_lab3_apply_lut2", C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p])
N.register("b200_sort_mgpu_splitters", C.c_int,
           [C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_int,
            C.c_void_p])
N.register("b200_sort_mgpu_scatter", C.c_int,
           [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
            C.c_size_t, C.c_void_p, C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_uint64, C.c_int,
            C.c_void_p])
N.register("b200_sort_mgpu_count", C.c_int,
           [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int, C.c_int, C.c_void_p])
N.register("b200_sort_devn", C.c_int,
           [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p])
N.register("b200_sort_ex", C.c_int,
           [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int, C.c_void_p, C.c_size_t,
            C.c_void_p])


def shard_rows(h: int, world: int) -> List[Tuple[int, int]]:
    """Contiguous, near-equal row bands [(y0, y1)] — every band non-empty when h >= world."""
    base, extra = divmod(h, world)
    out, y = [], 0
    for r in range(world):
        n = base + (1 if r < extra else 0)
        out.append((y, y + n))
        y += n
    return out


def _ptr_array(values: Sequence[int]):
    arr = (C.c_void_p * len(values))()
    for i, v in enumerate(values):
        arr[i] = v
    return arr


# ------------------------------------------------------------------------- lab1
class ShardedVectorSub:
    """lab1 over several GPUs: pure data parallelism — every rank subtracts its slice,
    nothing is exchanged (ref: lab1/src/main.cu:22-29 is a single-GPU grid-stride loop)."""

    def __init__(self, group: PeerGroup, n_total: int):
        from ..ops.vecsub import vecsub
        self._op = vecsub
        lo, hi = shard_rows(n_total, group.world)[group.rank]
        self.lo, self.hi = lo, hi
        dev = group.device
        self.a = torch.randn(hi - lo, dtype=torch.float64, device=dev)
        self.b = torch.randn(hi - lo, dtype=torch.float64, device=dev)
        self.c = torch.empty_like(self.a)
        self.launches_per_step = 1

    def step(self) -> torch.Tensor:
        return self._op(self.a, self.b, self.c)


# ------------------------------------------------------------------------- lab2
class ShardedRoberts:
    """Row-sharded Roberts filter.  The band lives in a symmetric buffer so that rank
    r-1's kernel can read this band's first row as its halo.

    Ordering contract: the fused kernel reads the NEIGHBOUR's band with plain loads and no
    in-kernel handshake, which is only safe while that band is not being written.  A caller
    that fills the bands once (the bench) needs one ``group.barrier()`` after the fill.  A
    caller that refills ``band`` between steps must order "every rank's upload finished"
    before any rank's ``step()``: pass ``refill_barrier=True`` and ``step()`` runs the
    stream-ordered cross-GPU barrier (``PeerGroup.device_barrier``, flags over peer memory,
    no host sync) in front of the kernel, or call ``group.device_barrier()`` yourself."""

    def __init__(self, group: PeerGroup, w: int, h_total: int, impl: str = "fused", variant: str = "auto",
                 refill_barrier: bool = False):
        assert impl in ("fused", "nccl")
        self.group, self.w, self.h_total, self.impl, self.variant = group, w, h_total, impl, variant
        self.refill_barrier = refill_barrier
        self.y0, self.y1 = shard_rows(h_total, group.world)[group.rank]
        self.h_local = self.y1 - self.y0
        self.sym_in: SymmetricBuffer = group.alloc(max(1, self.h_local) * w * 4)
        self.band = self.sym_in.view(torch.int32, self.h_local, w)
        self.out = torch.empty((self.h_local, w), dtype=torch.int32, device=group.device)
        self.halo = torch.empty(w, dtype=torch.int32, device=group.device) if impl == "nccl" else None
        self.launches_per_step = 1

    def free(self) -> None:
        """Release the symmetric band (call on every rank, after a barrier)."""
        self.band = None
        self.sym_in.free()

    def below_address(self) -> Optional[int]:
        r = self.group.rank
        return None if r == self.group.world - 1 else self.sym_in.peer_address(r + 1)

    def run_host(self, host_in: torch.Tensor, host_out: torch.Tensor, chunk_bytes: int = 256 << 20) -> torch.Tensor:
        """One end-to-end step through the product path.  This rank's band sits in pinned host
        memory (int32 [h_local, w]); it is copied into the symmetric device band in row
        chunks on a copy stream, each chunk is filtered as soon as it and the row after it
        are on the GPU, and the result is copied back on a third stream, so H2D, kernel
        and D2H of different chunks overlap.  The last chunk's kernel reads the row below
        the band straight from the next rank's GPU over NVLink, as ``step()`` does; two
        stream-ordered cross-GPU barriers (flags over peer memory, no host sync) bracket
        that: one after every rank's first chunk has landed (the neighbour's first row is
        valid), one after the last kernel (no rank refills its band while a neighbour may
        still be reading it).  Synchronises the calling stream before returning."""
        if self.impl != "fused":
            raise ValueError("run_host drives the fused (peer-memory halo) path")
        g, h, w = self.group, self.h_local, self.w
        src = host_in.view(torch.int32).reshape(h, w)
        dst = host_out.view(torch.int32).reshape(h, w)
        rows = max(1, min(h, chunk_bytes // (4 * w)))
        chunks = [(r0, min(h, r0 + rows)) for r0 in range(0, h, rows)]
        cur = torch.cuda.current_stream(g.device)
        if getattr(self, "_io_streams", None) is None:
            self._io_streams = (torch.cuda.Stream(device=g.device), torch.cuda.Stream(device=g.device))
        s_up, s_down = self._io_streams
        s_up.wait_stream(cur)
        s_down.wait_stream(cur)
        landed = []
        with torch.cuda.stream(s_up):
            for r0, r1 in chunks:
                self.band[r0:r1].copy_(src[r0:r1], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(s_up)
                landed.append(ev)
        peer_below = self.below_address()
        for k, (r0, r1) in enumerate(chunks):
            if k == 0 and g.world > 1:
                cur.wait_event(landed[0])
                g.device_barrier()
            cur.wait_event(landed[min(k + 1, len(chunks) - 1)])   # row r1 belongs to the next chunk
            below = self.band[r1] if r1 < h else peer_below
            roberts_op(self.band[r0:r1], self.out[r0:r1], below_row=below, variant=self.variant)
            done = torch.cuda.Event()
            done.record(cur)
            with torch.cuda.stream(s_down):
                s_down.wait_event(done)
                dst[r0:r1].copy_(self.out[r0:r1], non_blocking=True)
        if g.world > 1:
            g.device_barrier()
        cur.wait_stream(s_down)
        cur.synchronize()
        return host_out

    def step(self) -> torch.Tensor:
        g = self.group
        if self.impl == "fused":
            if self.refill_barrier and g.world > 1:
                g.device_barrier()          # neighbours' uploads are complete before our loads
            below = self.below_address()
        else:
            dist = g.dist
            ops = []
            if g.rank > 0:
                ops.append(dist.P2POp(dist.isend, self.band[0], g.rank - 1, group=g.pg))
            if g.rank < g.world - 1:
                ops.append(dist.P2POp(dist.irecv, self.halo, g.rank + 1, group=g.pg))
            if ops:
                for req in dist.batch_isend_irecv(ops):
                    req.wait()
            below = self.halo if g.rank < g.world - 1 else None
        return roberts_op(self.band, self.out, below_row=below, variant=self.variant)


# ------------------------------------------------------------------------- lab3
def fused_stats(img: torch.Tensor, w: int, y0: int, h_local: int, d_pts: torch.Tensor, d_off: torch.Tensor,
                nc: int, ws_ptrs_dev: torch.Tensor, rank: int, world: int, epoch: int,
                params: torch.Tensor, moments_out: Optional[torch.Tensor], phase_mask: int = 3) -> None:
    N.check(N.lib().b200_lab3_stats_allreduce(
        N.ptr(img), w, y0, h_local, N.ptr(d_pts), N.ptr(d_off), nc, N.ptr(ws_ptrs_dev), rank, world,
        epoch, N.ptr(params), N.ptr(moments_out), phase_mask, N.current_stream_ptr()),
        "b200_lab3_stats_allreduce")


def lut_build_allgather(lut_ptrs: Sequence[int], nc: int, params: torch.Tensor, ws_ptrs_dev: torch.Tensor,
                        rank: int, world: int, epoch: int, phase_mask: int = 3) -> None:
    N.check(N.lib().b200_lab3_lut_build_allgather(_ptr_array(lut_ptrs), nc, N.ptr(params), N.ptr(ws_ptrs_dev),
                                                  rank, world, epoch, phase_mask, N.current_stream_ptr()),
            "b200_lab3_lut_build_allgather")


def apply_lut(img: torch.Tensor, lut_addr: int, coarse: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Image pass against a finished colour table; ``coarse`` (uint8[b200_lab3_coarse_bytes()])
    turns on the two-level lookup (cell table in shared memory, see csrc/lab3_classify.cu)."""
    N.check(N.lib().b200_lab3_apply_lut(N.ptr(img), img.shape[0] * img.shape[1], C.c_void_p(lut_addr),
                                        N.ptr(coarse), N.current_stream_ptr()), "b200_lab3_apply_lut")
    return img


def lut_build2_allgather(lut_ptrs: Optional[Sequence[int]], mc_lut: int, nc: int, params: torch.Tensor,
                         ws_ptrs_dev: torch.Tensor, rank: int, world: int, epoch: int, phase_mask: int = 3) -> None:
    """Colour table + cell table built cooperatively, 16-byte stores: into every peer's buffer
    (``lut_ptrs``), or one multimem.st per piece on the multicast address ``mc_lut``."""
    arr = _ptr_array(lut_ptrs) if lut_ptrs is not None else None
    N.check(N.lib().b200_lab3_lut_build2_allgather(arr, C.c_void_p(mc_lut or 0), nc, N.ptr(params),
                                                   N.ptr(ws_ptrs_dev), rank, world, epoch, phase_mask,
                                                   N.current_stream_ptr()), "b200_lab3_lut_build2_allgather")


def apply_lut2(img: torch.Tensor, lut_addr: int) -> torch.Tensor:
    """Image pass against a table buffer that carries its cell table (b200_lab3_lut2_bytes())."""
    N.check(N.lib().b200_lab3_apply_lut2(N.ptr(img), img.shape[0] * img.shape[1], C.c_void_p(lut_addr),
                                         N.current_stream_ptr()), "b200_lab3_apply_lut2")
    return img


class ShardedClassifier:
    """Row-sharded lab3: global sample points, statistics all-reduced, local classify.
    With the fused implementation on several GPUs the 2^24-entry colour table is built
    cooperatively (each rank classifies 1/world of the colours and stores them into
    every rank's table over NVLink).  The table is rewritten every step; the in-kernel
    all-reduce of the next step's statistics is what orders a peer's rewrite after this
    rank's image pass, so always call ``step()`` (fit + classify), not classify alone."""

    def __init__(self, group: PeerGroup, w: int, h_total: int, classes, impl: str = "fused",
                 variant: str = "auto"):
        # impl: "fused" = product path: in-kernel all-reduce and cooperative table build over peer
        #         memory, through the NVSwitch (multimem.red / multimem.st on multicast buffers of
        #         parallel/vmm.py) when the system has NVLS, else with peer stores;
        #       "nvls"  = as "fused" but NVLS is required (raises without it);
        #       "p2p"   = as "fused" but never multicast (B200LABS_NVLS=0 does the same);
        #       "nccl"  = baseline: library all-reduce, every rank builds the whole table.
        import os
        assert impl in ("fused", "nccl", "nvls", "p2p")
        self.nvls = False
        self._acc_buf = self._lut_buf = None
        want_mc = (impl in ("fused", "nvls") and group.world > 1 and
                   (impl == "nvls" or os.environ.get("B200LABS_NVLS", "1") != "0"))
        if want_mc:
            from . import vmm
            try:
                self._acc_buf = vmm.alloc(group, 2 * 32 * 10 * 8, multicast=True, map_peers=False)
            except RuntimeError:
                self._acc_buf = None
            if self._acc_buf is not None and not self._acc_buf.mc_ptr:
                self._acc_buf.free()
                self._acc_buf = None
            self.nvls = self._acc_buf is not None
            if self.nvls:
                self._acc = self._acc_buf.view(torch.int64, 2 * 32 * 10)
                self._nvls_calls = 0
        if impl == "nvls" and not self.nvls:
            raise RuntimeError("NVLS multicast is not available on this system")
        if impl in ("nvls", "p2p"):
            impl = "fused"
        self.group, self.w, self.h_total, self.impl, self.variant = group, w, h_total, impl, variant
        self.y0, self.y1 = shard_rows(h_total, group.world)[group.rank]
        self.h_local = self.y1 - self.y0
        self.nc = len(classes)
        pts, off = CL.pack_points(classes, w=w, h=h_total)   # checked against the WHOLE image
        dev = group.device
        self.d_pts = torch.from_numpy(pts).to(dev) if pts.size else torch.zeros(2, dtype=torch.int32, device=dev)
        self.d_off = torch.from_numpy(off).to(dev)
        self.band = torch.empty((self.h_local, w), dtype=torch.int32, device=dev)
        self.params = torch.empty((self.nc, CL.N_PARAMS), dtype=torch.float64, device=dev)
        self.moments = torch.zeros((self.nc, CL.N_MOMENTS), dtype=torch.int64, device=dev)
        npix = self.h_local * w
        need = N.lib().b200_lab3_classify_scratch_bytes(npix, CL.CLASSIFY_VARIANTS[variant])
        self.scratch = torch.empty(need, dtype=torch.uint8, device=dev) if need else None
        self.ws = group.workspace()
        lut = need > 0
        # cooperative table build: any world size; the table buffer carries its cell table
        self.coop_lut = lut and impl == "fused" and group.world > 1
        self.sym_lut, self._lut_mc, self._lut_local = None, 0, 0
        if self.coop_lut:
            lut_bytes = int(N.lib().b200_lab3_lut2_bytes())
            if self.nvls:
                from . import vmm
                self._lut_buf = vmm.alloc(group, lut_bytes, multicast=True, map_peers=False)
                self._lut_mc, self._lut_local = self._lut_buf.mc_ptr, self._lut_buf.local_ptr
            if not self._lut_mc:
                self.sym_lut = group.alloc(lut_bytes)
                self._lut_local = self.sym_lut.local_ptr
        # kernels per step: stats(1) [+ finalize for nccl] + classify (lut: build [+ cells] + apply)
        self.launches_per_step = (1 if impl == "fused" else 2) + (2 if lut else 1)

    def fit(self) -> torch.Tensor:
        g = self.group
        if self.nvls:
            N.check(N.lib().b200_lab3_stats_nvls(
                N.ptr(self.band), self.w, self.y0, self.h_local, N.ptr(self.d_pts), N.ptr(self.d_off), self.nc,
                N.ptr(self.ws.ptrs_dev), C.c_void_p(self._acc_buf.mc_ptr), N.ptr(self._acc), g.rank,
                g.world, g.next_epoch(), self._nvls_calls & 1, N.ptr(self.params), N.ptr(self.moments),
                N.current_stream_ptr()), "b200_lab3_stats_nvls")
            self._nvls_calls += 1
        elif self.impl == "fused":
            fused_stats(self.band, self.w, self.y0, self.h_local, self.d_pts, self.d_off, self.nc,
                        self.ws.ptrs_dev, g.rank, g.world, g.next_epoch(), self.params, self.moments, 3)
        else:
            self.moments.zero_()
            CL.lab3_moments(self.band, self.d_pts, self.d_off, self.nc, self.moments, y0=self.y0)
            g.dist.all_reduce(self.moments, op=g.dist.ReduceOp.SUM, group=g.pg)
            CL.lab3_finalize(self.moments, self.params)
        return self.params

    def step(self) -> torch.Tensor:
        self.fit()
        if self.coop_lut:
            g = self.group
            lut_build2_allgather(self.sym_lut.peer_ptrs if self.sym_lut is not None else None, self._lut_mc,
                                 self.nc, self.params, self.ws.ptrs_dev, g.rank, g.world, g.next_epoch(), 3)
            return apply_lut2(self.band, self._lut_local)
        return CL.lab3_classify(self.band, self.params, variant=self.variant, scratch=self.scratch)

    def check(self) -> None:
        """Call at a point where the stream may be synchronised: raises if an in-kernel wait for a
        peer timed out during the steps so far (parameters and classes would then be garbage)."""
        torch.cuda.synchronize(self.group.device)
        if self.group.world > 1:
            raise_on_error(self.group, "ShardedClassifier")

    def free(self) -> None:
        if self.sym_lut is not None:
            self.sym_lut.free()
            self.sym_lut = None
        if self._lut_buf is not None:
            self._lut_buf.free()
            self._lut_buf = None
        if self._acc_buf is not None:
            self._acc = None
            self._acc_buf.free()
            self._acc_buf = None


# ------------------------------------------------------------------------- lab5
def mgpu_splitters(keys: torch.Tensor, samples_per_rank: int, ws_ptrs_dev: torch.Tensor, rank: int,
                   world: int, epoch: int, phase_mask: int = 3) -> None:
    N.check(N.lib().b200_sort_mgpu_splitters(
        N.ptr(keys), keys.numel(), ST.KEY_TYPES[keys.dtype], samples_per_rank, N.ptr(ws_ptrs_dev), rank,
        world, epoch, phase_mask, N.current_stream_ptr()), "b200_sort_mgpu_splitters")


def mgpu_scatter(keys: torch.Tensor, splitters_addr: int, bucket_ptrs: Sequence[int],
                 cursor_ptrs: Optional[Sequence[int]], capacity: int, ws: SymmetricBuffer, rank: int,
                 world: int, epoch: int, phase_mask: int = 3) -> None:
    N.check(N.lib().b200_sort_mgpu_scatter(
        N.ptr(keys), keys.numel(), ST.KEY_TYPES[keys.dtype], C.c_void_p(splitters_addr),
        _ptr_array(bucket_ptrs), _ptr_array(cursor_ptrs) if cursor_ptrs is not None else None, capacity,
        N.ptr(ws.ptrs_dev), _ptr_array(ws.peer_ptrs), rank, world, epoch, phase_mask,
        N.current_stream_ptr()), "b200_sort_mgpu_scatter")


def mgpu_count(keys: torch.Tensor, splitters_addr: int, capacity: int, ws: SymmetricBuffer, rank: int,
               world: int) -> None:
    N.check(N.lib().b200_sort_mgpu_count(N.ptr(keys), keys.numel(), ST.KEY_TYPES[keys.dtype],
                                         C.c_void_p(splitters_addr), capacity, N.ptr(ws.ptrs_dev), rank, world,
                                         N.current_stream_ptr()), "b200_sort_mgpu_count")


def sort_devn(bucket: torch.Tensor, tmp: torch.Tensor, count_addr: int, scratch: torch.Tensor) -> None:
    N.check(N.lib().b200_sort_devn(N.ptr(bucket), N.ptr(tmp), bucket.numel(), C.c_void_p(count_addr),
                                   ST.KEY_TYPES[bucket.dtype], N.ptr(scratch), scratch.numel(),
                                   N.current_stream_ptr()), "b200_sort_devn")


def samples_per_rank_for(world: int) -> int:
    """Samples each rank contributes to the splitter selection: as many as fit the 8192-entry
    sample area (the kernel pads the total to a power of two for its sorting network)."""
    return 8192 // world


ERR_TIMEOUT, ERR_SORT_OVERFLOW = 1, 2     # SymmError bits of the workspace's sticky error word


def raise_on_error(group: PeerGroup, what: str) -> int:
    """Reads this rank's sticky error word (a device-to-host copy: call at a point where the
    stream may be synchronised) and raises if an in-kernel wait for a peer timed out — after
    a timeout every number derived from that peer's data is garbage.  Returns the word."""
    err = group.ws_error()
    if err & ERR_TIMEOUT:
        raise RuntimeError(f"{what}: rank {group.rank} timed out waiting for a peer inside a kernel "
                           f"(error word {err:#x}); results of this step are invalid")
    return err


class ShardedSort:
    """Sample sort over the ranks (any world size up to 8).  After ``step()`` rank r holds the
    r-th slice of the globally sorted sequence in ``result()`` (= ``bucket[:count()]``).

    ``step()`` only enqueues work.  ``count()`` / ``result()`` are COLLECTIVE sync points: they
    read the error word of every rank, raise if a peer wait timed out, and if any
    destination bucket overflowed (the kernels never write out of bounds: they drop the excess
    and set the overflow bit) they grow the buckets on every rank and run the step again —
    doubling until nothing overflows, which must happen by capacity = all keys.  Keys are
    never lost; until ``count()``/``result()`` has returned, ``bucket`` must be treated as
    not final."""

    SLACK = 1.25

    def __init__(self, group: PeerGroup, n_local: int, dtype: torch.dtype = torch.int32, impl: str = "fused"):
        assert impl in ("fused", "nccl")
        self.group, self.n_local, self.dtype, self.impl = group, n_local, dtype, impl
        dev, world = group.device, group.world
        self.keys = torch.empty(n_local, dtype=dtype, device=dev)
        self.ws = group.workspace()
        self.samples = samples_per_rank_for(world)
        self._off = {k: group.ws_offset(k) for k in ("sort_splitters", "sort_recv_count", "sort_cursor", "error")}
        self.sym_bucket = None
        self._allocate(n_local if world == 1 else int(n_local * self.SLACK) + 8192)
        if impl == "nccl" and world > 1:
            self.send = torch.empty((world, n_local), dtype=dtype, device=dev)
            self.cursors = torch.zeros(world, dtype=torch.int64, device=dev)
            self.count_dev = torch.zeros(1, dtype=torch.int64, device=dev)
        # kernels per step: splitters + scatter + memset-free sort (hist, scan, 4 passes)
        # reserve=True (B200LABS_SORT_RESERVE=1): a count pass reserves one exact segment per
        # destination; default: every tile reserves its room with one remote atomic per
        # destination — measured faster (7.0 vs 2.0 + 6.4 ms for 2^30 keys per GPU, N = 2)
        import os
        self.reserve = os.environ.get("B200LABS_SORT_RESERVE", "0") == "1"
        self.launches_per_step = (6 if world == 1 else 9 if self.reserve else 8)
        self._count_host: Optional[int] = None
        self._settled = False
        self.retries = 0

    def _allocate(self, capacity: int) -> None:
        """(Re)allocates bucket, ping-pong buffer and sort scratch for ``capacity`` keys.  Collective."""
        g = self.group
        if self.sym_bucket is not None:
            self.bucket = self.tmp = self.scratch = None
            g.barrier()
            self.sym_bucket.free()
        self.capacity = int(capacity)
        self.sym_bucket = g.alloc(self.capacity * 4)
        self.bucket = self.sym_bucket.view(self.dtype, self.capacity)
        self.tmp = torch.empty(self.capacity, dtype=self.dtype, device=g.device)
        self.scratch = torch.empty(ST.sort_scratch_bytes(self.capacity, self.dtype), dtype=torch.uint8,
                                   device=g.device)

    def _settle(self) -> None:
        """Collective check of the step just run; grows the buckets and re-runs on overflow."""
        if self._settled:
            return
        g = self.group
        while True:
            torch.cuda.synchronize(g.device)
            err = raise_on_error(g, "ShardedSort") if g.world > 1 else 0
            overflow = g.all_reduce_max(1.0 if err & ERR_SORT_OVERFLOW else 0.0) > 0 if g.world > 1 else False
            if not overflow:
                break
            n_total = int(g.all_reduce_max(float(self.n_local))) * g.world      # upper bound on any bucket
            if self.capacity >= n_total:
                raise RuntimeError("ShardedSort: bucket overflow with capacity for every key (corrupt workspace?)")
            off = self._off["error"]                     # clear the sticky overflow bit, keep nothing else set
            self.ws.local[off: off + 8].view(torch.int64).zero_()
            self._allocate(min(2 * self.capacity, n_total))
            self.retries += 1
            self._count_host = None
            self._run()
        self._settled = True

    def step(self) -> None:
        self._count_host = None
        self._settled = False
        self._run()

    def _run(self) -> None:
        g, world, r = self.group, self.group.world, self.group.rank
        if world == 1:
            # out of place: pass 0 reads `keys`, the result lands in `bucket`
            N.check(N.lib().b200_sort_ex(N.ptr(self.keys), N.ptr(self.bucket), N.ptr(self.tmp), self.n_local,
                                         None, ST.KEY_TYPES[self.dtype], N.ptr(self.scratch),
                                         self.scratch.numel(), N.current_stream_ptr()), "b200_sort_ex")
            self._count_host = self.n_local
            return
        ws_local = self.ws.local_ptr
        mgpu_splitters(self.keys, self.samples, self.ws.ptrs_dev, r, world, g.next_epoch(), 3)
        if self.impl == "fused":
            if self.reserve:   # exact segments: count pass + one remote atomic per destination
                mgpu_count(self.keys, ws_local + self._off["sort_splitters"], self.capacity, self.ws, r, world)
            mgpu_scatter(self.keys, ws_local + self._off["sort_splitters"], self.sym_bucket.peer_ptrs, None,
                         self.capacity, self.ws, r, world, g.next_epoch(), 1 | 2 | (8 if self.reserve else 0))
            sort_devn(self.bucket, self.tmp, ws_local + self._off["sort_recv_count"], self.scratch)
            return
        # ---- NCCL baseline: partition into local send segments, exchange with all_to_all
        dist = g.dist
        self.cursors.zero_()
        seg = [self.send[d].data_ptr() for d in range(world)]
        cur = [self.cursors[d:].data_ptr() for d in range(world)]
        mgpu_scatter(self.keys, ws_local + self._off["sort_splitters"], seg, cur, self.n_local, self.ws, r,
                     world, g.next_epoch(), 1 | 4)
        recv_counts = torch.empty_like(self.cursors)
        dist.all_to_all_single(recv_counts, self.cursors, group=g.pg)
        send_n, recv_n = self.cursors.tolist(), recv_counts.tolist()
        total = int(sum(recv_n))
        need = int(g.all_reduce_max(float(total)))      # the exchange has not happened yet: just grow
        if need > self.capacity:
            self._allocate(need)
            self.retries += 1
        outs, o = [], 0
        for d in range(world):
            outs.append(self.bucket[o: o + recv_n[d]])
            o += recv_n[d]
        ins = [self.send[d, : send_n[d]] for d in range(world)]
        dist.all_to_all(outs, ins, group=g.pg)
        ST.sort_(self.bucket[:total], self.tmp[:total], self.scratch)
        self._count_host = total

    def free(self) -> None:
        self.bucket = None
        self.sym_bucket.free()

    def count(self) -> int:
        self._settle()
        if self._count_host is None:
            off = self._off["sort_recv_count"]
            self._count_host = int(self.ws.local[off: off + 8].view(torch.int64).item())
        return self._count_host

    def result(self) -> torch.Tensor:
        n = self.count()          # may re-allocate `bucket` (overflow retry): read it only afterwards
        return self.bucket[:n]
