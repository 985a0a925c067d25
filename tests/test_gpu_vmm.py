"""The VMM / multicast runtime (csrc/vmm.cu, parallel/vmm.py) on one GPU: create, export
as a file descriptor, import it back, map twice; argument checks of the self-test kernels.
The multi-process side (descriptors over unix sockets, peers, a real multicast group)
is tests/mp_worker.py."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

from cuda_mpi_openmp_b200 import _native as N
from cuda_mpi_openmp_b200.parallel import SoloPeerGroup
from cuda_mpi_openmp_b200.parallel import vmm

pytestmark = pytest.mark.gpu


def test_vmm_create_export_import_aliases_same_memory(cuda_device):
    caps = vmm.query()
    assert caps["vmm"], "B200 + driver 580 must support fd-exportable VMM allocations"
    gran = vmm.granularity(1)
    assert gran >= 4096 and gran & (gran - 1) == 0
    h = vmm._create(gran)
    a = vmm._map(h, gran)
    fd = vmm._export(h)
    h2 = vmm._import(fd)
    os.close(fd)
    b = vmm._map(h2, gran)
    assert a != b
    ta = vmm._wrap(a, gran, cuda_device).view(torch.int64)
    tb = vmm._wrap(b, gran, cuda_device).view(torch.int64)
    ta.copy_(torch.arange(gran // 8, device=cuda_device))
    torch.cuda.synchronize()
    assert torch.equal(tb, torch.arange(gran // 8, device=cuda_device))   # second mapping, same bytes
    tb[7] = -5
    torch.cuda.synchronize()
    assert int(ta[7]) == -5
    del ta, tb
    L = N.lib()
    for va, hh in ((b, h2), (a, h)):
        N.check(L.b200_vmm_unmap(C.c_void_p(va), gran), "unmap")
        N.check(L.b200_vmm_release(hh), "release")


def test_vmm_errors_are_reported_not_fatal(cuda_device):
    h = C.c_uint64(0)
    rc = N.lib().b200_vmm_create(12345, C.byref(h))      # not a multiple of the granularity
    assert rc != 0
    assert b"cuMemCreate" in N.lib().b200_last_error() or b"memCreate" in N.lib().b200_last_error()
    # the library is still usable afterwards
    assert vmm.granularity(1) > 0


def test_solo_buffer_has_no_multicast(cuda_device):
    # a multicast object needs at least two devices: the multimem paths are covered by the
    # multi-process test (tests/mp_worker.py: MULTIMEM_OK, NVLS_STATS_OK)
    g = SoloPeerGroup(cuda_device)
    buf = vmm.alloc(g, 1 << 16)
    assert buf.nbytes % vmm.granularity(1) == 0 and buf.peer_ptrs == [buf.local_ptr]
    assert buf.mc_ptr == 0
    loc = buf.view(torch.int64, 8192)
    assert int(loc.abs().sum()) == 0
    loc += 3
    torch.cuda.synchronize()
    assert int(loc.sum()) == 3 * 8192
    del loc
    buf.free()


def test_selftest_argument_checks(cuda_device):
    """The ping-pong needs two GPUs (one launch must never wait on another on the same GPU);
    on a single GPU only the argument checks run."""
    out = torch.zeros(2, dtype=torch.int64, device=cuda_device)
    rc = N.lib().b200_selftest_pingpong(N.ptr(out), N.ptr(out), 1, 0, 0, 0, N.ptr(out), N.current_stream_ptr())
    assert rc != 0 and b"iters" in N.lib().b200_last_error()
    rc = N.lib().b200_selftest_multimem(None, None, 16, 1, 0, N.current_stream_ptr())
    assert rc != 0 and b"output" in N.lib().b200_last_error()
