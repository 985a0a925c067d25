"""b200_sort_noalloc — the sort entry point for callers that are not allowed to
allocate GPU memory (an embedded plug-in whose host application's pool owns the
device memory, a latency-critical loop that cannot take cudaMalloc/cudaFree's
device-wide synchronisation, code under a no-allocation-after-start-up rule).

It is called here exactly as such a caller would: through the C ABI of
libb200labs.so with raw device addresses, passing only the input and the output
buffer.  Its result must equal b200_sort's (the entry point that takes caller-
provided scratch) bit for bit — the same tolerance the other sort tests use — on
two calls in a row: the second call finds the library's built-in scratch array
full of the first call's counters and look-back words and must not be disturbed
by them.  The file sorts last on purpose: if it fails, a `-x` run has already
executed every other test."""
import ctypes as C

import numpy as np
import pytest
import torch

from cuda_mpi_openmp_b200 import _native as N

pytestmark = pytest.mark.gpu

MAX_KEYS = 1 << 20            # B200_SORT_NOALLOC_MAX_KEYS
KEY_I32, KEY_F32, KEY_U8, KEY_U32 = 0, 1, 2, 3


def _bind(lib):
    fn = lib.b200_sort_noalloc
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
    return fn


def _sort_with_caller_scratch(lib, keys: torch.Tensor, key_type: int) -> torch.Tensor:
    """The existing entry point: scratch and second buffer allocated by the caller."""
    out = keys.clone()
    tmp = torch.empty_like(out)
    need = lib.b200_sort_scratch_bytes(out.numel(), key_type)
    scratch = torch.empty(need, dtype=torch.uint8, device=out.device)
    N.check(lib.b200_sort(N.ptr(out), N.ptr(tmp), out.numel(), key_type, N.ptr(scratch), need,
                          N.current_stream_ptr()), "b200_sort")
    return out


def _make(rng, n, key_type, device):
    if key_type == KEY_U8:
        k = rng.integers(0, 256, n, dtype=np.int64).astype(np.uint8)
    elif key_type == KEY_F32:
        k = (rng.standard_normal(n) * 1e4).astype(np.float32)
    else:
        k = rng.integers(-2 ** 31, 2 ** 31 - 1, n, dtype=np.int64).astype(np.int32)
    return torch.from_numpy(k).to(device)


@pytest.mark.parametrize("key_type", [KEY_I32, KEY_F32, KEY_U8, KEY_U32])
def test_two_calls_in_a_row_match_the_scratch_taking_entry_point(cuda_device, lib, key_type):
    sort_noalloc = _bind(lib)
    rng = np.random.default_rng(2024 + key_type)
    stream = N.current_stream_ptr()
    # first call: the largest size allowed; second call: smaller, different data, so
    # every counter and look-back word left behind by the first one is stale
    for n in (MAX_KEYS, 300_007):
        keys = _make(rng, n, key_type, cuda_device)
        want = _sort_with_caller_scratch(lib, keys, key_type)
        if key_type == KEY_U32:                  # same bits, unsigned order: check the oracle too
            assert np.array_equal(want.cpu().numpy().view(np.uint32),
                                  np.sort(keys.cpu().numpy().view(np.uint32)))
        out = torch.empty_like(keys)
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info(cuda_device)
        rc = sort_noalloc(N.ptr(keys), N.ptr(out), n, key_type, stream)
        assert rc == 0, lib.b200_last_error().decode()
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info(cuda_device)
        assert torch.equal(out.view(torch.uint8), want.view(torch.uint8)), (key_type, n)
        if n != MAX_KEYS:                        # (the very first call may load the module lazily)
            assert free_after == free_before     # the call took no device memory
    # in place (input and output alias) is allowed too
    keys = _make(rng, 70_001, key_type, cuda_device)
    want = _sort_with_caller_scratch(lib, keys, key_type)
    assert sort_noalloc(N.ptr(keys), N.ptr(keys), keys.numel(), key_type, stream) == 0
    torch.cuda.synchronize()
    assert torch.equal(keys.view(torch.uint8), want.view(torch.uint8))


def test_buffers_that_are_only_four_byte_aligned(cuda_device, lib):
    """A caller's buffers need not be 16-byte aligned (that only switches the vectorised
    histogram loads off): input and output both start 4 bytes into an allocation."""
    sort_noalloc = _bind(lib)
    rng = np.random.default_rng(77)
    n = 123_457
    backing_in = _make(rng, n + 1, KEY_I32, cuda_device)
    backing_out = torch.empty_like(backing_in)
    keys, out = backing_in[1:], backing_out[1:]
    assert keys.data_ptr() % 16 == 4 and out.data_ptr() % 16 == 4
    want = _sort_with_caller_scratch(lib, keys, KEY_I32)
    for _ in range(2):
        out.zero_()
        assert sort_noalloc(N.ptr(keys), N.ptr(out), n, KEY_I32, N.current_stream_ptr()) == 0
        torch.cuda.synchronize()
        assert torch.equal(out, want)


def test_sizes_beyond_the_built_in_scratch_are_refused(cuda_device, lib):
    sort_noalloc = _bind(lib)
    keys = torch.zeros(MAX_KEYS + 1, dtype=torch.int32, device=cuda_device)
    out = torch.full_like(keys, 5)
    rc = sort_noalloc(N.ptr(keys), N.ptr(out), keys.numel(), KEY_I32, N.current_stream_ptr())
    assert rc != 0
    msg = lib.b200_last_error().decode()
    assert "exceeds" in msg and str(MAX_KEYS) in msg
    torch.cuda.synchronize()
    assert bool((out == 5).all())                # nothing was launched
    assert sort_noalloc(N.ptr(keys), N.ptr(out), 0, KEY_I32, N.current_stream_ptr()) == 0
