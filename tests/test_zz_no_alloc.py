"""GPU: gs2d_cumsum_tiles_hit_static -- the scan with its scratch in a file-scope device array of the library
(no allocation anywhere in the call, no counterpart in the
reference) -- against the existing entry point gs2d_cumsum_tiles_hit with caller scratch, through the C ABI,
on two calls in a row (the array keeps the previous call's contents), and the size limit.
Collected last on purpose."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
cuda = torch.device("cuda:0")


def _existing(lib, nth):
    from gsvc_b200 import _lib
    n = nth.numel()
    cum = torch.empty_like(nth)
    total = torch.zeros(1, dtype=torch.int32, device=cuda)
    nbytes = lib.gs2d_scan_scratch_bytes(n)
    scratch = torch.empty((nbytes,), dtype=torch.uint8, device=cuda)
    _lib.check(lib.gs2d_cumsum_tiles_hit(n, nth.data_ptr(), cum.data_ptr(), total.data_ptr(), None, scratch.data_ptr(),
                                         nbytes, torch.cuda.current_stream().cuda_stream), "cumsum_tiles_hit")
    return cum, total


def test_static_scratch_scan_matches_the_existing_entry_point_twice_in_a_row():
    from gsvc_b200 import _lib
    lib = _lib.load()
    _lib.require_device(0)
    torch.cuda.set_device(0)
    g = torch.Generator().manual_seed(7)
    # two different inputs of different sizes back to back: the second call meets the first one's flags
    for n in (1_000_003, 50_000, 257, 1):
        nth = torch.randint(0, 12, (n,), generator=g, dtype=torch.int32).to(cuda)
        want, want_total = _existing(lib, nth)
        for _ in range(2):
            cum = torch.full((n,), -1, dtype=torch.int32, device=cuda)      # outputs allocated BEFORE the call
            total = torch.full((1,), -1, dtype=torch.int32, device=cuda)
            torch.cuda.synchronize()
            before = torch.cuda.memory_allocated()
            rc = lib.gs2d_cumsum_tiles_hit_static(n, nth.data_ptr(), cum.data_ptr(), total.data_ptr(), None,
                                                  torch.cuda.current_stream().cuda_stream)
            assert rc == 0, lib.gs2d_last_error()
            assert torch.cuda.memory_allocated() == before                  # nothing came from torch's allocator
            torch.cuda.synchronize()
            assert torch.equal(cum, want)                                   # integers: bit-exact
            assert torch.equal(total, want_total)
            assert torch.equal(cum, torch.cumsum(nth, 0, dtype=torch.int32))


def test_static_scratch_scan_rejects_sizes_beyond_its_array():
    from gsvc_b200 import _lib
    lib = _lib.load()
    limit = 1 << 22                                                          # GS2D_CUMSUM_STATIC_MAX_N (include/gs2d.h)
    n = limit + 1
    nth = torch.ones(n, dtype=torch.int32, device=cuda)
    cum = torch.empty_like(nth)
    rc = lib.gs2d_cumsum_tiles_hit_static(n, nth.data_ptr(), cum.data_ptr(), None, None,
                                          torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.gs2d_last_error().decode()
    assert "exceeds" in msg and str(limit) in msg
    # exactly at the limit it works
    rc = lib.gs2d_cumsum_tiles_hit_static(limit, nth.data_ptr(), cum.data_ptr(), None, None,
                                          torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.gs2d_last_error()
    torch.cuda.synchronize()
    assert int(cum[limit - 1]) == limit and int(cum[0]) == 1


def test_static_scratch_scan_empty_input():
    from gsvc_b200 import _lib
    lib = _lib.load()
    total = torch.full((1,), 5, dtype=torch.int32, device=cuda)
    assert lib.gs2d_cumsum_tiles_hit_static(0, None, None, total.data_ptr(), None,
                                            torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()
    assert int(total) == 0
