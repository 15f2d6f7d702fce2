"""gi2d_scan_tiles_static: the allocation-free variant of gi2d_scan_tiles (scratch = a file-scope device array of the
library, see include/gi2d.h), called through the C ABI and compared with the existing entry point on the same inputs.
Integer work: the bar is bit-exact.  The file sorts last on purpose."""
import numpy as np
import pytest
import torch

from gaussianimage_b200._lib import check, lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
MAX_N = 1 << 22          # GI2D_SCAN_STATIC_MAX_N
ERR_STATIC_LIMIT = 3     # GI2D_ERR_STATIC_LIMIT


def _existing(nth):
    L = lib()
    N = nth.numel()
    cum = torch.empty(N, dtype=torch.int32, device=DEV)
    total = torch.zeros(1, dtype=torch.int32, device=DEV)
    nbytes = L.gi2d_scan_workspace_bytes(N)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=DEV)
    check(L.gi2d_scan_tiles(N, nth.data_ptr(), cum.data_ptr(), total.data_ptr(), ws.data_ptr(), nbytes,
                            torch.cuda.current_stream().cuda_stream), "scan_tiles")
    torch.cuda.synchronize()
    return cum, total


def _static(nth, cum, total):
    """outputs are the caller's (allocated before the call); the entry point itself allocates nothing"""
    check(lib().gi2d_scan_tiles_static(nth.numel(), nth.data_ptr(), cum.data_ptr(), total.data_ptr(),
                                       torch.cuda.current_stream().cuda_stream), "scan_tiles_static")


@pytest.mark.parametrize("N", [1, 1000, 70000, 1_000_000, MAX_N])
def test_static_scan_equals_existing_entry_point_twice_in_a_row(N):
    g = torch.Generator().manual_seed(N)
    a = torch.randint(0, 40, (N,), generator=g, dtype=torch.int32).to(DEV)
    b = torch.randint(0, 7, (N,), generator=g, dtype=torch.int32).to(DEV)      # second call: other data, same scratch
    cum = [torch.full((N,), -1, dtype=torch.int32, device=DEV) for _ in range(2)]
    tot = [torch.full((1,), -1, dtype=torch.int32, device=DEV) for _ in range(2)]
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats(DEV)["allocation.all.allocated"]
    _static(a, cum[0], tot[0])           # two calls in a row on one stream: the second finds the first one's flags in the scratch
    _static(b, cum[1], tot[1])
    after = torch.cuda.memory_stats(DEV)["allocation.all.allocated"]
    torch.cuda.synchronize()
    assert after == before               # (torch's allocator was not involved either)
    for x, c, t in ((a, cum[0], tot[0]), (b, cum[1], tot[1])):
        want_cum, want_total = _existing(x)
        assert torch.equal(c, want_cum)
        assert torch.equal(t, want_total)
        ref = np.cumsum(x.cpu().numpy().astype(np.int64))
        assert int(t.item()) == int(ref[-1]) and np.array_equal(c.cpu().numpy().astype(np.int64), ref)


def test_static_scan_rejects_sizes_beyond_its_scratch():
    L = lib()
    nth = torch.zeros(8, dtype=torch.int32, device=DEV)
    cum = torch.full((8,), -1, dtype=torch.int32, device=DEV)
    st = torch.cuda.current_stream().cuda_stream
    rc = L.gi2d_scan_tiles_static(MAX_N + 1, nth.data_ptr(), cum.data_ptr(), 0, st)     # rejected before anything is touched
    assert rc == ERR_STATIC_LIMIT
    assert b"static scratch" in L.gi2d_status_string(rc)
    assert L.gi2d_scan_tiles_static(0, nth.data_ptr(), cum.data_ptr(), 0, st) == 1      # GI2D_ERR_INVALID
    torch.cuda.synchronize()
    assert bool((cum == -1).all())
