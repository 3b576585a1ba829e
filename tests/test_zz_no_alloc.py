"""knc_scan_cells_noalloc -- the entry point for callers that must not allocate device memory
(programs that gave the whole GPU to their own pool, latency-critical loops): same exclusive scan as
the grid build behind knc_prepare, but with its tile-state scratch in a fixed array inside the library.

Collected last on purpose (file name): with `pytest -x` every other test has already run if this fails.
The check is against the EXISTING entry point -- the d_ptrs that knc_prepare_ex computed from the very
same d_counters -- and against the CPU oracle's pointers.  Integer work, so the comparison is exact equality.
Two calls in a row on different grids, because the scratch is shared and still holds the previous call's tile
states; once on a stream of the caller's; at the size limit and just around a tile boundary; and with the
device's free memory and the memory pools watched across a call, since allocating nothing is the point.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_CELLS = 16777216  # KNC_NOALLOC_MAX_CELLS


class _DevArray:
    """Zero-copy view of raw device memory for torch (CUDA array interface)."""

    def __init__(self, ptr, count, typestr="<i4"):
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2}


def _scan(kn, counters, out, stream=None):
    rc = kn.lib().knc_scan_cells_noalloc(C.c_void_p(counters.data_ptr()), counters.numel(), C.c_void_p(out.data_ptr()),
                                         C.c_void_p(stream) if stream else None)
    assert rc == 0, kn.lib().knc_last_error()


def _scan_and_compare(kn, oracle, torch, pts, k, stream=None):
    with kn.KnProblem(pts, k=k, density=3.1) as p:  # existing entry point: knc_prepare_ex builds counters and ptrs
        info = p.info
        cells = info.dimx * info.dimy * (info.z_end - info.z_begin)
        counters = torch.as_tensor(_DevArray(info.d_counters, cells), device="cuda")
        want = torch.as_tensor(_DevArray(info.d_ptrs, cells + 1), device="cuda").clone()
        out = torch.full((cells + 1,), -1, dtype=torch.int32, device="cuda")  # caller-owned output
        torch.cuda.synchronize()
        _scan(kn, counters, out, stream)
        assert torch.equal(out, want)                                   # the sibling scan of knc_prepare
        assert np.array_equal(out.cpu().numpy(), oracle.Grid(pts).ptrs)  # the oracle's exclusive prefix sum (knearests.cu:40-48)
        assert int(out[-1]) == info.n and int(out[0]) == 0
        return cells


def test_noalloc_scan_equals_prepare_scan_twice_in_a_row(kn, oracle):
    import torch
    big = oracle.generate_uniform(1_500_000, seed=21)    # 79^3 cells = 121 tiles
    small = oracle.generate_uniform(40_000, seed=22)     # 23^3 cells =   3 tiles
    a = _scan_and_compare(kn, oracle, torch, big, 8)
    b = _scan_and_compare(kn, oracle, torch, small, 8)   # stale states of the big call must not leak in
    side = torch.cuda.Stream()                           # the caller's own stream instead of the NULL stream
    c = _scan_and_compare(kn, oracle, torch, big, 8, stream=side.cuda_stream)
    assert a == c == 79 ** 3 and b == 23 ** 3


def test_noalloc_scan_allocates_nothing(kn):
    """Free device memory (cudaMemGetInfo) and what the default and the current memory pool hold do not move across
    a call.  The first call of the process may load the kernel's module; it is made before the measurement, as the
    header asks callers to do at start-up."""
    import torch
    rt = C.CDLL("libcudart.so.12")
    counters = torch.randint(0, 5, (300_000,), dtype=torch.int32, device="cuda")
    out = torch.empty(counters.numel() + 1, dtype=torch.int32, device="cuda")
    _scan(kn, counters, out)  # set-up call

    def snapshot():
        torch.cuda.synchronize()
        free, total = C.c_size_t(), C.c_size_t()
        assert rt.cudaMemGetInfo(C.byref(free), C.byref(total)) == 0
        pools = []
        for getter in ("cudaDeviceGetDefaultMemPool", "cudaDeviceGetMemPool"):
            pool, used, reserved = C.c_void_p(), C.c_uint64(), C.c_uint64()
            assert getattr(rt, getter)(C.byref(pool), 0) == 0
            assert rt.cudaMemPoolGetAttribute(pool, 7, C.byref(used)) == 0      # cudaMemPoolAttrUsedMemCurrent
            assert rt.cudaMemPoolGetAttribute(pool, 5, C.byref(reserved)) == 0  # cudaMemPoolAttrReservedMemCurrent
            pools.append((used.value, reserved.value))
        return free.value, pools, torch.cuda.memory_allocated(), torch.cuda.memory_reserved()

    before = snapshot()
    _scan(kn, counters, out)
    assert snapshot() == before
    want = torch.cat([torch.zeros(1, dtype=torch.int64, device="cuda"), torch.cumsum(counters.to(torch.int64), 0)])
    assert torch.equal(out.to(torch.int64), want)


@pytest.mark.parametrize("cells", [MAX_CELLS, 4096 * 37 + 1, 4096, 4095, 1])
def test_noalloc_scan_at_the_limit_and_around_a_tile_boundary(kn, cells):
    import torch
    counters = torch.ones(cells, dtype=torch.int32, device="cuda")
    out = torch.full((cells + 1,), -1, dtype=torch.int32, device="cuda")
    _scan(kn, counters, out)
    assert torch.equal(out, torch.arange(cells + 1, dtype=torch.int32, device="cuda"))


def test_noalloc_scan_rejects_oversize_and_null(kn):
    L = kn.lib()
    dummy = C.c_void_p(256)  # never dereferenced: the size check comes first
    assert L.knc_scan_cells_noalloc(dummy, MAX_CELLS + 1, dummy, None) == 1  # KNC_ERR_INVALID
    assert b"KNC_NOALLOC_MAX_CELLS" in L.knc_last_error()
    assert L.knc_scan_cells_noalloc(dummy, 0, dummy, None) == 1
    assert L.knc_scan_cells_noalloc(None, 10, dummy, None) == 1
