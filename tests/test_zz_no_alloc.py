"""knc_scan_cells_noalloc -- the entry point for callers that must not allocate device memory
(programs that gave the whole GPU to their own pool, latency-critical loops): same exclusive scan as
the grid build behind knc_prepare, but with its tile-state scratch in a fixed array inside the library.

Collected last on purpose (file name): with `pytest -x` every other test has already run if this fails.
The check is against the EXISTING entry point: the d_ptrs that knc_prepare_ex computed from the very
same d_counters.  Integer work, so the comparison is exact equality.  Two calls in a row on different
grids, because the scratch is shared and still holds the previous call's tile states.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


class _DevArray:
    """Zero-copy view of raw device memory for torch (CUDA array interface)."""

    def __init__(self, ptr, count, typestr="<i4"):
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2}


def _scan_and_compare(kn, torch, pts, k):
    with kn.KnProblem(pts, k=k) as p:  # existing entry point: knc_prepare_ex builds counters and ptrs
        info = p.info
        cells = info.dimx * info.dimy * (info.z_end - info.z_begin)
        counters = torch.as_tensor(_DevArray(info.d_counters, cells), device="cuda")
        want = torch.as_tensor(_DevArray(info.d_ptrs, cells + 1), device="cuda").clone()
        out = torch.full((cells + 1,), -1, dtype=torch.int32, device="cuda")  # caller-owned output
        torch.cuda.synchronize()
        rc = kn.lib().knc_scan_cells_noalloc(C.c_void_p(counters.data_ptr()), cells, C.c_void_p(out.data_ptr()), None)
        assert rc == 0, kn.lib().knc_last_error()
        assert torch.equal(out, want)
        assert int(out[-1]) == info.n and int(out[0]) == 0
        return cells


def test_noalloc_scan_equals_prepare_scan_twice_in_a_row(kn, oracle):
    import torch
    big = oracle.generate_uniform(1_500_000, seed=21)    # 79^3 cells = 121 tiles
    small = oracle.generate_uniform(40_000, seed=22)     # 23^3 cells =   3 tiles
    a = _scan_and_compare(kn, torch, big, 8)
    b = _scan_and_compare(kn, torch, small, 8)           # stale states of the big call must not leak in
    c = _scan_and_compare(kn, torch, big, 8)
    assert a == c == 79 ** 3 and b == 23 ** 3


def test_noalloc_scan_rejects_oversize_and_null(kn):
    L = kn.lib()
    dummy = C.c_void_p(256)  # never dereferenced: the size check comes first
    assert L.knc_scan_cells_noalloc(dummy, 16777216 + 1, dummy, None) == 1  # KNC_ERR_INVALID
    assert b"KNC_NOALLOC_MAX_CELLS" in L.knc_last_error()
    assert L.knc_scan_cells_noalloc(dummy, 0, dummy, None) == 1
    assert L.knc_scan_cells_noalloc(None, 10, dummy, None) == 1
