"""cuda_knearests_b200 -- ctypes binding of the B200-native exact kNN engine.

The product is the C/C++/CUDA library ``libknearests_b200.so`` (sources in ``csrc/``, headers in
``include/``); the reference is compiled code, so its host side is C++ too (``csrc/api.cu``).  This
module is only the thin Python door that tests, ``bench.py`` and Python callers use: it mirrors the
reference interface one to one (``knearests.h``: kn_prepare / kn_solve / kn_get_points /
kn_get_knearests / kn_get_permutation / kn_print_stats / kn_free) and adds nothing to the data path.

There is NO fallback: if the shared library is missing or no GPU is usable, calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get("KN_LIB_PATH", HERE / "libknearests_b200.so"))  # override: A/B builds
EMPTY = 0xFFFFFFFF
MAX_K = 256

# every symbol include/knearests_c.h declares (tests check that the library exports all of them)
C_ABI_SYMBOLS = [
    "knc_prepare", "knc_prepare_ex", "knc_solve", "knc_solve_async", "knc_solve_finish", "knc_halo_flag_ptr",
    "knc_solve_to_host", "knc_free", "knc_trim", "knc_get_points", "knc_get_permutation",
    "knc_get_knearests", "knc_get_knearests_original", "knc_get_stats", "knc_get_info", "knc_get_timings",
    "knc_grid_dim", "knc_fast_dims", "knc_last_error", "knc_version", "knc_generate_uniform_device", "knc_scan_cells_noalloc",
    "knc_slab_route", "knc_get_knearests_global", "knc_set_index_offset", "knc_get_permutation_global",
    "knc_get_dists", "knc_query_points", "knc_get_tags",
    "knc_route_begin", "knc_route_scatter", "knc_route_end", "knc_route_begin_async", "knc_route_scatter_async",
    "knc_peer_alloc", "knc_peer_open", "knc_peer_close", "knc_peer_free", "knc_prepare_multi", "knc_solve_multi",
]
# the reference's own (C++-mangled) entry points, knearests.h:21-29
CXX_ABI_SYMBOLS = [
    "_Z10kn_prepareP6float3i", "_Z8kn_solveP10kn_problem", "_Z7kn_freePP10kn_problem",
    "_Z13kn_get_pointsP10kn_problem", "_Z16kn_get_knearestsP10kn_problem",
    "_Z18kn_get_permutationP10kn_problem", "_Z14kn_print_statsP10kn_problem",
]


def fast_density(k: int) -> float:
    """Points per cell at which the query kernels run fastest on a CUBIC grid on B200 for uniform clouds (measured at
    100 M points, profiles/r1_notes.md): the first search radius -- a sphere expected to hold k + 2.5 sqrt(k) + 2
    points, knn_query.cu -- should stay just under 2 cell widths, so that rows three cells away never enter the row
    table.  1.0 / 1.0 / 1.5 / 2.1 for k = 8 / 20 / 32 / 50.  (Round 1's bench grid; the library default is now the
    non-cubic grid of fast_dims / knc_fast_dims, and density = 3.1 gives the reference's, knearests.cu:249.)  The
    choice changes speed, never results."""
    import math
    lam = k + 2.5 * math.sqrt(k) + 2.0
    cells_in_radius_2 = 4.0 / 3.0 * math.pi * 8.0
    return max(1.0, math.ceil(1.005 * lam / cells_in_radius_2 * 10.0 - 1e-9) / 10.0)


class KnError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"knearests_b200 error {code}: {message}")
        self.code = code


class Options(C.Structure):  # knc_options
    _fields_ = [("struct_size", C.c_size_t), ("k", C.c_int), ("density", C.c_float),
                ("points_on_device", C.c_int), ("grid_dim", C.c_int), ("z_begin", C.c_int),
                ("z_end", C.c_int), ("own_z_begin", C.c_int), ("own_z_end", C.c_int),
                ("search_scale", C.c_float), ("algorithm", C.c_int), ("point_stride", C.c_int),
                ("grid_dims", C.c_int * 3), ("stream", C.c_void_p), ("query_hint", C.c_int),
                ("query_begin_hint", C.c_int), ("query_count_hint", C.c_int)]


class Info(C.Structure):  # knc_info
    _fields_ = [("n", C.c_int), ("k", C.c_int), ("dimx", C.c_int), ("dimy", C.c_int), ("dimz", C.c_int),
                ("z_begin", C.c_int), ("z_end", C.c_int), ("query_begin", C.c_int), ("query_count", C.c_int),
                ("solved", C.c_int), ("d_points", C.c_void_p), ("d_points4", C.c_void_p),
                ("d_permutation", C.c_void_p), ("d_counters", C.c_void_p), ("d_ptrs", C.c_void_p),
                ("d_knearests", C.c_void_p)]


class Stats(C.Structure):  # knc_stats
    _fields_ = [("cells", C.c_longlong), ("min_per_cell", C.c_int), ("max_per_cell", C.c_int),
                ("avg_per_cell", C.c_double), ("total", C.c_longlong), ("histogram", C.c_longlong * 64)]


class Timings(C.Structure):  # knc_timings
    _fields_ = [("build_ms", C.c_float), ("solve_ms", C.c_float), ("h2d_ms", C.c_float),
                ("restarts", C.c_uint), ("max_radius2", C.c_float), ("handed_over_blocks", C.c_uint),
                ("second_retries", C.c_uint)]


def build(verbose: bool = False) -> Path:
    """Compile libknearests_b200.so for sm_100a in-tree (nvcc cross-compiles without a GPU)."""
    subprocess.run(["make", "-C", str(HERE / "csrc")], check=True,
                   stdout=None if verbose else subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def lib() -> C.CDLL:
    """Load the CUDA library.  Raises if it has not been built: there is no other implementation."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise ImportError(f"{LIB_PATH} is missing -- build it with `python -c 'import __graft_entry__ as g; "
                              f"g.build()'` or `make -C cuda_knearests_b200/csrc`; there is no CPU fallback")
        L = C.CDLL(str(LIB_PATH))
        vp = C.c_void_p
        L.knc_prepare.argtypes = [vp, C.c_int, C.c_int, C.POINTER(vp)]
        L.knc_prepare_ex.argtypes = [vp, C.c_int, C.POINTER(Options), C.POINTER(vp)]
        for nm in ("knc_solve", "knc_solve_async", "knc_solve_finish", "knc_free"):
            getattr(L, nm).argtypes = [vp]
        L.knc_halo_flag_ptr.argtypes = [vp, C.POINTER(vp)]
        L.knc_route_begin_async.argtypes = [vp, C.c_int, C.c_uint, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int, vp, vp,
                                            C.POINTER(vp)]
        L.knc_route_scatter_async.argtypes = [vp, C.POINTER(vp), C.POINTER(C.c_longlong), vp]
        for nm in ("knc_get_points", "knc_get_permutation", "knc_get_knearests", "knc_get_knearests_original"):
            getattr(L, nm).argtypes = [vp, vp]
        L.knc_solve_to_host.argtypes = [vp, vp, C.c_int]
        L.knc_get_dists.argtypes = [vp, vp]
        L.knc_query_points.argtypes = [vp, vp, C.c_int, vp, vp]
        L.knc_get_stats.argtypes = [vp, C.POINTER(Stats)]
        L.knc_get_info.argtypes = [vp, C.POINTER(Info)]
        L.knc_get_timings.argtypes = [vp, C.POINTER(Timings)]
        L.knc_grid_dim.argtypes = [C.c_int, C.c_float]
        L.knc_fast_dims.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_int)]
        L.knc_generate_uniform_device.argtypes = [vp, C.c_longlong, C.c_longlong, C.c_uint64]
        L.knc_scan_cells_noalloc.argtypes = [vp, C.c_longlong, vp, vp]
        L.knc_slab_route.argtypes = [vp, C.c_int, C.c_uint, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int, vp,
                                     C.c_longlong, C.POINTER(C.c_longlong)]
        L.knc_get_knearests_global.argtypes = [vp, vp, C.c_int, vp, vp]
        L.knc_route_begin.argtypes = [vp, C.c_int, C.c_uint, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int,
                                      C.POINTER(vp), C.POINTER(C.c_longlong)]
        L.knc_route_scatter.argtypes = [vp, C.POINTER(vp), C.POINTER(C.c_longlong)]
        L.knc_route_end.argtypes = [vp]
        L.knc_peer_alloc.argtypes = [C.c_size_t, C.POINTER(vp), C.c_char_p]
        L.knc_peer_open.argtypes = [C.c_char_p, C.POINTER(vp)]
        L.knc_peer_close.argtypes = [vp]
        L.knc_peer_free.argtypes = [vp]
        L.knc_set_index_offset.argtypes = [vp, C.c_longlong]
        L.knc_get_permutation_global.argtypes = [vp, vp, C.c_int, vp]
        L.knc_get_tags.argtypes = [vp, C.POINTER(vp)]
        L.knc_prepare_multi.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]
        L.knc_solve_multi.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, vp]
        L.knc_last_error.restype = C.c_char_p
        L.knc_version.restype = C.c_char_p
        _lib = L
    return _lib


def _check(rc: int) -> None:
    if rc != 0:
        raise KnError(rc, lib().knc_last_error().decode())


def grid_dim(n: int, density: float = 3.1) -> int:
    return int(lib().knc_grid_dim(int(n), float(density)))


def fast_dims(n: int, k: int = 0) -> tuple:
    """(dimx, dimy, dimz) of the grid the library chooses by default (knc_fast_dims): y/z cells a little wider than
    the first search radius, narrow x cells, about one point per cell."""
    d = (C.c_int * 3)()
    _check(lib().knc_fast_dims(int(n), int(k), d))
    return int(d[0]), int(d[1]), int(d[2])


def version() -> str:
    return lib().knc_version().decode()


class KnProblem:
    """One point cloud on one GPU -- the reference's ``kn_problem`` handle (knearests.h:3-16).

    ``points`` is either a host array (anything numpy can view as float32 [n,3]; copied to the device,
    as kn_prepare does, knearests.cu:308-311) or a CUDA ``torch.Tensor`` float32 [n,3] (used in place);
    a CUDA tensor [n,4] is read as 16-byte records x, y, z, tag (knc_options.point_stride = 4; see tags_ptr).
    """

    def __init__(self, points, k: int = 0, density: float = 0.0, grid_dim: int = 0, slab=None,
                 search_scale: float = 0.0, algorithm: int = 0, grid_dims=None, stream: int = 0, query_hint=None):
        """stream: a cudaStream_t (integer) of the caller -- all work of the handle then runs there and the constructor
        does not wait for the build (device input only); query_hint = (query_begin, query_count) of a slab handle."""
        self._h = C.c_void_p()
        opt = Options()
        opt.struct_size = C.sizeof(Options)
        opt.k = int(k)
        opt.density = float(density)
        opt.grid_dim = int(grid_dim)
        if grid_dims is not None:
            opt.grid_dims[0], opt.grid_dims[1], opt.grid_dims[2] = (int(v) for v in grid_dims)
        opt.stream = int(stream) if stream else None
        if query_hint is not None:
            opt.query_hint, opt.query_begin_hint, opt.query_count_hint = 1, int(query_hint[0]), int(query_hint[1])
        opt.search_scale = float(search_scale)
        opt.algorithm = int(algorithm)
        if slab is not None:
            opt.z_begin, opt.z_end, opt.own_z_begin, opt.own_z_end = (int(v) for v in slab)
        if hasattr(points, "is_cuda"):
            import torch
            if not points.is_cuda or points.dtype != torch.float32 or not points.is_contiguous():
                raise ValueError("device input must be a contiguous CUDA float32 tensor [n,3] or [n,4]")
            if not stream:
                torch.cuda.current_stream(points.device).synchronize()  # the library works on its own stream
            stride = 4 if (points.dim() == 2 and points.shape[1] == 4) else 3
            opt.point_stride = stride
            n = points.numel() // stride
            opt.points_on_device = 1
            self._keep = points
            ptr = C.c_void_p(points.data_ptr())
        else:
            arr = np.ascontiguousarray(points, np.float32).reshape(-1, 3)
            n = arr.shape[0]
            self._keep = arr
            ptr = C.c_void_p(arr.ctypes.data)
        _check(lib().knc_prepare_ex(ptr, n, C.byref(opt), C.byref(self._h)))
        self._keep = points if opt.points_on_device else None  # host input is not retained (kn_prepare)
        self.info = self._info()

    # -- reference interface -------------------------------------------------------------------------
    def solve(self) -> "KnProblem":  # kn_solve, knearests.cu:348
        _check(lib().knc_solve(self._h))
        self.info = self._info()
        return self

    def solve_async(self) -> "KnProblem":  # first half of kn_solve: enqueue only
        _check(lib().knc_solve_async(self._h))
        return self

    def solve_finish(self) -> "KnProblem":  # second half: wait, read the diagnostics, raise on a halo miss
        _check(lib().knc_solve_finish(self._h))
        self.info = self._info()
        return self

    def halo_flag_ptr(self) -> int:
        """Device address of the count of this slab's queries that reached beyond their halo planes."""
        ptr = C.c_void_p()
        _check(lib().knc_halo_flag_ptr(self._h, C.byref(ptr)))
        return int(ptr.value)

    def solve_to_host(self, out: np.ndarray | None = None, chunks: int = 8) -> np.ndarray:
        """kn_solve + kn_get_knearests overlapped (knc_solve_to_host); `out` should be page-locked."""
        if out is None:
            out = np.empty((self.info.query_count, self.info.k), np.uint32)
        _check(lib().knc_solve_to_host(self._h, out.ctypes.data, int(chunks)))
        self.info = self._info()
        return out

    def get_dists(self) -> np.ndarray:
        """Squared distances of the rows (the reference's GPU API returns none; its kd-tree does)."""
        out = np.empty((self.info.query_count, self.info.k), np.float32)
        _check(lib().knc_get_dists(self._h, out.ctypes.data))
        return out

    def tags_ptr(self) -> int:
        """Device address of the input records' 4th words in sorted order (knc_get_tags; [n, 4] input only)."""
        ptr = C.c_void_p()
        _check(lib().knc_get_tags(self._h, C.byref(ptr)))
        return int(ptr.value)

    def query_points(self, xyz, with_d2: bool = False):
        """kNN of points that are not in the cloud (kd_tree.cpp:168-192 semantics)."""
        q = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        rows = np.empty((q.shape[0], self.info.k), np.uint32)
        d2 = np.empty((q.shape[0], self.info.k), np.float32) if with_d2 else None
        _check(lib().knc_query_points(self._h, q.ctypes.data, q.shape[0], rows.ctypes.data,
                                      d2.ctypes.data if with_d2 else None))
        return (rows, d2) if with_d2 else rows

    def get_points(self) -> np.ndarray:  # kn_get_points, knearests.cu:410
        out = np.empty((self.info.n, 3), np.float32)
        _check(lib().knc_get_points(self._h, out.ctypes.data))
        return out

    def get_permutation(self) -> np.ndarray:  # kn_get_permutation, knearests.cu:420
        out = np.empty(self.info.n, np.uint32)
        _check(lib().knc_get_permutation(self._h, out.ctypes.data))
        return out

    def get_knearests(self, out: np.ndarray | None = None) -> np.ndarray:  # kn_get_knearests, knearests.cu:430
        if out is None:
            out = np.empty((self.info.query_count, self.info.k), np.uint32)
        _check(lib().knc_get_knearests(self._h, out.ctypes.data))
        return out

    def get_knearests_original(self) -> np.ndarray:  # the remap of test_knearests.cu:155-160
        out = np.empty((self.info.n, self.info.k), np.uint32)
        _check(lib().knc_get_knearests_original(self._h, out.ctypes.data))
        return out

    def stats(self) -> dict:  # kn_print_stats, knearests.cu:440
        s = Stats()
        _check(lib().knc_get_stats(self._h, C.byref(s)))
        return {"cells": s.cells, "min": s.min_per_cell, "max": s.max_per_cell, "avg": s.avg_per_cell,
                "total": s.total, "histogram": list(s.histogram)}

    def free(self) -> None:  # kn_free, knearests.cu:396
        if self._h:
            lib().knc_free(self._h)
            self._h = C.c_void_p()

    # -- extras ----------------------------------------------------------------------------------------
    def timings(self) -> dict:
        t = Timings()
        _check(lib().knc_get_timings(self._h, C.byref(t)))
        return {"build_ms": t.build_ms, "solve_ms": t.solve_ms, "h2d_ms": t.h2d_ms, "restarts": t.restarts,
                "max_radius2": t.max_radius2, "handed_over_blocks": t.handed_over_blocks,
                "second_retries": t.second_retries}

    def _info(self) -> Info:
        i = Info()
        _check(lib().knc_get_info(self._h, C.byref(i)))
        return i

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.free()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def prepare_multi(points, k: int = 0, ngpus: int = 0) -> KnProblem:
    """kn_prepare over `ngpus` ranks of this process (knc_prepare_multi; <= 0: every device, more ranks than devices
    share devices).  The handle answers solve / get_points / get_permutation / get_knearests / stats / free."""
    arr = np.ascontiguousarray(points, np.float32).reshape(-1, 3)
    self = KnProblem.__new__(KnProblem)
    self._h = C.c_void_p()
    self._keep = None
    _check(lib().knc_prepare_multi(C.c_void_p(arr.ctypes.data), arr.shape[0], int(k), int(ngpus), C.byref(self._h)))
    self.info = self._info()
    return self


# function-style mirror of knearests.h, for code that reads like the reference's test driver
def kn_prepare(points, numpoints: int | None = None, k: int = 0) -> KnProblem:
    pts = np.asarray(points, np.float32).reshape(-1, 3) if not hasattr(points, "is_cuda") else points
    if numpoints is not None and not hasattr(points, "is_cuda"):
        pts = pts[:numpoints]
    return KnProblem(pts, k=k)


def kn_solve(kn: KnProblem) -> None:
    kn.solve()


def kn_get_points(kn: KnProblem) -> np.ndarray:
    return kn.get_points()


def kn_get_knearests(kn: KnProblem) -> np.ndarray:
    return kn.get_knearests()


def kn_get_permutation(kn: KnProblem) -> np.ndarray:
    return kn.get_permutation()


def kn_print_stats(kn: KnProblem) -> None:
    s = kn.stats()
    print(f"Grid:  points per cell: {s['min']} (min), {s['max']} (max), {s['avg']:g} avg, total {s['total']}",
          file=sys.stderr)
    for c, v in enumerate(s["histogram"]):
        if v:
            print(f"[{c}] => {v}", file=sys.stderr)


def kn_free(kn: KnProblem) -> None:
    kn.free()


def generate_uniform_device(n: int, seed: int = 20171004, first: int = 0, device=None):
    """Synthetic uniform cloud of SURVEY 8d, generated on the GPU into a torch tensor [n,3]."""
    import torch
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    with torch.cuda.device(dev):
        t = torch.empty((n, 3), dtype=torch.float32, device=dev)
        torch.cuda.current_stream().synchronize()
        _check(lib().knc_generate_uniform_device(C.c_void_p(t.data_ptr()), first, n, seed))
    return t
