"""oracle -- CPU checkers for the cuda_knearests hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package, and only as the checker or as the timed CPU
arm.  The product (``cuda_knearests_b200``) never imports it and has no CPU fallback.

Two layers, both reached through ctypes:

* ``libknn_oracle.so``  -- our plain-C restatement of the reference algorithm
  (``oracle/knn_oracle.c``; every function cites the reference file:line it follows).
* ``_ref/*.so``         -- the UNMODIFIED reference, compiled from ``/root/reference`` by
  ``oracle/Makefile`` (kd_tree.cpp for the CPU arm, knearests.cu for sm_100a, the .xyz loader).
  These pin the restatement; they are git-ignored build outputs that travel to the GPU box.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
REF_DIR = HERE / "_ref"
EMPTY = np.uint32(0xFFFFFFFF)
REFERENCE_DENSITY = np.float32(3.1)  # knearests.cu:249

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")


def build(ref: bool = True) -> None:
    """Compile the checkers (``make -C oracle``).  Building the checker is not using it."""
    subprocess.run(["make", "-C", str(HERE), "all" if ref else "libknn_oracle.so"], check=True,
                   stdout=subprocess.DEVNULL)


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        so = HERE / "libknn_oracle.so"
        if not so.exists():
            build(ref=False)
        L = C.CDLL(str(so))
        L.kno_grid_dim.argtypes = [C.c_int, C.c_float]
        L.kno_grid_dim.restype = C.c_int
        L.kno_cell_from_point.argtypes = [C.c_int] * 3 + [C.c_float] * 3
        L.kno_cell_from_point.restype = C.c_int
        L.kno_build.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, C.c_int, _i32p, _i32p, _f32p, _u32p]
        L.kno_build.restype = None
        L.kno_ring_table.argtypes = [C.c_int, C.c_int, C.c_int, _i32p, _f32p]
        L.kno_ring_table.restype = C.c_int
        L.kno_knearest_refwalk.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, _i32p, _i32p, _f32p,
                                           C.c_int, _i32p, _f32p, C.c_int, _u32p]
        L.kno_knearest_refwalk.restype = C.c_int
        L.kno_knn_exact.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, _i32p, _f32p, C.c_int, _u32p,
                                    C.c_void_p]
        L.kno_knn_exact.restype = None
        L.kno_fast_dims.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_int)]
        L.kno_fast_dims.restype = None
        L.kno_brute.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _u32p, C.c_void_p]
        L.kno_brute.restype = None
        L.kno_brute_external.argtypes = [_f32p, C.c_int, _f32p, C.c_int, C.c_int, _u32p, C.c_void_p]
        L.kno_brute_external.restype = None
        L.kno_canonical_rows.argtypes = [_u32p, C.c_void_p, _f32p, C.c_int, C.c_int, _u32p, _f32p]
        L.kno_canonical_rows.restype = None
        L.kno_load_xyz.argtypes = [C.c_char_p, _f32p, C.c_int, C.c_int]
        L.kno_load_xyz.restype = C.c_int
        L.kno_normalize.argtypes = [_f32p, C.c_int]
        L.kno_normalize.restype = None
        L.kno_generate_uniform.argtypes = [C.c_uint64, C.c_long, C.c_long, _f32p]
        L.kno_generate_uniform.restype = None
        L.kno_distance_gpu.argtypes = [C.c_float] * 6
        L.kno_distance_gpu.restype = C.c_float
        L.kno_distance_kdtree.argtypes = [C.c_float] * 6
        L.kno_distance_kdtree.restype = C.c_float
        _lib = L
    return _lib


# ------------------------------------------------------------------------------ restatement
def grid_dim(n: int, density: float = 3.1) -> int:
    return int(lib().kno_grid_dim(int(n), float(np.float32(density))))


def generate_uniform(n: int, seed: int = 20171004, first: int = 0) -> np.ndarray:
    xyz = np.empty((n, 3), np.float32)
    lib().kno_generate_uniform(seed, first, n, xyz)
    return xyz


def load_xyz(path: str, normalize: bool = True) -> np.ndarray:
    with open(path) as f:
        cap = int(f.readline().split()[0]) + 16
    xyz = np.empty((cap, 3), np.float32)
    n = lib().kno_load_xyz(str(path).encode(), xyz, cap, int(normalize))
    if n < 0 or n > cap:
        raise IOError(f"cannot load {path} (n={n})")
    return xyz[:n].copy()


def fast_dims(n: int, k: int) -> tuple:
    """The engine's default grid resolution per axis, restated (kno_fast_dims)."""
    d = (C.c_int * 3)()
    lib().kno_fast_dims(int(n), int(k), d)
    return int(d[0]), int(d[1]), int(d[2])


class Grid:
    """Deterministic grid build (knearests.cu:32-60 made canonical, see knn_oracle.c).  The grid is the reference's
    cube for `density` points per cell (knearests.cu:249) unless `dim` (cube) or `dims` (x, y, z) name it."""

    def __init__(self, xyz: np.ndarray, density: float = 3.1, dim: int | None = None, dims=None):
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        self.n = n = xyz.shape[0]
        if dims is not None:
            self.dims = tuple(int(v) for v in dims)
        else:
            D = int(dim) if dim else grid_dim(n, density)
            self.dims = (D, D, D)
        self.dim = self.dims[0] if self.dims[0] == self.dims[1] == self.dims[2] else None
        dx, dy, dz = self.dims
        cells = dx * dy * dz
        self.counters = np.zeros(cells, np.int32)
        self.ptrs = np.zeros(cells + 1, np.int32)
        self.sorted_xyz = np.empty((n, 3), np.float32)
        self.permutation = np.empty(n, np.uint32)
        lib().kno_build(xyz, n, dx, dy, dz, self.counters, self.ptrs, self.sorted_xyz, self.permutation)

    def knn_exact(self, K: int, with_d2: bool = False):
        """Engine contract: rows by sorted position, ascending (d2_gpu, sorted position)."""
        out = np.empty((self.n, K), np.uint32)
        d2 = np.empty((self.n, K), np.float32) if with_d2 else None
        dx, dy, dz = self.dims
        lib().kno_knn_exact(dx, dy, dz, self.n, self.ptrs, self.sorted_xyz, K, out,
                            d2.ctypes.data if with_d2 else None)
        return (out, d2) if with_d2 else out

    def knn_refwalk(self, K: int):
        """Faithful restatement of the reference's knearest kernel (knearests.cu:93-148)."""
        D = self.dim
        if D is None:
            raise ValueError("the reference's walk is defined on its cubic grid only")
        if D < 16:  # knearests.cu:255-258
            raise ValueError("reference does not support low number of input points")
        offs = np.empty(16 ** 4, np.int32)
        dists = np.empty(16 ** 4, np.float32)
        m = lib().kno_ring_table(D, D, D, offs, dists)
        out = np.empty((self.n, K), np.uint32)
        exhausted = lib().kno_knearest_refwalk(D, D, D, self.n, self.ptrs, self.counters, self.sorted_xyz,
                                               m, offs, dists, K, out)
        return out, exhausted


def brute(xyz: np.ndarray, K: int, q0: int = 0, q1: int | None = None, formula: str = "gpu",
          with_d2: bool = False):
    xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
    n = xyz.shape[0]
    q1 = n if q1 is None else q1
    out = np.empty((q1 - q0, K), np.uint32)
    d2 = np.empty((q1 - q0, K), np.float32) if with_d2 else None
    lib().kno_brute(xyz, n, q0, q1, K, 0 if formula == "gpu" else 1, out,
                    d2.ctypes.data if with_d2 else None)
    return (out, d2) if with_d2 else out


def brute_external(xyz: np.ndarray, queries: np.ndarray, K: int, with_d2: bool = False):
    """K nearest of `xyz` for arbitrary query points (kd_tree.cpp:168-192 semantics), order (d2_gpu, index)."""
    xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
    queries = np.ascontiguousarray(queries, np.float32).reshape(-1, 3)
    out = np.empty((queries.shape[0], K), np.uint32)
    d2 = np.empty((queries.shape[0], K), np.float32) if with_d2 else None
    lib().kno_brute_external(xyz, xyz.shape[0], queries, queries.shape[0], K, out,
                             d2.ctypes.data if with_d2 else None)
    return (out, d2) if with_d2 else out


def canonical_rows(knn: np.ndarray, permutation: np.ndarray | None, xyz_orig: np.ndarray):
    """Rows by original index, original indices, ascending (d2_gpu, original index)  (SURVEY 8c)."""
    knn = np.ascontiguousarray(knn, np.uint32)
    n, K = knn.shape
    xyz_orig = np.ascontiguousarray(xyz_orig, np.float32).reshape(-1, 3)
    idx = np.empty((n, K), np.uint32)
    d2 = np.empty((n, K), np.float32)
    if permutation is not None:
        permutation = np.ascontiguousarray(permutation, np.uint32)
    lib().kno_canonical_rows(knn, permutation.ctypes.data if permutation is not None else None,
                             xyz_orig, n, K, idx, d2)
    return idx, d2


def compare_canonical(idx_a, d2_a, idx_b, d2_b, ulp_slack: int = 0):
    """SURVEY 8c rule.  Rows must be identical, except that entries whose d2 equals the row's k-th
    d2 (within ``ulp_slack`` ULPs -- 0 vs the reference GPU code, 2 vs the kd-tree arithmetic,
    SURVEY F11) may differ: the reference breaks such ties by thread arrival order (F7).
    Returns (n_rows_identical, n_rows_boundary_tie_only, bad_rows)."""
    same = (idx_a == idx_b).all(axis=1)
    bad = []
    tie_only = 0
    for r in np.nonzero(~same)[0]:
        a, b = set(idx_a[r].tolist()), set(idx_b[r].tolist())
        kth = max(d2_a[r, -1], d2_b[r, -1])
        lo = kth
        for _ in range(ulp_slack):
            lo = np.nextafter(np.float32(lo), np.float32(0))
        ok = True
        for side_idx, side_d2, other in ((idx_a[r], d2_a[r], b), (idx_b[r], d2_b[r], a)):
            for v, d in zip(side_idx.tolist(), side_d2.tolist()):
                if v not in other and not (d >= lo):
                    ok = False
        if ok and len(a ^ b) > 0:
            tie_only += 1
        elif ok:  # same set, order differs only inside equal-d2 runs -> identical after canonical sort
            bad.append(int(r))
        else:
            bad.append(int(r))
    return int(same.sum()), tie_only, bad


# ------------------------------------------------------------------------------ the reference itself
def have_ref(name: str) -> bool:
    return (REF_DIR / name).exists()


class RefKdTree:
    """The reference's CPU arm: kd_tree.cpp unmodified, driven as test_knearests.cu:197-212."""

    def __init__(self):
        so = REF_DIR / "libref_kdtree.so"
        if not so.exists():
            raise FileNotFoundError(f"{so} missing: run `make -C oracle` where /root/reference exists")
        self.L = C.CDLL(str(so))
        self.L.ref_kdtree_knn.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, C.c_int, _i32p, C.c_void_p,
                                          C.c_void_p, C.c_int]
        self.L.ref_kdtree_knn.restype = C.c_int
        self.L.ref_kdtree_max_threads.restype = C.c_int
        self.L.ref_kdtree_knn_passes.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, C.c_int, _i32p, C.c_void_p,
                                                 C.c_int, C.c_int]
        self.L.ref_kdtree_knn_passes.restype = C.c_int

    def max_threads(self) -> int:
        return int(self.L.ref_kdtree_max_threads())

    def knn(self, xyz: np.ndarray, K: int, q0: int = 0, q1: int | None = None, threads: int = 0,
            with_d2: bool = False):
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        n = xyz.shape[0]
        q1 = n if q1 is None else q1
        out = np.empty((q1 - q0, K), np.int32)
        d2 = np.empty((q1 - q0, K), np.float32) if with_d2 else None
        times = (C.c_double * 2)()
        used = self.L.ref_kdtree_knn(xyz, n, q0, q1, K, out, d2.ctypes.data if with_d2 else None,
                                     C.addressof(times), threads)
        info = {"build_s": times[0], "query_s": times[1], "threads": used}
        return (out, d2, info) if with_d2 else (out, info)

    def knn_passes(self, xyz: np.ndarray, K: int, q0: int, q1: int, threads: int, passes: int):
        """One tree build, `passes` query passes over the points [q0, q1): (rows of the last pass, info)."""
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        out = np.empty((q1 - q0, K), np.int32)
        times = (C.c_double * (1 + passes))()
        used = self.L.ref_kdtree_knn_passes(xyz, xyz.shape[0], q0, q1, K, out, C.addressof(times), threads, passes)
        return out, {"build_s": times[0], "query_s": [times[1 + i] for i in range(passes)], "threads": used}


def ref_load_xyz(path: str) -> np.ndarray:
    """The reference's own loader + normaliser (test_knearests.cu:40-80), for pinning load_xyz."""
    L = C.CDLL(str(REF_DIR / "libref_loader.so"))
    L.ref_load_xyz.argtypes = [C.c_char_p, _f32p, C.c_int]
    L.ref_load_xyz.restype = C.c_int
    with open(path) as f:
        cap = int(f.readline().split()[0]) + 16
    xyz = np.empty((cap, 3), np.float32)
    n = L.ref_load_xyz(str(path).encode(), xyz, cap)
    if n < 0:
        raise IOError(path)
    return xyz[:n].copy()


class _RefProblem(C.Structure):  # knearests.h:3-16
    _fields_ = [("allocated_points", C.c_int), ("dimx", C.c_int), ("dimy", C.c_int), ("dimz", C.c_int),
                ("num_cell_offsets", C.c_int), ("d_cell_offsets", C.c_void_p),
                ("d_cell_offset_dists", C.c_void_p), ("d_cell_max", C.c_void_p),
                ("d_permutation", C.c_void_p), ("d_counters", C.c_void_p), ("d_ptrs", C.c_void_p),
                ("d_globcounter", C.c_void_p), ("d_stored_points", C.c_void_p),
                ("d_knearests", C.c_void_p)]


class RefGpu:
    """The reference's GPU arm: knearests.cu unmodified for sm_100a, one library per K (K is a
    compile-time macro there, params.h:4).  Needs a GPU; symbols are C++-mangled."""

    def __init__(self, K: int):
        so = REF_DIR / f"libref_knearests_k{K}.so"
        if not so.exists():
            raise FileNotFoundError(f"{so} missing")
        self.K = K
        self.L = L = C.CDLL(str(so))
        self.prepare = L._Z10kn_prepareP6float3i
        self.prepare.argtypes = [_f32p, C.c_int]
        self.prepare.restype = C.POINTER(_RefProblem)
        self.solve = L._Z8kn_solveP10kn_problem
        self.solve.argtypes = [C.POINTER(_RefProblem)]
        self.solve.restype = None
        self.free = L._Z7kn_freePP10kn_problem
        self.free.argtypes = [C.POINTER(C.POINTER(_RefProblem))]
        self.free.restype = None
        for nm in ("_Z16kn_get_knearestsP10kn_problem", "_Z18kn_get_permutationP10kn_problem",
                   "_Z13kn_get_pointsP10kn_problem"):
            fn = getattr(L, nm)
            fn.argtypes = [C.POINTER(_RefProblem)]
            fn.restype = C.c_void_p
        self.libc = C.CDLL(None)
        self.libc.free.argtypes = [C.c_void_p]

    def _take(self, fn_name: str, kn, count: int, dtype):
        p = getattr(self.L, fn_name)(kn)
        arr = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(count * np.dtype(dtype).itemsize,))
        out = arr.view(dtype).copy()
        self.libc.free(p)  # caller frees, test_knearests.cu:169-170
        return out

    def run(self, xyz: np.ndarray, fetch: bool = True):
        """kn_prepare + kn_solve (+ accessors) exactly as test_knearests.cu:143-171.
        Returns dict(knn[n,K] sorted space, permutation[n], points[n,3], dim, prepare_s, solve_s)."""
        import time
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        n = xyz.shape[0]
        t0 = time.perf_counter()
        kn = self.prepare(xyz, n)
        t1 = time.perf_counter()
        self.solve(kn)
        t2 = time.perf_counter()
        res = {"dim": kn.contents.dimx, "prepare_s": t1 - t0, "solve_s": t2 - t1, "n": n}
        if fetch:
            res["knn"] = self._take("_Z16kn_get_knearestsP10kn_problem", kn, n * self.K, np.uint32).reshape(n, self.K)
            res["permutation"] = self._take("_Z18kn_get_permutationP10kn_problem", kn, n, np.uint32)
            res["points"] = self._take("_Z13kn_get_pointsP10kn_problem", kn, n * 3, np.float32).reshape(n, 3)
        pk = C.POINTER(_RefProblem)(kn.contents)
        self.free(C.byref(pk))
        return res
