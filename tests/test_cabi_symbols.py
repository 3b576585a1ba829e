"""The C-ABI library loads without a GPU and exports every symbol include/*.h declares."""
import ctypes
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def declared_c_symbols():
    text = (ROOT / "include" / "knearests_c.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(knc_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(kn):
    L = kn.lib()
    names = declared_c_symbols()
    assert len(names) >= 15
    for nm in names:
        assert hasattr(L, nm), f"{nm} declared in include/knearests_c.h but not exported"
    assert sorted(kn.C_ABI_SYMBOLS) == names  # the Python binding knows exactly the declared set


def test_library_exports_reference_mangled_names(kn):
    # knearests.h keeps the reference's C++ signatures, so the mangled names must be the reference's
    # (nm of the reference's knearests.o)
    L = kn.lib()
    for nm in kn.CXX_ABI_SYMBOLS:
        assert hasattr(L, nm), nm


def test_public_struct_layout_matches_reference():
    # knearests.h:3-16 -- 5 ints then 9 pointers; a consumer compiled against the reference header
    # reads the same offsets
    text = (ROOT / "include" / "knearests.h").read_text()
    body = text[text.index("typedef struct {"):text.index("} kn_problem;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = re.findall(r"\b(allocated_points|dimx|dimy|dimz|num_cell_offsets|d_cell_offsets|d_cell_offset_dists|"
                        r"d_cell_max|d_permutation|d_counters|d_ptrs|d_globcounter|d_stored_points|d_knearests)\b", body)
    assert fields == ["allocated_points", "dimx", "dimy", "dimz", "num_cell_offsets", "d_cell_offsets",
                      "d_cell_offset_dists", "d_cell_max", "d_permutation", "d_counters", "d_ptrs",
                      "d_globcounter", "d_stored_points", "d_knearests"]


def test_version_and_grid_dim_need_no_gpu(kn, oracle):
    assert "sm_100a" in kn.version()
    # knearests.cu:249 at the BASELINE sizes
    for n, d in [(20626, 19), (300000, 46), (900000, 66), (10_000_000, 148), (100_000_000, 318),
                 (1_000_000_000, 686), (125_000_000, 343), (11545, 16), (1, 1)]:
        assert kn.grid_dim(n) == d == oracle.grid_dim(n)
    for n in np.random.default_rng(0).integers(1, 2 ** 31 - 1, 200):
        assert kn.grid_dim(int(n), 2.0) == oracle.grid_dim(int(n), 2.0)


def test_argument_errors_are_reported_not_fatal(kn):
    L = kn.lib()
    h = ctypes.c_void_p()
    pts = np.zeros((4, 3), np.float32)
    assert L.knc_prepare(None, 4, 8, ctypes.byref(h)) == 1  # KNC_ERR_INVALID
    assert b"NULL" in L.knc_last_error()
    assert L.knc_prepare(pts.ctypes.data, 0, 8, ctypes.byref(h)) == 1
    assert L.knc_prepare(pts.ctypes.data, 4, 100000, ctypes.byref(h)) == 1
    assert b"KNC_MAX_K" in L.knc_last_error()
    assert L.knc_solve(None) == 1
    assert L.knc_free(None) == 0
    with pytest.raises(kn.KnError):
        kn.KnProblem(pts, k=8, slab=(0, 4, 1, 3))  # slab without grid_dim


def test_bench_grid_density_rule():
    """bench.py's grid density (points per cell) for the BASELINE.json k values: first search radius just under two
    cell widths; never below 1 point per cell."""
    import math
    import cuda_knearests_b200 as K
    assert [K.fast_density(k) for k in (8, 20, 32, 50)] == [1.0, 1.0, 1.5, 2.1]
    for k in (20, 32, 50, 100):
        lam = k + 2.5 * math.sqrt(k) + 2.0
        r0_cells = (3.0 * lam / (4.0 * math.pi * K.fast_density(k))) ** (1.0 / 3.0)
        assert 1.9 < r0_cells < 2.0


def test_product_never_touches_the_oracle():
    """oracle/ is test infrastructure: nothing under cuda_knearests_b200/ may import, link or open it
    (bench.py's CPU baseline legs, smoke() and tests/ are the only users)."""
    import pathlib
    import re
    root = pathlib.Path(__file__).resolve().parents[1] / "cuda_knearests_b200"
    for f in list(root.glob("*.py")) + list((root / "csrc").glob("*.cu")) + list((root / "csrc").glob("*.h")) + \
            [root / "csrc" / "Makefile"]:
        text = f.read_text()
        assert not re.search(r"\boracle\b|_ref/|libkn_oracle|libref_", text), f"{f} refers to the oracle"


def test_ctypes_structs_match_the_header_as_compiled(kn, tmp_path):
    """The Python door restates knc_options / knc_info / knc_stats / knc_timings by hand: sizes and field offsets are
    compared with what a C compiler makes of include/knearests_c.h (a field added on one side only would corrupt memory)."""
    import ctypes as C
    import subprocess
    structs = {"knc_options": kn.Options, "knc_info": kn.Info, "knc_stats": kn.Stats, "knc_timings": kn.Timings}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "knearests_c.h"', 'int main(void) {']
    for cname, cls in structs.items():
        lines.append(f'  printf("{cname} %zu\\n", sizeof({cname}));')
        for fname, *_ in cls._fields_:
            lines.append(f'  printf("{cname}.{fname} %zu\\n", offsetof({cname}, {fname}));')
    lines += ['  return 0;', '}']
    src = tmp_path / "abi.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "abi"
    inc = Path(__file__).resolve().parent.parent / "include"
    subprocess.run(["gcc", "-std=c99", "-I", str(inc), str(src), "-o", str(exe)], check=True)
    got = dict(l.split() for l in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines())
    for cname, cls in structs.items():
        assert int(got[cname]) == C.sizeof(cls), cname
        for fname, *_ in cls._fields_:
            assert int(got[f"{cname}.{fname}"]) == getattr(cls, fname).offset, f"{cname}.{fname}"
