"""The drop-in boundary at the C++ level: a caller written against the reference's interface only
(tests/cpp/dropin_main.cu: params.h + knearests.h, the reference driver's call sequence) is compiled with
nvcc, linked against libknearests_b200.so and run on the reference's fixture.  Its output must be the
oracle's rows.  __graft_entry__.build() compiles the binary; the test compiles it itself if it is missing."""
import subprocess
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
BIN = ROOT / "tests" / "cpp" / "dropin_test"


def build_dropin():
    lib = ROOT / "cuda_knearests_b200"
    subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O2", f"-I{ROOT / 'include'}",
                    str(ROOT / "tests" / "cpp" / "dropin_main.cu"), "-o", str(BIN), f"-L{lib}", "-lknearests_b200",
                    "-Xlinker", f"-rpath={lib}"], check=True)


def test_dropin_binary_links_against_the_reference_symbols():
    """No GPU needed: the caller's undefined symbols are the reference's mangled names, and they resolve."""
    if not BIN.exists():
        build_dropin()
    out = subprocess.run(["nm", "-u", str(BIN)], capture_output=True, text=True, check=True).stdout
    for sym in ("_Z10kn_prepareP6float3i", "_Z8kn_solveP10kn_problem", "_Z16kn_get_knearestsP10kn_problem",
                "_Z18kn_get_permutationP10kn_problem", "_Z13kn_get_pointsP10kn_problem",
                "_Z14kn_print_statsP10kn_problem", "_Z7kn_freePP10kn_problem"):
        assert sym in out


@pytest.mark.gpu
def test_dropin_caller_gets_the_oracle_rows(oracle, pts20k, tmp_path):
    if not BIN.exists():
        build_dropin()
    import os
    pts20k.tofile(tmp_path / "points.f32")
    # (1) on the reference's grid (KN_GRID_DENSITY=3.1: 19^3 cells for this fixture): kn_print_stats
    #     prints the reference's numbers; (2) on the library's own grid: other cell statistics, the same neighbours
    for env, g in (({"KN_GRID_DENSITY": "3.1"}, oracle.Grid(pts20k)),
                   ({}, oracle.Grid(pts20k, dims=oracle.fast_dims(len(pts20k), 50)))):
        r = subprocess.run([str(BIN), str(tmp_path / "points.f32"), str(tmp_path / "neighbors.u32")],
                           capture_output=True, text=True, env={**os.environ, **env})
        assert r.returncode == 0, r.stderr
        if env:
            assert "Grid:  points per cell: 0 (min), 8 (max)" in r.stderr  # kn_print_stats, as the reference prints it
            assert "[3] => 2091" in r.stderr
        else:
            assert "Grid:  points per cell: " in r.stderr
        got = np.fromfile(tmp_path / "neighbors.u32", np.uint32).reshape(-1, 50)
        want = np.empty_like(got)
        want[g.permutation] = g.permutation[g.knn_exact(50)]
        assert np.array_equal(got, want)


@pytest.mark.gpu
def test_cli_driver_loads_xyz_like_the_reference_and_prints_its_lines(oracle, tmp_path):
    """tools/kn_cli: .xyz in, the reference driver's kind of stderr lines, neighbour table out."""
    cli = ROOT / "tools" / "kn_cli"
    if not cli.exists():
        import __graft_entry__ as G
        G.build()
    rng = np.random.default_rng(11)
    raw = (rng.random((30000, 3)) * [2.0, 1.0, 0.5]).astype(np.float32)
    xyz = tmp_path / "cloud.xyz"
    with open(xyz, "w") as f:
        f.write(f"{len(raw)}\n")
        for p in raw:
            f.write(f"{p[0]:.9g} {p[1]:.9g} {p[2]:.9g}\n")
    r = subprocess.run([str(cli), str(xyz), str(tmp_path / "nb.u32"), str(tmp_path / "pts.f32")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    # the lines scripts around the reference's driver parse: bbox (test_knearests.cu:77), num_cell_offsets (ungated,
    # knearests.cu:288), GPU memory used / kn_firstbuild / kn_solve (IF_VERBOSE, :342,200,376), Max visited ring
    # (ungated, :389), the statistics of kn_print_stats (:461-464)
    for needle in ("bbox [", "num_cell_offsets = 121", "GPU memory used: ", "kn_firstbuild: ", "kn_solve: ",
                   "Max visited ring: ", "Grid:  points per cell:"):
        assert needle in r.stderr, needle
    import re
    a, b = re.search(r"Max visited ring: ([0-9.eE+-]+) / ([0-9.eE+-]+)", r.stderr).groups()
    assert float(b) == 5.0 and 1.0 < float(a) < 6.0
    # second run: the cloud comes from the binary cache written beside the .xyz file, same bytes, same rows
    assert (tmp_path / "cloud.xyz.f32").exists()
    r2 = subprocess.run([str(cli), str(xyz), str(tmp_path / "nb2.u32"), str(tmp_path / "pts2.f32")],
                        capture_output=True, text=True)
    assert r2.returncode == 0 and "binary cache" in r2.stderr and "binary cache" not in r.stderr
    assert (tmp_path / "pts2.f32").read_bytes() == (tmp_path / "pts.f32").read_bytes()
    assert (tmp_path / "nb2.u32").read_bytes() == (tmp_path / "nb.u32").read_bytes()
    pts = np.fromfile(tmp_path / "pts.f32", np.float32).reshape(-1, 3)
    assert np.array_equal(pts, oracle.load_xyz(str(xyz)))  # loader == restated reference loader, bit for bit
    got = np.fromfile(tmp_path / "nb.u32", np.uint32).reshape(-1, 50)
    g = oracle.Grid(pts, dims=oracle.fast_dims(len(pts), 50))
    want = np.empty_like(got)
    want[g.permutation] = g.permutation[g.knn_exact(50)]
    assert np.array_equal(got, want)
