"""The oracle against the reference's golden vectors and against the reference itself (CPU, no GPU).

This is what pins oracle/knn_oracle.c: (1) the committed vectors in tests/golden/ -- made by the
reference's own loader and kd-tree from the reference's only fixture, pts20K.xyz -- and (2), when
oracle/_ref/ is present (it is built from /root/reference by oracle/Makefile and travels to the GPU
box), the unmodified reference kd_tree.cpp run live on fresh synthetic clouds.
"""
import json

import numpy as np
import pytest

from conftest import GOLDEN, assert_sets_match_golden_hash, set_hash


def test_grid_facts_match_fixture(oracle, pts20k):
    # D = 19, occupancy histogram of the 6 859 cells (what kn_print_stats prints)
    facts = json.loads((GOLDEN / "pts20k_facts.json").read_text())
    g = oracle.Grid(pts20k)
    assert g.dim == facts["dim"] == 19
    hist = np.bincount(g.counters)
    assert {str(i): int(c) for i, c in enumerate(hist)} == facts["occupancy_histogram"]
    assert int(g.ptrs[-1]) == pts20k.shape[0]
    # permutation is a bijection (test_knearests.cu:162-168) and the layout is the stable counting sort
    assert np.array_equal(np.sort(g.permutation), np.arange(pts20k.shape[0], dtype=np.uint32))
    assert np.array_equal(g.sorted_xyz, pts20k[g.permutation])
    for c in np.nonzero(g.counters > 1)[0][:500]:
        seg = g.permutation[g.ptrs[c]:g.ptrs[c + 1]]
        assert np.all(np.diff(seg.astype(np.int64)) > 0)


@pytest.mark.parametrize("K", [8, 20, 50])
def test_exact_rows_match_reference_kdtree_golden(oracle, pts20k, K):
    """Set equality with the reference kd-tree on every row (test_knearests.cu:216-231)."""
    g = oracle.Grid(pts20k)
    golden = np.load(GOLDEN / f"pts20k_kdtree_k{K}_rowhash.npy")
    ties = assert_sets_match_golden_hash(oracle, pts20k, g, K, golden)
    # the fixture has a handful of exact k/k+1 ties at k=8 (lattice-like points on the cube faces)
    assert ties <= {8: 3, 20: 3, 50: 1}[K]


def test_exact_rows_order_matches_golden_rows(oracle, pts20k):
    g = oracle.Grid(pts20k)
    idx, d2 = oracle.canonical_rows(g.knn_exact(50), g.permutation, pts20k)
    gold = np.load(GOLDEN / "pts20k_kdtree_k50_rows512.npy").astype(np.uint32)
    assert np.array_equal(np.sort(idx[:512], axis=1), np.sort(gold, axis=1))
    # nearest-first order agrees wherever the distances are distinct
    strict = np.all(np.diff(d2[:512], axis=1) > 0, axis=1)
    assert strict.sum() > 400
    assert np.array_equal(idx[:512][strict], gold[strict])


@pytest.mark.parametrize("K", [8, 50])
def test_refwalk_restatement_agrees_with_exact(oracle, pts20k, K):
    """The line-by-line restatement of the reference kernel (knearests.cu:93-148) and the engine
    contract give the same canonical rows on the fixture (no boundary ties there)."""
    g = oracle.Grid(pts20k)
    walk, exhausted = g.knn_refwalk(K)
    assert exhausted == 0
    ia, da = oracle.canonical_rows(g.knn_exact(K), g.permutation, pts20k)
    ib, db = oracle.canonical_rows(walk, g.permutation, pts20k)
    same, ties, bad = oracle.compare_canonical(ia, da, ib, db)
    assert bad == [] and same + ties == pts20k.shape[0]


def test_exact_equals_brute_force(oracle):
    pts = oracle.generate_uniform(6000, seed=1)
    g = oracle.Grid(pts, dim=8)
    for K in (1, 8, 33):
        assert np.array_equal(g.knn_exact(K), oracle.brute(g.sorted_xyz, K))


def test_generator_is_on_the_24bit_lattice(oracle):
    a = oracle.generate_uniform(1000, seed=20171004)
    b = oracle.generate_uniform(400, seed=20171004, first=600)
    assert np.array_equal(a[600:], b)  # counter based: any slice on its own
    assert a.min() >= 0 and a.max() < 1000
    lattice = a.astype(np.float64) / 1000 * 2 ** 24
    assert np.abs(lattice - np.round(lattice)).max() < 0.51  # 1000*u rounds once in f32


def test_loader_restatement(oracle, tmp_path):
    """kno_load_xyz follows test_knearests.cu:15-80; checked against the reference loader when built."""
    rng = np.random.default_rng(3)
    raw = rng.random((2000, 3)).astype(np.float32) * np.float32([1.0, 0.5, 0.25])
    path = tmp_path / "cloud.xyz"
    with open(path, "w") as f:
        f.write(f"{raw.shape[0]}\n")
        for p in raw:
            f.write(f"{p[0]:.9g} {p[1]:.9g} {p[2]:.9g}\n\n")  # blank lines are skipped (:50)
    mine = oracle.load_xyz(str(path))
    assert mine.shape == raw.shape
    assert 0.9 < mine[:, 0].min() < 1.1 and 998.9 < mine[:, 0].max() < 999.1  # 0.1 % margin (:28-37)
    assert mine[:, 2].max() < 260
    if oracle.have_ref("libref_loader.so"):
        assert np.array_equal(mine, oracle.ref_load_xyz(str(path)))


def test_against_live_reference_kdtree(oracle):
    """oracle/_ref/libref_kdtree.so = the reference's kd_tree.cpp, unmodified.  120 k uniform points."""
    if not oracle.have_ref("libref_kdtree.so"):
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    pts = oracle.generate_uniform(120000, seed=2)
    g = oracle.Grid(pts)
    kd = oracle.RefKdTree()
    for K in (20, 50):
        ia, da = oracle.canonical_rows(g.knn_exact(K), g.permutation, pts)
        ref, _ = kd.knn(pts, K)
        ib, db = oracle.canonical_rows(ref.astype(np.uint32), None, pts)
        same, ties, bad = oracle.compare_canonical(ia, da, ib, db, ulp_slack=2)
        assert bad == [] and same + ties == pts.shape[0]
        assert ties < 20


def test_edge_cases_oracle(oracle):
    # fewer points than k: rows padded with UINT_MAX (knearests.cu:108 sentinel)
    pts = oracle.generate_uniform(5, seed=7)
    rows = oracle.Grid(pts, dim=2).knn_exact(8)
    assert (rows[:, :4] != oracle.EMPTY).all() and (rows[:, 4:] == oracle.EMPTY).all()
    # coincident points are neighbours of each other at distance 0, self excluded by position (:123)
    dup = np.repeat(oracle.generate_uniform(50, seed=9), 3, axis=0)
    g = oracle.Grid(dup, dim=3)
    rows, d2 = g.knn_exact(4, with_d2=True)
    assert (d2[:, :2] == 0).all() and (d2[:, 2] > 0).all()
    assert (rows != np.arange(150, dtype=np.uint32)[:, None]).all()


def test_exact_equals_brute_force_on_random_shapes(oracle):
    """The grid-accelerated oracle against plain brute force on clouds that break the uniform assumptions:
    clusters, planes, lines, duplicates, lattices, points outside [0,1000]^3 (clamped into border cells)."""
    from test_random_differential_gpu import KINDS, make_cloud
    rng = np.random.default_rng(77)
    for case in range(70):
        kind = KINDS[case % len(KINDS)]
        n = int(rng.choice([1, 2, 5, 60, 400, 2500]))
        k = int(rng.choice([1, 3, 8, 20, 50]))
        density = float(rng.choice([0.5, 1.0, 3.1, 10.0]))
        pts = np.ascontiguousarray(make_cloud(rng, kind, n), np.float32)
        g = oracle.Grid(pts, density=density)
        assert np.array_equal(g.knn_exact(k), oracle.brute(g.sorted_xyz, k)), (case, kind, n, k, density)


def test_exact_walk_on_non_cubic_grids_equals_brute_force(oracle):
    """The engine's default grid is not a cube (narrow x cells, y/z cells as wide as the search radius): the oracle's
    exact walk generalises the reference's Chebyshev rings to box shells of equal thickness in space, and must give
    the brute-force rows on any resolution triple -- including degenerate ones."""
    pts = oracle.generate_uniform(40_000, seed=3)
    want_idx, _ = oracle.canonical_rows(oracle.brute(pts, 20), None, pts)
    for dims in [oracle.fast_dims(len(pts), 20), oracle.fast_dims(len(pts), 50), (103, 12, 12), (7, 40, 19), (1, 1, 1),
                 (300, 5, 5), (2, 3, 500)]:
        g = oracle.Grid(pts, dims=dims)
        assert g.dims == tuple(dims) and int(g.ptrs[-1]) == len(pts)
        idx, _ = oracle.canonical_rows(g.knn_exact(20), g.permutation, pts)
        assert np.array_equal(idx, want_idx), dims


def test_default_grid_rule_is_the_same_in_library_and_oracle(kn, oracle):
    """knc_fast_dims (host code of the library, no GPU needed) against the oracle's restatement of the rule."""
    for n in (1, 2, 100, 11_545, 20_626, 300_000, 10_000_000, 100_000_000, 125_000_000, 1_000_000_000, 2_000_000_000):
        for k in (1, 8, 20, 32, 50, 100, 256):
            assert kn.fast_dims(n, k) == oracle.fast_dims(n, k), (n, k)
    dx, dy, dz = kn.fast_dims(100_000_000, 20)
    assert dy == dz and 0.9 < dx * dy * dz / 1e8 < 1.1               # about one cell per point
    r0 = (3 * (20 + 2.5 * 20 ** 0.5 + 2) / (4 * np.pi * 0.1)) ** (1 / 3)
    assert 1.0 < (1000 / dy) / r0 < 1.05                              # y/z cells just wider than the first radius
