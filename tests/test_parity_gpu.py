"""GPU parity: the CUDA engine, called through its C ABI, against the oracle -- bit exact.

Grid build (layout, permutation) and neighbour rows are integer/index work: every comparison here is
array equality, no tolerance.  The engine's rows are defined by the total order
(d2, sorted position) with d2 = fma(dz,dz,fma(dx,dx,dy*dy)) (knearests.cu:125 as compiled), which
oracle.Grid.knn_exact computes independently on the CPU.
"""
import numpy as np
import pytest

from conftest import GOLDEN, assert_sets_match_golden_hash, set_hash

pytestmark = pytest.mark.gpu


def run_engine(kn, pts, K, **kw):
    with kn.KnProblem(pts, k=K, **kw) as p:
        p.solve()
        return {"knn": p.get_knearests(), "perm": p.get_permutation(), "points": p.get_points(),
                "stats": p.stats(), "dims": (p.info.dimx, p.info.dimy, p.info.dimz), "timings": p.timings(),
                "info": p.info, "k": K}


def default_grid(oracle, pts, K):
    """The oracle's build on the grid the library chooses when the caller names none (knc_fast_dims, restated in
    oracle/knn_oracle.c:kno_fast_dims)."""
    return oracle.Grid(pts, dims=oracle.fast_dims(len(pts), K))


REFERENCE_GRID = {"density": 3.1}  # the reference's cube, knearests.cu:249 (engine and oracle take the same keyword)


def assert_layout_equals_oracle(oracle, pts, res, **gridkw):
    g = oracle.Grid(pts, **gridkw) if gridkw else default_grid(oracle, pts, res["k"])
    assert res["dims"] == g.dims
    assert np.array_equal(res["perm"], g.permutation)          # stable counting sort, knearests.cu:51-60
    assert np.array_equal(res["points"], g.sorted_xyz)
    hist = np.bincount(g.counters, minlength=64)
    assert res["stats"]["histogram"][:63] == hist[:63].tolist()  # what kn_print_stats prints
    assert res["stats"]["total"] == pts.shape[0]
    return g


@pytest.mark.parametrize("grid", [{}, REFERENCE_GRID], ids=["fast_grid", "reference_grid"])
@pytest.mark.parametrize("K", [1, 8, 13, 20, 32, 50, 100, 200, 256])
def test_pts20k_rows_bit_exact(kn, oracle, pts20k, K, grid):
    res = run_engine(kn, pts20k, K, **grid)
    g = assert_layout_equals_oracle(oracle, pts20k, res, **grid)
    if grid:
        assert g.dim == 19  # the reference's grid for its own fixture
    assert np.array_equal(res["knn"], g.knn_exact(K))


@pytest.mark.parametrize("K", [8, 20, 50])
def test_pts20k_sets_match_reference_kdtree_golden(kn, oracle, pts20k, K):
    """The reference's own acceptance test (test_knearests.cu:194-232): set equality with kd_tree.cpp."""
    res = run_engine(kn, pts20k, K)
    idx, _ = oracle.canonical_rows(res["knn"], res["perm"], pts20k)
    golden = np.load(GOLDEN / f"pts20k_kdtree_k{K}_rowhash.npy")
    mism = np.nonzero(set_hash(idx) != golden)[0]
    # the only rows that may differ are the fixture's exact k/k+1 ties, checked row by row on the oracle
    g = default_grid(oracle, pts20k, K)
    assert np.array_equal(res["knn"], g.knn_exact(K))
    assert len(mism) == assert_sets_match_golden_hash(oracle, pts20k, g, K, golden)


@pytest.mark.parametrize("n,K,seed", [(11545, 50, 1), (100_000, 20, 2), (300_000, 50, 20171004),
                                      (1_000_000, 20, 20171004), (1_000_000, 8, 1), (500_000, 32, 2)])
@pytest.mark.parametrize("grid", [{}, REFERENCE_GRID, {"algorithm": 4}, {"algorithm": 6}, {"algorithm": 6, **REFERENCE_GRID}],
                         ids=["fast_grid", "reference_grid", "fast_grid_no_tile", "fast_grid_second_chance", "reference_grid_second_chance"])
def test_uniform_rows_bit_exact(kn, oracle, n, K, seed, grid):
    """fast_grid: the default grid, tile kernel (+ select kernel for blocks it hands over, + retry kernel);
    reference_grid: the reference's cube, select kernel; fast_grid_no_tile: the default grid without the tile kernel;
    *_second_chance: the retry list goes through the select kernel's second, larger sphere whatever its length
    (by default only a long list does -- clouds of millions of points)."""
    pts = oracle.generate_uniform(n, seed=seed)
    res = run_engine(kn, pts, K, **grid)
    g = assert_layout_equals_oracle(oracle, pts, res, **{k: v for k, v in grid.items() if k != "algorithm"})
    assert np.array_equal(res["knn"], g.knn_exact(K))
    if n >= 1_000_000:  # restarts are a cost, not an error; smaller clouds are mostly domain boundary
        assert res["timings"]["restarts"] < n // 20


@pytest.mark.parametrize("density", [0.5, 1.0, 2.0, 8.0])
def test_other_grid_densities(kn, oracle, density):
    pts = oracle.generate_uniform(200_000, seed=5)
    res = run_engine(kn, pts, 20, density=density)
    g = assert_layout_equals_oracle(oracle, pts, res, density=density)
    assert np.array_equal(res["knn"], g.knn_exact(20))


@pytest.mark.parametrize("K", [8, 20, 50, 77])
def test_general_kernel_alone_gives_the_same_rows(kn, oracle, K):
    """algorithm=1 runs every query through the general kernel (the one that also serves the retries
    of the fast kernel); the total order makes both paths produce identical rows."""
    pts = oracle.generate_uniform(120_000, seed=13)
    a = run_engine(kn, pts, K)
    b = run_engine(kn, pts, K, algorithm=1)
    assert np.array_equal(a["knn"], b["knn"])
    assert np.array_equal(a["knn"], default_grid(oracle, pts, K).knn_exact(K))


def test_search_radius_guess_does_not_change_rows(kn, oracle):
    pts = oracle.generate_uniform(150_000, seed=14)
    want = default_grid(oracle, pts, 20).knn_exact(20)
    for scale in (0.3, 0.7, 1.0, 1.6, 3.0, 6.0):  # tiny guess: everything retried; huge: beyond the row table
        r = run_engine(kn, pts, 20, search_scale=scale)
        assert np.array_equal(r["knn"], want), scale


def test_large_cloud_banded_build_is_the_same_stable_sort(kn, oracle):
    """From 20 M points on, the build first splits the input into z-bands (L2 locality); the layout must
    still be the deterministic stable counting sort, and identical to the build without the pre-pass."""
    n = 21_000_000
    pts = oracle.generate_uniform(n, seed=99)
    r = run_engine(kn, pts, 8)
    g = assert_layout_equals_oracle(oracle, pts, r)
    assert np.array_equal(r["knn"], g.knn_exact(8))
    plain = run_engine(kn, pts, 8, algorithm=3)  # no z-band pre-pass
    for key in ("knn", "perm", "points"):
        assert np.array_equal(r[key], plain[key])


def test_device_input_equals_host_input(kn, oracle):
    import torch
    n = 250_000
    dev = kn.generate_uniform_device(n, seed=20171004)
    host = oracle.generate_uniform(n, seed=20171004)
    assert np.array_equal(dev.cpu().numpy(), host)  # device generator == oracle generator, bit for bit
    a = run_engine(kn, host, 20)
    b = run_engine(kn, dev, 20)
    for key in ("knn", "perm", "points"):
        assert np.array_equal(a[key], b[key])
    # unaligned device pointer takes the scalar load path of the count kernel
    shifted = torch.empty(3 * n + 1, dtype=torch.float32, device="cuda")[1:].view(n, 3)
    shifted.copy_(dev)
    c = run_engine(kn, shifted, 20)
    assert np.array_equal(a["knn"], c["knn"]) and np.array_equal(a["perm"], c["perm"])


def test_solve_is_idempotent_and_deterministic(kn, oracle):
    pts = oracle.generate_uniform(150_000, seed=11)
    with kn.KnProblem(pts, k=20) as p:
        a = p.solve().get_knearests()
        b = p.solve().get_knearests()  # kn_solve may be called twice
    c = run_engine(kn, pts, 20)["knn"]
    assert np.array_equal(a, b) and np.array_equal(a, c)


def test_solve_to_host_streams_the_same_rows(kn, oracle):
    import torch
    pts = oracle.generate_uniform(333_333, seed=9)
    want = run_engine(kn, pts, 20)["knn"]
    for chunks in (1, 3, 8, 64):
        pinned = torch.empty(want.shape, dtype=torch.int32, pin_memory=True)
        with kn.KnProblem(pts, k=20) as p:
            got = p.solve_to_host(pinned.numpy().view(np.uint32), chunks=chunks)
            assert np.array_equal(got, want)
            assert np.array_equal(p.get_knearests(), want)  # rows also stay on the device
    with kn.KnProblem(pts, k=20) as p:
        assert np.array_equal(p.solve_to_host(), want)      # pageable buffer: same result, no overlap


def test_distances_and_external_query_points(kn, oracle):
    """distances of the rows, and queries that are not data points (what the reference's kd-tree
    offers, kd_tree.cpp:168-192).  Distances are bit-exact (same f32 expression), rows are exact."""
    pts = oracle.generate_uniform(80_000, seed=17)
    g = default_grid(oracle, pts, 20)
    rng = np.random.default_rng(5)
    with kn.KnProblem(pts, k=20) as p:
        p.solve()
        rows, d2 = p.get_knearests(), p.get_dists()
        want_rows, want_d2 = g.knn_exact(20, with_d2=True)
        assert np.array_equal(rows, want_rows)
        assert np.array_equal(d2.view(np.uint32), want_d2.view(np.uint32))
        q = np.concatenate([rng.random((3000, 3)).astype(np.float32) * 1000,   # anywhere in the domain
                            g.sorted_xyz[:200] + np.float32(1e-3),               # right next to data points
                            g.sorted_xyz[200:300],                                # exactly ON data points
                            np.float32([[-50, 500, 500], [1200, 1200, 1200]])])   # outside the domain
        got, got_d2 = p.query_points(q, with_d2=True)
        want, want_q_d2 = oracle.brute_external(g.sorted_xyz, q, 20, with_d2=True)
        assert np.array_equal(got, want)
        assert np.array_equal(got_d2.view(np.uint32), want_q_d2.view(np.uint32))
        assert (got[3200:3300, 0] == np.arange(200, 300)).all()  # a query on a data point finds it at distance 0
    with kn.KnProblem(pts, k=50) as p:  # works before knc_solve too (another k: another default grid, other positions)
        q = rng.random((500, 3)).astype(np.float32) * 1000
        assert np.array_equal(p.query_points(q), oracle.brute_external(p.get_points(), q, 50))


def test_original_space_remap(kn, oracle):
    pts = oracle.generate_uniform(60_000, seed=4)
    with kn.KnProblem(pts, k=8) as p:
        p.solve()
        knn, perm, orig = p.get_knearests(), p.get_permutation(), p.get_knearests_original()
    want = np.empty_like(orig)
    want[perm] = perm[knn]  # test_knearests.cu:155-160
    assert np.array_equal(orig, want)
    # the reference's checks: permutation is a bijection (:162-168), no duplicate in a row (:174-191)
    assert np.array_equal(np.sort(perm), np.arange(len(perm), dtype=np.uint32))
    srt = np.sort(orig, axis=1)
    assert (np.diff(srt.astype(np.int64), axis=1) > 0).all()


# ------------------------------------------------------------------------------------ edge cases
def test_single_point_and_fewer_points_than_k(kn, oracle):
    one = np.array([[500.0, 500.0, 500.0]], np.float32)
    r = run_engine(kn, one, 8)
    assert r["knn"].shape == (1, 8) and (r["knn"] == kn.EMPTY).all()
    pts = oracle.generate_uniform(7, seed=3)
    r = run_engine(kn, pts, 20)
    g = default_grid(oracle, pts, 20)
    assert np.array_equal(r["knn"], g.knn_exact(20))
    assert (r["knn"][:, 6:] == kn.EMPTY).all() and (r["knn"][:, :6] != kn.EMPTY).all()


def test_small_clouds_below_reference_minimum(kn, oracle):
    # the reference exits below D = 16 (knearests.cu:255-258); the engine accepts any n >= 1
    for n in (2, 33, 500, 5000):
        pts = oracle.generate_uniform(n, seed=n)
        r = run_engine(kn, pts, 8)
        g = assert_layout_equals_oracle(oracle, pts, r)
        assert np.array_equal(r["knn"], g.knn_exact(8))


def test_coincident_points(kn, oracle):
    # duplicates are neighbours of each other at distance 0; only the point itself is excluded (:123)
    base = oracle.generate_uniform(20_000, seed=8)
    pts = np.concatenate([base, base[:5000], base[:100], base[:100]])
    r = run_engine(kn, pts, 8)
    g = assert_layout_equals_oracle(oracle, pts, r)
    assert np.array_equal(r["knn"], g.knn_exact(8))


def test_points_outside_the_unit_range_are_clamped(kn, oracle):
    # knearests.cu:21-30: out-of-range points land in border cells; still found, still exact
    pts = oracle.generate_uniform(30_000, seed=6)
    pts[:300] = pts[:300] * 1.2 - 100.0
    pts[300:320] = [-50.0, 1050.0, 500.0]
    r = run_engine(kn, pts, 20)
    g = assert_layout_equals_oracle(oracle, pts, r)
    assert np.array_equal(r["knn"], g.knn_exact(20))


def test_presorted_and_repeated_input_takes_the_aggregated_atomics(kn, oracle):
    """Input that is already in cell order (every warp of the count kernel sees runs of equal cells: the
    warp-aggregated atomic path) and input with long runs of one repeated point."""
    base = oracle.generate_uniform(120_000, seed=23)
    g0 = default_grid(oracle, base, 8)
    for pts in (np.ascontiguousarray(g0.sorted_xyz),                      # cell-ordered input
                np.ascontiguousarray(np.repeat(base[:3000], 40, axis=0))):  # 40 copies of each point, adjacent
        r = run_engine(kn, pts, 8)
        g = assert_layout_equals_oracle(oracle, pts, r)
        assert np.array_equal(r["knn"], g.knn_exact(8))


def test_clustered_cloud_overfull_cells_and_restarts(kn, oracle):
    # far from uniform: most of the cloud in two tiny blobs (cells with thousands of points -> the
    # block-wide ordering path of the build), the rest sparse (radius guess far too small -> restarts)
    rng = np.random.default_rng(5)
    blob1 = (rng.normal(0, 0.4, (6000, 3)) + [200, 300, 400]).astype(np.float32)
    blob2 = (rng.normal(0, 2.0, (3000, 3)) + [800, 800, 100]).astype(np.float32)
    sparse = oracle.generate_uniform(3000, seed=12)
    pts = np.ascontiguousarray(np.concatenate([blob1, sparse, blob2])[rng.permutation(12000)])
    r = run_engine(kn, pts, 20)
    g = assert_layout_equals_oracle(oracle, pts, r)
    assert r["stats"]["max"] > 1000
    assert np.array_equal(r["knn"], g.knn_exact(20))
    assert r["timings"]["restarts"] > 0


@pytest.mark.parametrize("K,algorithm", [(20, 0), (20, 6), (50, 6), (8, 0)])
def test_clusters_on_a_background_local_density(kn, oracle, K, algorithm):
    """400 k points: 60 Gaussian clusters (sigma 6) on a 10 % uniform background.  The first radius comes from the MEAN
    density: inside a cluster its sphere overflows every buffer, in the background it is nearly empty.  The select
    kernel sizes the sphere by the local density there (algorithm 6: also in its second-chance pass, whatever the
    length of the retry list); rows must not change."""
    rng = np.random.default_rng(77)
    n = 400_000
    centers = rng.random((60, 3)) * 1000
    pts = centers[rng.integers(0, 60, n)] + rng.normal(0, 6.0, (n, 3))
    back = rng.random(n) < 0.1
    pts[back] = rng.random((int(back.sum()), 3)) * 1000
    pts = np.ascontiguousarray(np.clip(pts, 0, 999.99), np.float32)
    r = run_engine(kn, pts, K, algorithm=algorithm)
    g = assert_layout_equals_oracle(oracle, pts, r)
    assert np.array_equal(r["knn"], g.knn_exact(K))
    # queries that are not members of the cloud, on the same data
    q = np.concatenate([rng.random((2000, 3)).astype(np.float32) * 1000, g.sorted_xyz[::400] + np.float32(0.01)])
    with kn.KnProblem(pts, k=K, algorithm=algorithm) as p:
        got, got_d2 = p.query_points(q, with_d2=True)
    want, want_d2 = oracle.brute_external(g.sorted_xyz, q, K, with_d2=True)
    assert np.array_equal(got, want)
    assert np.array_equal(got_d2.view(np.uint32), want_d2.view(np.uint32))


def test_planar_and_lattice_clouds_with_many_ties(kn, oracle):
    # integer lattice: massive distance ties, including at the k/k+1 border -> the stated tie-break
    ax = np.arange(0, 40, dtype=np.float32) * 25.0 + 12.5
    lattice = np.stack(np.meshgrid(ax, ax, ax, indexing="ij"), -1).reshape(-1, 3)
    r = run_engine(kn, lattice, 20)
    g = assert_layout_equals_oracle(oracle, lattice, r)
    assert np.array_equal(r["knn"], g.knn_exact(20))
    plane = oracle.generate_uniform(40_000, seed=2)
    plane[:, 2] = 500.0
    r = run_engine(kn, plane, 8)
    g = assert_layout_equals_oracle(oracle, plane, r)
    assert np.array_equal(r["knn"], g.knn_exact(8))


@pytest.mark.parametrize("bits", [7, 9, 11, 13])
def test_quantised_coordinates_many_ties_and_near_ties(kn, oracle, bits):
    """Coordinates on a 2^bits lattice: squared distances collide exactly or differ in their last bits
    everywhere, including at the k/k+1 border and across the select kernel's sorting rounds -- the cases
    in which a truncated-key comparison must hand the query to the exact path."""
    raw = oracle.generate_uniform(200_000, seed=40 + bits)
    pts = (np.floor(raw / 1000.0 * 2 ** bits) * (1000.0 / 2 ** bits)).astype(np.float32)
    for K in (8, 20):
        r = run_engine(kn, pts, K)
        g = assert_layout_equals_oracle(oracle, pts, r)
        assert np.array_equal(r["knn"], g.knn_exact(K))


def test_slab_mode_rows_equal_whole_grid_rows(kn, oracle):
    """z-slab handles: planes [zb,ze) held, queries for [ozb,oze) only; the rows are the
    whole-grid rows, shifted by the slab's first sorted position."""
    pts = oracle.generate_uniform(400_000, seed=20171004)
    K = 20
    for dims, slabs in [((51, 51, 51), [(0, 14, 0, 10), (6, 30, 10, 26), (35, 51, 39, 51)]),   # the reference's cube
                        (oracle.fast_dims(len(pts), K), [(0, 9, 0, 6), (4, 20, 7, 17), (26, 36, 29, 36)])]:
        _check_slabs(kn, oracle, pts, K, dims, slabs)


def _check_slabs(kn, oracle, pts, K, dims, slabs):
    g = oracle.Grid(pts, dims=dims)
    D, plane = dims[2], dims[0] * dims[1]
    assert slabs[-1][1] == D
    full = g.knn_exact(K)
    cz = np.clip(np.floor(pts[:, 2] * np.float32(D) / np.float32(1000)).astype(int), 0, D - 1)
    for (zb, ze, ozb, oze) in slabs:
        keep = (cz >= zb) & (cz < ze)
        sub = np.ascontiguousarray(pts[keep])
        with kn.KnProblem(sub, k=K, grid_dims=dims, slab=(zb, ze, ozb, oze)) as p:
            p.solve()
            rows, perm, info = p.get_knearests(), p.get_permutation(), p.info
        # global sorted positions of the slab: contiguous, starting at ptrs[zb * plane]
        base = int(g.ptrs[zb * plane])
        q0 = int(g.ptrs[ozb * plane]) - base
        q1 = int(g.ptrs[oze * plane]) - base
        assert (info.query_begin, info.query_count) == (q0, q1 - q0)
        want = full[base + q0: base + q1].astype(np.int64) - base
        assert np.array_equal(rows.astype(np.int64), want)
        assert np.array_equal(np.nonzero(keep)[0][perm], g.permutation[base: base + sub.shape[0]])


def test_fourth_word_of_16_byte_records_is_carried_into_sorted_order(kn, oracle):
    """knc_get_tags (no reference counterpart): [n, 4] device records x, y, z, tag  ->  tags[s] = tag of the point at
    sorted position s, i.e. tag_in[permutation[s]]; aligned and unaligned record arrays, with and without the
    z-band pre-pass; xyz input has no tags."""
    import torch
    from cuda_knearests_b200 import slabs
    n = 300_000
    pts = oracle.generate_uniform(n, seed=5)
    tag = np.random.default_rng(5).integers(0, 2**32, n, dtype=np.uint32)
    rec = np.empty((n, 4), np.float32)
    rec[:, :3] = pts
    rec[:, 3] = tag.view(np.float32)
    flat = torch.zeros(4 * n + 1, dtype=torch.float32, device="cuda")
    for shift in (0, 1):  # 1: records start 4 bytes off a 16-byte boundary
        view = flat[shift: shift + 4 * n].view(n, 4)
        view.copy_(torch.from_numpy(rec))
        with kn.KnProblem(view, k=8) as p:
            perm = p.get_permutation()
            got = torch.as_tensor(slabs._DevView(p.tags_ptr(), (n,)), device="cuda").cpu().numpy().view(np.uint32)
            assert np.array_equal(got, tag[perm])
            assert np.array_equal(p.get_points(), pts[perm])
    with kn.KnProblem(torch.from_numpy(pts).cuda(), k=8) as p:
        with pytest.raises(kn.KnError) as e:
            p.tags_ptr()
        assert e.value.code == 4  # KNC_ERR_STATE


def test_slab_mode_reports_insufficient_halo(kn, oracle):
    pts = oracle.generate_uniform(100_000, seed=3)
    D = oracle.grid_dim(len(pts))
    cz = np.clip(np.floor(pts[:, 2] * np.float32(D) / np.float32(1000)).astype(int), 0, D - 1)
    sub = np.ascontiguousarray(pts[(cz >= 10) & (cz < 20)])
    with kn.KnProblem(sub, k=20, grid_dim=D, slab=(10, 20, 10, 20)) as p:  # no halo at all
        with pytest.raises(kn.KnError) as e:
            p.solve()
        assert e.value.code == 5  # KNC_ERR_HALO


# ------------------------------------------------------------------------ BASELINE sizes: properties
def test_ten_million_points_properties_and_sampled_brute_force(kn, oracle):
    """configs[2]: 10 M uniform, k = 20.  Too big for the O(n) oracle per query on all rows, so:
    size-independent properties on everything + exact brute force on sampled queries."""
    import torch
    n, K = 10_000_000, 20
    dev = kn.generate_uniform_device(n, seed=20171004)
    with kn.KnProblem(dev, k=K) as p:
        p.solve()
        knn = p.get_knearests()
        perm = p.get_permutation()
        pts = p.get_points()
        assert (p.info.dimx, p.info.dimy, p.info.dimz) == oracle.fast_dims(n, K)
    # permutation is a bijection (test_knearests.cu:162-168)
    seen = np.zeros(n, np.uint8)
    seen[perm] = 1
    assert seen.all()
    host = dev.cpu().numpy()
    assert np.array_equal(pts[:100000], host[perm[:100000]])
    assert (knn != kn.EMPTY).all() and (knn != np.arange(n, dtype=np.uint32)[:, None]).all()
    # rows ascending in (d2, position), d2 recomputed with the reference formula, on 1 M sampled rows
    rows = np.random.default_rng(0).choice(n, 200_000, replace=False)
    rows.sort()
    q = pts[rows].astype(np.float32)
    c = pts[knn[rows]]
    d = c - q[:, None, :]
    dx, dy, dz = d[..., 0].astype(np.float64), d[..., 1].astype(np.float64), d[..., 2].astype(np.float64)
    f32 = np.float32
    t = (dx * dx + (dy * dy).astype(f32).astype(np.float64)).astype(f32)   # fma(dx,dx,rn(dy*dy))
    d2 = (dz * dz + t.astype(np.float64)).astype(f32)                       # fma(dz,dz,t)
    assert (np.diff(d2, axis=1) >= 0).all()
    tie = np.diff(d2, axis=1) == 0
    assert (np.diff(knn[rows].astype(np.int64), axis=1)[tie] > 0).all()
    # exact brute force over all 10 M candidates for 96 sampled queries
    for q0 in (0, 4_999_968, n - 32):
        want = oracle.brute(pts, K, q0, q0 + 32)
        assert np.array_equal(knn[q0:q0 + 32], want)
