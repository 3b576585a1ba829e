"""Randomised differential test: many small clouds of different shapes, sizes, k and grid densities, every one
compared bit for bit with the oracle (layout and rows).  Seeds are fixed; failures print the case."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def make_cloud(rng, kind, n):
    if kind == "uniform":
        return rng.random((n, 3)) * 1000
    if kind == "clustered":
        centres = rng.random((max(1, n // 200), 3)) * 1000
        return centres[rng.integers(0, len(centres), n)] + rng.normal(0, rng.choice([0.05, 2.0, 30.0]), (n, 3))
    if kind == "planar":
        p = rng.random((n, 3)) * 1000
        p[:, rng.integers(0, 3)] = rng.random() * 1000
        return p
    if kind == "line":
        t = rng.random((n, 1))
        return t * rng.random((1, 3)) * 1000
    if kind == "duplicates":
        base = rng.random((max(1, n // 5), 3)) * 1000
        return base[rng.integers(0, len(base), n)]
    if kind == "lattice":
        return np.floor(rng.random((n, 3)) * 12) * 80.0 + 20.0
    if kind == "out_of_range":
        return rng.random((n, 3)) * 1400 - 200
    raise ValueError(kind)


KINDS = ["uniform", "clustered", "planar", "line", "duplicates", "lattice", "out_of_range"]


@pytest.mark.parametrize("seed", range(16))
def test_random_clouds_match_the_oracle(kn, oracle, seed):
    rng = np.random.default_rng(1000 + seed)
    for case in range(30):
        kind = KINDS[(seed + case) % len(KINDS)]
        n = int(rng.choice([1, 2, 3, 17, 100, 999, 5000, 20000, 60000]))
        k = int(rng.choice([1, 2, 8, 13, 20, 32, 50, 64, 100]))
        density = float(rng.choice([0.0, 0.0, 0.5, 1.0, 3.1, 10.0]))  # 0: the library's own (non-cubic) grid
        alg = int(rng.choice([0, 0, 6, 1, 2, 4]))  # 6: second-chance pass of the select kernel over any retry list
        pts = np.ascontiguousarray(make_cloud(rng, kind, n), np.float32)
        tag = f"seed={seed} case={case} kind={kind} n={n} k={k} density={density} algorithm={alg}"
        with kn.KnProblem(pts, k=k, density=density, algorithm=alg) as p:
            p.solve()
            rows, perm, spts = p.get_knearests(), p.get_permutation(), p.get_points()
            d2 = p.get_dists()
        g = oracle.Grid(pts, density=density) if density > 0 else oracle.Grid(pts, dims=oracle.fast_dims(n, k))
        assert np.array_equal(perm, g.permutation), tag
        assert np.array_equal(spts, g.sorted_xyz), tag
        want, want_d2 = g.knn_exact(k, with_d2=True)
        assert np.array_equal(rows, want), tag
        assert np.array_equal(d2.view(np.uint32), want_d2.view(np.uint32)), tag


@pytest.mark.parametrize("seed", range(4))
def test_random_slabs_external_queries_and_streaming(kn, oracle, seed):
    """Random z-slab handles (with the halo the driver would give them), external query points and the
    streamed solve, on random cloud shapes -- against the whole-grid oracle."""
    import torch
    from cuda_knearests_b200 import slabs
    rng = np.random.default_rng(500 + seed)
    for case in range(8):
        kind = ["uniform", "clustered", "planar", "out_of_range"][(seed + case) % 4]
        n = int(rng.choice([3000, 20000, 90000]))
        k = int(rng.choice([8, 20, 32]))
        pts = np.ascontiguousarray(make_cloud(rng, kind, n), np.float32)
        tag = f"seed={seed} case={case} kind={kind} n={n} k={k}"
        dims = oracle.fast_dims(n, k) if case % 2 else (oracle.grid_dim(n),) * 3
        g = oracle.Grid(pts, dims=dims)
        D, plane = dims[2], dims[0] * dims[1]
        full = g.knn_exact(k)
        # streamed solve == plain solve == oracle
        pinned = torch.empty(full.shape, dtype=torch.int32, pin_memory=True)
        with kn.KnProblem(pts, k=k, grid_dims=dims) as p:
            got = p.solve_to_host(pinned.numpy().view(np.uint32), chunks=int(rng.integers(1, 9)))
            assert np.array_equal(got, full), tag
            q = np.ascontiguousarray(make_cloud(rng, "out_of_range", 300), np.float32)
            assert np.array_equal(p.query_points(q), oracle.brute_external(g.sorted_xyz, q, k)), tag
        # a random slab with generous halo: rows of its owned planes == whole-grid rows (shifted)
        if D < 8:
            continue
        ozb = int(rng.integers(0, D - 2))
        oze = int(rng.integers(ozb + 1, D + 1))
        cz = np.clip(np.floor(pts[:, 2] * np.float32(D) / np.float32(1000)).astype(int), 0, D - 1)
        for halo in (slabs.halo_planes(k, n, D), D):  # the driver's halo may be too thin for odd clouds: then it must say so
            zb, ze = max(0, ozb - halo), min(D, oze + halo)
            sub = np.ascontiguousarray(pts[(cz >= zb) & (cz < ze)])
            if len(sub) == 0:
                break
            base = int(g.ptrs[zb * plane])
            q0, q1 = int(g.ptrs[ozb * plane]) - base, int(g.ptrs[oze * plane]) - base
            with kn.KnProblem(sub, k=k, grid_dims=dims, slab=(zb, ze, ozb, oze)) as p:
                try:
                    p.solve()
                except kn.KnError as e:
                    assert e.code == 5 and halo < D, tag  # KNC_ERR_HALO, only possible with the thin halo
                    continue
                rows = p.get_knearests().astype(np.int64)
            want = full[base + q0: base + q1].astype(np.int64)
            want = np.where(want == 0xFFFFFFFF, want, want - base)
            assert np.array_equal(rows, want), tag + f" slab=({zb},{ze},{ozb},{oze})"
            break
