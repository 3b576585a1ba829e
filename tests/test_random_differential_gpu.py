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
        density = float(rng.choice([0.0, 0.5, 1.0, 3.1, 10.0]))
        alg = int(rng.choice([0, 0, 0, 1, 2]))
        pts = np.ascontiguousarray(make_cloud(rng, kind, n), np.float32)
        tag = f"seed={seed} case={case} kind={kind} n={n} k={k} density={density} algorithm={alg}"
        with kn.KnProblem(pts, k=k, density=density, algorithm=alg) as p:
            p.solve()
            rows, perm, spts = p.get_knearests(), p.get_permutation(), p.get_points()
            d2 = p.get_dists()
        g = oracle.Grid(pts, density=density if density > 0 else 3.1)
        assert np.array_equal(perm, g.permutation), tag
        assert np.array_equal(spts, g.sorted_xyz), tag
        want, want_d2 = g.knn_exact(k, with_d2=True)
        assert np.array_equal(rows, want), tag
        assert np.array_equal(d2.view(np.uint32), want_d2.view(np.uint32)), tag
