"""bench.py's after-the-run checker (verify_against_oracle) on the CPU: fed with the oracle's own grid and rows it
must report zero mismatches, and it must notice a wrong neighbour, a wrong permutation entry and a shifted cell."""
import sys
from types import SimpleNamespace

import numpy as np

from conftest import ROOT

sys.path.insert(0, str(ROOT))


def _case(oracle, n=30_000, k=8, dims=None):
    import bench
    import torch
    pts = oracle.generate_uniform(n, seed=bench.SEED)
    g = oracle.Grid(pts, density=1.0)
    rows = torch.from_numpy(g.knn_exact(k).astype(np.int64).astype(np.int32, casting="unsafe"))
    perm = torch.from_numpy(g.permutation.astype(np.int32, casting="unsafe"))
    ptrs = torch.from_numpy(g.ptrs)
    info = SimpleNamespace(query_count=n, query_begin=0, dimx=g.dim, dimy=g.dim, dimz=g.dim, z_begin=0, z_end=g.dim, n=n)
    return bench, torch, rows, perm, ptrs, info, n, k


def test_verify_accepts_the_oracle_rows(oracle):
    bench, torch, rows, perm, ptrs, info, n, k = _case(oracle)
    v = bench.verify_against_oracle(torch, rows, perm, ptrs, info, 0, n, k, samples=48)
    assert v["queries"] >= 40 and v["mismatches"] == 0


def test_verify_notices_a_wrong_row_and_a_wrong_layout(oracle):
    bench, torch, rows, perm, ptrs, info, n, k = _case(oracle)
    bad = rows.clone()
    bad[0, 0] = (int(bad[0, 0]) + 7) % n      # row 0 is always sampled
    assert bench.verify_against_oracle(torch, bad, perm, ptrs, info, 0, n, k, samples=8)["rows_wrong"] >= 1
    badperm = perm.clone()
    badperm[0] = perm[1]
    badperm[1] = perm[0]
    assert bench.verify_against_oracle(torch, rows, badperm, ptrs, info, 0, n, k, samples=8)["mismatches"] >= 1
