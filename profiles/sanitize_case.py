"""Small end-to-end cases that walk every kernel once (written for compute-sanitizer, one tool per call:
   compute-sanitizer --tool memcheck python profiles/sanitize_case.py [big]; the tool is closed on this pool in round 2,
   so the script only serves as a quick walk-through there).
   `big` adds a 4.2 M-point cloud (band split of the build, tile kernel at scale); leave it out for racecheck."""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import cuda_knearests_b200 as K  # noqa: E402

pts = np.load(Path(__file__).resolve().parent.parent / "tests" / "golden" / "pts20k_norm.npy")
for k, alg in ((50, 0), (20, 0), (20, 2), (13, 0), (8, 1), (20, 4), (20, 6), (50, 6), (32, 6), (8, 6)):
    with K.KnProblem(pts, k=k, algorithm=alg) as p:
        p.solve()
        rows = p.get_knearests()
        assert rows.shape == (len(pts), k)
        p.get_dists()
        p.query_points(pts[:100] + np.float32(0.5))
rng = np.random.default_rng(0)
# clusters on a background: cells of hundreds of points (block-wide ordering of the build), local-density spheres
centers = rng.random((8, 3)) * 1000
cl = centers[rng.integers(0, 8, 60_000)] + rng.normal(0, 1.5, (60_000, 3))
cl[:6000] = rng.random((6000, 3)) * 1000
cl = np.ascontiguousarray(np.clip(cl, 0, 999.9), np.float32)
for k, alg in ((20, 0), (20, 6), (50, 6)):
    with K.KnProblem(cl, k=k, algorithm=alg) as p:
        p.solve()
        assert p.stats()["max"] > 32
big = (rng.random((300_000, 3)) * 1000).astype(np.float32)
for dens in (1.0, 0.0):
    with K.KnProblem(big, k=20, density=dens) as p:
        p.solve()
        p.get_knearests_original()
if len(sys.argv) > 1 and sys.argv[1] == "big":
    huge = (rng.random((4_200_000, 3)) * 1000).astype(np.float32)
    for k in (8, 20):
        with K.KnProblem(huge, k=k) as p:
            p.solve()
            assert p.timings()["restarts"] > 0
print("sanitize case done")
