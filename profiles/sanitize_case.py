"""Small end-to-end case for compute-sanitizer (one tool per gpurun call):
   compute-sanitizer --tool memcheck python profiles/sanitize_case.py"""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import cuda_knearests_b200 as K  # noqa: E402

pts = np.load(Path(__file__).resolve().parent.parent / "tests" / "golden" / "pts20k_norm.npy")
for k, alg in ((50, 0), (20, 0), (20, 2), (13, 0), (8, 1)):
    with K.KnProblem(pts, k=k, algorithm=alg) as p:
        p.solve()
        rows = p.get_knearests()
        assert rows.shape == (len(pts), k)
        p.get_dists()
        p.query_points(pts[:100] + np.float32(0.5))
rng = np.random.default_rng(0)
big = (rng.random((300_000, 3)) * 1000).astype(np.float32)
with K.KnProblem(big, k=20, density=1.0) as p:
    p.solve()
    p.get_knearests_original()
print("sanitize case done")
