"""Two measurements outside the headline: external query points (knc_query_points, general kernel) and a clustered
cloud.  python profiles/extra_cases.py > gpurun_out/r2_extra_cases.txt"""
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import cuda_knearests_b200 as K  # noqa: E402

rng = np.random.default_rng(7)
# (1) queries that are not members of the cloud: 100 M uniform points, k = 20, 1 M uniform query points (host in, host out)
pts = K.generate_uniform_device(100_000_000)
with K.KnProblem(pts, k=20) as p:
    q = (rng.random((1_000_000, 3)) * 1000).astype(np.float32)
    p.query_points(q[:1000])
    for m in (100_000, 1_000_000):
        t0 = time.perf_counter()
        rows = p.query_points(q[:m])
        dt = time.perf_counter() - t0
        print(f"external queries: {m} points on the 100 M cloud, k = 20: {1e3 * dt:.2f} ms wall clock, host to host "
              f"({m / dt / 1e6:.1f} M queries/s)")
del pts

# (2) clustered cloud: 20 M points, 2000 Gaussian clusters (sigma 8 units) + 10 % uniform background, against 20 M uniform
def clustered(n):
    centers = rng.random((2000, 3)) * 1000
    c = centers[rng.integers(0, len(centers), n)] + rng.normal(0, 8.0, (n, 3))
    u = rng.random((n, 3)) * 1000
    m = rng.random(n) < 0.1
    c[m] = u[m]
    return np.clip(c, 0, 999.999).astype(np.float32)

for name, cloud in (("uniform", (rng.random((20_000_000, 3)) * 1000).astype(np.float32)), ("clustered", clustered(20_000_000))):
    for k in (20,):
        t = None
        for _ in range(3):
            with K.KnProblem(cloud, k=k) as p:
                p.solve()
                t = p.timings()
                st = p.stats()
        print(f"{name} 20 M points, k = {k}: build {t['build_ms']:.2f} ms, query {t['solve_ms']:.2f} ms, restarts {t['restarts']}, "
              f"handed-over blocks {t['handed_over_blocks']}, second retries {t['second_retries']}, max points per cell {st['max']}")
