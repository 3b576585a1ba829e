"""One or more prepare + solve passes on a synthetic uniform cloud, nothing else: the short command that ncu wraps.
   python profiles/run_case.py --n 100000000 --k 20 [--algorithm A] [--density D] [--steps S]"""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import cuda_knearests_b200 as K  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=100_000_000)
ap.add_argument("--k", type=int, default=20)
ap.add_argument("--algorithm", type=int, default=0)
ap.add_argument("--density", type=float, default=0.0)
ap.add_argument("--steps", type=int, default=2)
ap.add_argument("--search-scale", type=float, default=0.0)
ap.add_argument("--yz-margin", type=float, default=0.0, help="y/z cell width over the first radius (library: 1.02)")
ap.add_argument("--cells-per-point", type=float, default=1.0)
ap.add_argument("--no-solve", action="store_true", help="grid build only")
a = ap.parse_args()
dims = None
if a.yz_margin > 0:
    import math
    r0 = (3 * (a.k + 2.5 * math.sqrt(a.k) + 2) / (4 * math.pi * a.n / 1e9)) ** (1 / 3)
    dyz = int(1000 / (a.yz_margin * r0))
    dims = (max(1, round(a.cells_per_point * a.n / dyz / dyz)), dyz, dyz)
pts = K.generate_uniform_device(a.n)
for _ in range(a.steps):
    with K.KnProblem(pts, k=a.k, density=a.density, algorithm=a.algorithm, search_scale=a.search_scale, grid_dims=dims) as p:
        if not a.no_solve:
            p.solve()
        t = p.timings()
print({kk: (round(v, 3) if isinstance(v, float) else v) for kk, v in t.items()}, (p.info.dimx, p.info.dimy, p.info.dimz))
