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
a = ap.parse_args()
pts = K.generate_uniform_device(a.n)
for _ in range(a.steps):
    with K.KnProblem(pts, k=a.k, density=a.density, algorithm=a.algorithm) as p:
        p.solve()
        t = p.timings()
print({kk: (round(v, 3) if isinstance(v, float) else v) for kk, v in t.items()}, (p.info.dimx, p.info.dimy, p.info.dimz))
