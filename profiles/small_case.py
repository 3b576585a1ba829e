"""kn_prepare + kn_solve on a small cloud, a few times: the command ncu wraps for the small BASELINE configs.
   python profiles/small_case.py pts20k|u300k|u900k [k]"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import cuda_knearests_b200 as K  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "pts20k"
k = int(sys.argv[2]) if len(sys.argv) > 2 else 50
if name == "pts20k":
    pts = np.load(ROOT / "tests/golden/pts20k_norm.npy")
else:
    n = {"u300k": 300_000, "u900k": 900_000}[name]
    import torch
    pts = K.generate_uniform_device(n).cpu().numpy()
for _ in range(3):
    with K.KnProblem(pts, k=k) as p:
        p.solve()
        t = p.timings()
print(name, k, {kk: (round(v, 4) if isinstance(v, float) else v) for kk, v in t.items()}, (p.info.dimx, p.info.dimy, p.info.dimz))
