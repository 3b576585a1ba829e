"""Times the reference's own knearests.cu (oracle/_ref, sm_100a build) and this engine side by side on the
BASELINE configs that fit one GPU; prints a markdown table.   python profiles/ref_gpu_table.py"""
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import cuda_knearests_b200 as K  # noqa: E402
import oracle as O  # noqa: E402
from test_parity_reference_gpu import blue_noise_substitute  # noqa: E402

cases = [("pts20K.xyz (reference fixture)", np.load(ROOT / "tests/golden/pts20k_norm.npy"), 50),
         ("300 k uniform (substitute for pts300K.xyz)", O.generate_uniform(300_000), 50),
         ("300 k jittered grid (substitute for 300k_blue_cube.xyz)", blue_noise_substitute(67, 3), 50),
         ("900 k jittered grid (substitute for 900k_blue_cube.xyz)", blue_noise_substitute(97, 4), 50),
         ("10 M uniform", O.generate_uniform(10_000_000), 50),
         ("10 M uniform", O.generate_uniform(10_000_000), 20)]
print("| cloud | k | reference kn_solve ms | engine kn_solve ms | ratio | engine build ms |")
print("|---|---|---|---|---|---|")
for name, pts, k in cases:
    ref = O.RefGpu(k)
    ref.run(pts[: min(len(pts), 100_000)], fetch=False)
    r = min((ref.run(pts, fetch=False)["solve_s"] for _ in range(3))) * 1e3
    best = None
    for _ in range(4):
        with K.KnProblem(pts, k=k) as p:
            p.solve()
            t = p.timings()
        best = t if best is None or t["solve_ms"] < best["solve_ms"] else best
    print(f"| {name} ({len(pts)} pts) | {k} | {r:.2f} | {best['solve_ms']:.3f} | {r / best['solve_ms']:.1f}x | {best['build_ms']:.3f} |")
