"""Regenerates the golden vectors in this directory FROM THE REFERENCE ITSELF.

Run in the build container only (needs /root/reference and `make -C oracle`):

    python tests/golden/make_golden.py

What is recorded (all from the reference's only fixture, pts20K.xyz, N = 20 626):
  pts20k_norm.npy            f32 [N,3]  the fixture after the reference's OWN loader+normaliser
                                        (test_knearests.cu:40-80 via oracle/_ref/libref_loader.so)
  pts20k_kdtree_k{8,20,50}_rowhash.npy  u64 [N]  order-independent hash of every neighbour SET the
                                        reference kd-tree returns (kd_tree.cpp via test_knearests.cu:197-212)
  pts20k_kdtree_k50_rows512.npy  u16 [512,50]  full rows (nearest first) of the first 512 queries
  pts20k_facts.json          D, per-cell occupancy histogram (what kn_print_stats would print)
The GPU box has no /root/reference: tests read only these files there.
"""
import json
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parent.parent))
import oracle as O  # noqa: E402


def set_hash(rows: np.ndarray) -> np.ndarray:
    """Order-independent 64-bit hash of each row's index set: sum of splitmix64(idx) mod 2^64."""
    x = rows.astype(np.uint64) + np.uint64(0x9E3779B97F4A7C15)
    with np.errstate(over="ignore"):
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        x = x ^ (x >> np.uint64(31))
        return x.sum(axis=1, dtype=np.uint64)


def main():
    O.build(ref=True)
    src = "/root/reference/pts20K.xyz"
    pts = O.ref_load_xyz(src)
    assert pts.shape == (20626, 3)
    np.save(HERE / "pts20k_norm.npy", pts)
    kd = O.RefKdTree()
    for K in (8, 20, 50):
        rows, _ = kd.knn(pts, K)
        np.save(HERE / f"pts20k_kdtree_k{K}_rowhash.npy", set_hash(rows))
        if K == 50:
            np.save(HERE / "pts20k_kdtree_k50_rows512.npy", rows[:512].astype(np.uint16))
    g = O.Grid(pts)
    facts = {"n": int(pts.shape[0]), "dim": int(g.dim),
             "occupancy_histogram": {str(i): int(c) for i, c in enumerate(np.bincount(g.counters))},
             "bbox_min": float(pts.min()), "bbox_max": float(pts.max()),
             "source": "ssloy/cuda_knearests pts20K.xyz through the reference loader"}
    (HERE / "pts20k_facts.json").write_text(json.dumps(facts, indent=1) + "\n")
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()
