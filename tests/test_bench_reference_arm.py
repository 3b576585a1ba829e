"""bench.py --impl reference (the reference's CPU path, kd_tree.cpp unmodified, no GPU needed): the contract of its JSON
line on a tiny cloud, and that only rank 0 works under torchrun."""
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def run(extra_env=None, *args):
    env = dict(os.environ)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        env.pop(k, None)
    env.update(extra_env or {})
    return subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--points", "200000",
                           "--cpu-queries", "50000", "--steps", "1", "--warmup", "0", *args],
                          capture_output=True, text=True, env=env, timeout=300)


def test_reference_arm_line():
    r = run()
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line["impl"] == "reference"
    assert line["metric"] == "kNN queries/sec" and line["unit"] == "queries/s" and line["higher_is_better"] is True
    assert line["value"] > 0 and line["ms_per_step"] > 0
    assert line["config"]["n_points"] == 200000 and line["config"]["k"] == 20
    cb = line["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and cb["value"] == line["value"] and "kd-tree" in cb["sample"]
    e = line["e2e"]
    assert e["value"] == line["value"] and e["unit"] == line["unit"]
    assert e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0


def test_reference_arm_other_ranks_do_nothing():
    # torchrun exports OMP_NUM_THREADS=1 and starts every rank: rank 0 alone runs the arm (with an explicit thread count)
    r = run({"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2", "OMP_NUM_THREADS": "1"}, "--gpus", "2")
    assert r.returncode == 0, r.stderr[-2000:]
    assert not [l for l in r.stdout.splitlines() if l.startswith("{")]
