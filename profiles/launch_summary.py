"""Per-kernel launch counts, mean durations and shares from an `ncu --metrics gpu__time_duration.sum --csv` log.
usage: python profiles/launch_summary.py gpurun_out/launches.csv "<title line>" > profiles/rN_launches_summary.txt"""
import csv
import re
import sys
from collections import OrderedDict


def main(path, title):
    lines = [l for l in open(path) if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    acc = OrderedDict()
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"^void |<unnamed>::|\(.*$", "", r["Kernel Name"])
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        ms = v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1e-6)
        a = acc.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += ms
    total = sum(a[1] for a in acc.values())
    print(f"# {title}")
    print("# cold-cache, serialised: compare SHARES, not absolutes")
    for name, (cnt, ms) in acc.items():
        print(f"{name:55s} launches={cnt} avg_ms={ms / cnt:8.3f} share={100 * ms / total:5.1f}%")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "")
