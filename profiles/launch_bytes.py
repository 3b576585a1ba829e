"""Per launch: duration and DRAM bytes from an `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv` log.
usage: python profiles/launch_bytes.py gpurun_out/launches.csv [first N launches]"""
import csv
import re
import sys
from collections import OrderedDict

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
acc = OrderedDict()
for r in csv.DictReader(lines):
    name = re.sub(r"^void |<unnamed>::|\(.*$", "", r["Kernel Name"])
    acc.setdefault((int(r["ID"]), name), {})[r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
limit = int(sys.argv[2]) if len(sys.argv) > 2 else len(acc)
print(f"{'kernel':50s} {'ms':>8s} {'read GB':>8s} {'write GB':>8s} {'TB/s':>6s}")
for (i, name), v in list(acc.items())[:limit]:
    ms = v["gpu__time_duration.sum"] * 1e-6
    rd, wr = v.get("dram__bytes_read.sum", 0) / 1e9, v.get("dram__bytes_write.sum", 0) / 1e9
    print(f"{name[:50]:50s} {ms:8.3f} {rd:8.2f} {wr:8.2f} {(rd + wr) / ms:6.2f}")
