"""Groups the SASS listing of an .ncu-rep into runs of equal execution count: a quick per-loop cost table.
usage: python profiles/segments.py gpurun_out/x.ncu-rep"""
import csv
import subprocess
import sys

src = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
h = rows[1]
data = rows[2:]
ix = {n: i for i, n in enumerate(h)}
tot = sum(int(r[ix["Instructions Executed"]]) for r in data)
segs, cur = [], None
for n, r in enumerate(data):
    ie = int(r[ix["Instructions Executed"]])
    th = float(r[ix["Avg. Threads Executed"]])
    sm = int(r[ix["# Samples"]])
    if ie == 0:
        continue
    if cur and abs(ie - cur["ie"]) <= 0.25 * max(ie, cur["ie"]):
        cur["n"] += 1; cur["sum"] += ie; cur["samp"] += sm; cur["thr"] += th * ie; cur["end"] = n
    else:
        if cur:
            segs.append(cur)
        cur = {"start": n, "end": n, "ie": ie, "n": 1, "sum": ie, "samp": sm, "thr": th * ie,
               "first": r[ix["Source"]].strip()[:50]}
segs.append(cur)
samp_tot = sum(s["samp"] for s in segs)
print(f"# total warp instructions {tot / 1e6:.1f} M, stall samples {samp_tot}")
print("# lines      n-instr  exec-each(M)  sum(M)   %instr  lanes  %samples  first instruction")
for s in segs:
    if s["sum"] > 0.004 * tot:
        print(f"{s['start']:4d}-{s['end']:4d} {s['n']:5d} {s['ie'] / 1e6:10.2f} {s['sum'] / 1e6:9.1f} {100 * s['sum'] / tot:6.1f} "
              f"{s['thr'] / s['sum']:6.1f} {100 * s['samp'] / samp_tot:7.1f}   {s['first']}")
