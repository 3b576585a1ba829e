"""Turns an .ncu-rep brought back in gpurun_out/ into the small text summaries kept in profiles/.
usage: python profiles/summarize.py gpurun_out/query_r1.ncu-rep profiles/r1_query_v1"""
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__inst_executed_pipe_lsu.sum",
        "smsp__inst_executed_op_shared_ld.sum", "smsp__inst_executed_op_shared_st.sum"]


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    head, unit = rows[0], rows[1]
    with open(out + "_metrics.txt", "w") as f:
        for r in rows[2:]:
            f.write(f"# kernel: {r[head.index('Kernel Name')]}\n")
            for i, n in enumerate(head):
                if n in KEYS or ("warp_issue_stalled" in n and n.endswith("per_warp_active.pct")):
                    f.write(f"{n} = {r[i]} {unit[i]}\n")
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(src.splitlines()))
    h = rows[1]
    ix = {n: i for i, n in enumerate(h)}
    with open(out + "_sass_hot.txt", "w") as f:
        tot = sum(int(r[ix["Instructions Executed"]]) for r in rows[2:])
        thr = sum(int(r[ix["Thread Instructions Executed"]]) for r in rows[2:])
        f.write(f"# warp instructions {tot}, thread instructions {thr}, active lanes/instruction {thr / max(tot, 1):.2f}\n")
        f.write("# line  warp-inst(M)  avg-active-lanes  stall-samples  SASS\n")
        hot = max(int(r[ix["Instructions Executed"]]) for r in rows[2:])
        wv = ix.get("L1 Wavefronts Shared")
        f.write("# the hot loop = every instruction executed at least half as often as the most executed one\n")
        for n, r in enumerate(rows[2:]):
            ie = int(r[ix["Instructions Executed"]])
            if 2 * ie < hot:
                continue
            smem = f" smem-wavefronts/inst {float(r[wv] or 0) / max(ie, 1):4.1f}" if wv is not None and float(r[wv] or 0) > 0 else ""
            f.write(f"{n:5d} {ie / 1e6:10.1f} {float(r[ix['Avg. Threads Executed']]):6.1f} {int(r[ix['# Samples']]):8d}  "
                    f"{r[ix['Source']].strip()}{smem}\n")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
