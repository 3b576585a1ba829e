#!/usr/bin/env python
"""bench.py -- kNN queries/s of the B200 engine on the reference's headline workload.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--n POINTS] [--k K]

Workload (BASELINE.json configs[3], the one the metric's target is quoted on; fits one GPU):
synthetic uniform cloud of 100 M points in [0,1000)^3 (counter-based generator of SURVEY 8d,
seed 20171004), k = 20.  One STEP = one full pass of the hot path over the cloud:
kn_prepare (grid build) + kn_solve (all-points kNN query), through the C ABI.

  value   queries/s with the input already resident in HBM (device pointer in, rows left on the
          device): N / wall time per step, wall time bracketed by barrier + synchronize.
  e2e     same step through the reference-facing call with HOST buffers: pinned host points in
          (H2D inside), kn_solve, neighbour rows + permutation copied back to pinned host memory.
  roofline  the query kernel (the dominant kernel): algorithmic bytes (12 + 4k per query)
          over its CUDA-event duration, against the measured HBM copy peak of MEASURED_PEAKS.json.
  cpu_baseline  the reference's kd_tree.cpp (oracle/_ref, unmodified) on the host cores, on a bounded
          sample of the same workload.

--impl reference times the reference's own CPU implementation of the path (kd_tree.cpp driven as
test_knearests.cu:197-212, all host threads) on a bounded sample of the same workload.

N > 1 (torchrun, one rank per GPU): the cloud is cut into z-slabs of cell planes; every rank builds
and queries its slab + halo; strong scaling (total work fixed).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

SEED = 20171004
METRIC = "kNN queries/sec"
UNIT = "queries/s"
KERNELS_PER_STEP = 11  # z-band split (3), count, scan, scatter, order, order_big, gather, query (select), query (retry)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", "--points", dest="n", type=int, default=100_000_000)  # --points: torchrun chokes on "--n"
    ap.add_argument("--k", type=int, default=20)
    ap.add_argument("--cpu-sample", type=int, default=16_000_000, help="points of the CPU baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--density", type=float, default=0.0,
                    help="points per grid cell handed to knc_prepare_ex (library default: the reference's 3.1); "
                         "0 = cuda_knearests_b200.fast_density(k): 1.0 at k=20, ~14 %% faster on B200 than 3.1, "
                         "profiles/r1_notes.md")
    ap.add_argument("--algorithm", type=int, default=0, help="knc_options.algorithm (0 auto, 1 general, 2 insertion kernel)")
    ap.add_argument("--no-ref-gpu", action="store_true", help="skip timing the reference's own knearests.cu on this GPU")
    args = ap.parse_args()
    if args.density <= 0:
        from cuda_knearests_b200 import fast_density
        args.density = fast_density(args.k)
    return args


def measured_peak_gbs():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"  # B200_PROFILING.md fallback


class ClockSampler:
    """Samples SM clock and throttle reasons DURING the timed region (NVML, 20 ms period)."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap", 0x80: "hw_power_brake_slowdown"}

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(int(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                mask = int(self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.02)

    def __enter__(self):
        if self.nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t:
            self._t.join()

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


# ------------------------------------------------------------------------------------------ CPU arm
def cpu_reference_run(n_sample: int, k: int, steps: int = 1, warmup: int = 0):
    """The reference's CPU path (kd_tree.cpp unmodified, oracle/_ref) on a bounded sample: the first
    `n_sample` points of the workload's generator, rescaled to fill [0,1000)^3 at the same density
    law (uniform), all host threads.  queries/s = sample queries / (tree build + query) seconds."""
    import numpy as np
    import oracle as O
    kd = O.RefKdTree()
    kind = "reference"
    pts = O.generate_uniform(n_sample, seed=SEED)
    runs = []
    for it in range(warmup + steps):
        _, info = kd.knn(pts, k)
        if it >= warmup:
            runs.append(info)
    build = float(np.mean([r["build_s"] for r in runs]))
    query = float(np.mean([r["query_s"] for r in runs]))
    return {"value": n_sample / (build + query), "unit": UNIT, "cores": runs[0]["threads"], "kind": kind,
            "sample": f"{n_sample} uniform points (same generator, seed {SEED}), k={k}: kd-tree build "
                      f"{build:.2f} s single-threaded + {n_sample} queries {query:.2f} s on {runs[0]['threads']} threads",
            "build_s": build, "query_s": query, "query_only_value": n_sample / query}


def main_reference(args, rank: int):
    if rank != 0:
        return
    n_sample = min(args.cpu_sample, args.n)
    t0 = time.perf_counter()
    cb = cpu_reference_run(n_sample, args.k, steps=max(1, args.steps), warmup=min(args.warmup, 1))
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": min(args.warmup, 1), "ms_per_step": 1e3 * (cb["build_s"] + cb["query_s"]),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": workload_config(args, n_sample_note=n_sample),
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "wall_s": time.perf_counter() - t0}
    print(json.dumps(line), flush=True)


def workload_config(args, n_sample_note=None):
    c = {"workload": f"synthetic uniform {args.n} points in [0,1000)^3 (splitmix64 counter generator, seed {SEED}), "
                     f"k={args.k}: kn_prepare + kn_solve", "n_points": args.n, "k": args.k,
         "grid_density": args.density,
         "l2_policy": "inputs larger than L2 (points 12 B x n, rows 4k B x n >> 126 MB); no flush needed"}
    if n_sample_note:
        c["cpu_sample_points"] = n_sample_note
    return c


# ------------------------------------------------------------------------------------------ GPU arm
def main_ours(args, rank: int, world: int, local_rank: int):
    import numpy as np
    import torch
    import torch.distributed as dist

    import cuda_knearests_b200 as K
    K.lib()  # raises if the CUDA library is missing: no fallback
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    n, k = args.n, args.k

    if world > 1:
        from cuda_knearests_b200 import slabs
        return slabs.bench_multi(args, rank, world, local_rank, METRIC, UNIT, SEED, workload_config,
                                 ClockSampler, measured_peak_gbs)

    pts = K.generate_uniform_device(n, seed=SEED, device=dev)
    torch.cuda.synchronize()

    def step_device():
        p = K.KnProblem(pts, k=k, density=args.density, algorithm=args.algorithm)
        p.solve()
        t = p.timings()
        p.free()
        return t

    for _ in range(args.warmup):
        step_device()
    torch.cuda.synchronize()
    timings = []
    with ClockSampler(local_rank) as clk:
        t0 = time.perf_counter()
        for _ in range(args.steps):
            timings.append(step_device())
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
    ms_per_step = 1e3 * wall / args.steps
    value = n / (wall / args.steps)
    solve_ms = float(np.mean([t["solve_ms"] for t in timings]))
    build_ms = float(np.mean([t["build_ms"] for t in timings]))

    peak, peak_kind = measured_peak_gbs()
    algo_bytes = (12 + 4 * k) * n
    achieved = algo_bytes / (solve_ms * 1e-3) / 1e9
    traffic = None
    tf = ROOT / "profiles" / "roofline_traffic.json"
    if tf.exists():
        try:
            traffic = json.loads(tf.read_text()).get(f"n{n}_k{k}")
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "kernel": "kn_query_select_kernel (+ retry kernel)", "kernel_ms": solve_ms,
                "algorithmic_bytes": algo_bytes, "peak_kind": f"{peak_kind} copy bandwidth (burst)"}

    # ---- end to end: host buffers in, host buffers out, through the reference-facing call ----------
    e2e = None
    if not args.no_e2e:
        host_in = torch.empty((n, 3), dtype=torch.float32, pin_memory=True)
        host_in.copy_(pts)
        torch.cuda.synchronize()
        host_in_np = host_in.numpy()
        host_rows = torch.empty((n, k), dtype=torch.int32, pin_memory=True)
        host_perm = torch.empty((n,), dtype=torch.int32, pin_memory=True)
        rows_np = host_rows.numpy().view(np.uint32)
        perm_np = host_perm.numpy().view(np.uint32)
        import ctypes as C

        def step_e2e():
            p = K.KnProblem(host_in_np, k=k, density=args.density, algorithm=args.algorithm)   # kn_prepare: H2D + grid build
            p.solve_to_host(rows_np, chunks=8)                        # kn_solve + kn_get_knearests, overlapped
            K._check(K.lib().knc_get_permutation(p._h, perm_np.ctypes.data))    # kn_get_permutation
            p.free()

        e2e_steps = max(2, min(args.steps, 3))
        step_e2e()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            step_e2e()
        torch.cuda.synchronize()
        e2e_wall = (time.perf_counter() - t0) / e2e_steps
        e2e = {"value": n / e2e_wall, "unit": UNIT, "h2d_bytes_per_step": 12 * n,
               "d2h_bytes_per_step": 4 * k * n + 4 * n, "ms_per_step": 1e3 * e2e_wall, "steps": e2e_steps,
               "call": "knc_prepare(host xyz) + knc_solve_to_host(rows, 8 chunks) + knc_get_permutation, pinned host buffers"}
        del host_in, host_rows, host_perm

    cpu_baseline = None
    if not args.no_cpu_baseline:
        cb = cpu_reference_run(min(args.cpu_sample, n), k)
        cpu_baseline = {kk: cb[kk] for kk in ("value", "unit", "cores", "kind", "sample")}

    # ---- the reference's OWN GPU kernel (knearests.cu, unmodified, compiled for sm_100a: oracle/_ref) on
    # the same cloud and the same GPU, timed around its blocking kn_prepare / kn_solve (informational)
    reference_gpu = None
    if not args.no_ref_gpu:
        try:
            import oracle as O
            if O.have_ref(f"libref_knearests_k{k}.so"):
                host_pts = pts.cpu().numpy()
                del pts
                torch.cuda.empty_cache()
                ref = O.RefGpu(k)
                ref.run(host_pts[: min(n, 1_000_000)], fetch=False)  # context / module warm-up
                r = ref.run(host_pts, fetch=False)
                reference_gpu = {"kernel": "knearests.cu (reference, sm_100a build)", "n_points": n, "k": k,
                                 "kn_solve_ms": 1e3 * r["solve_s"], "kn_prepare_ms_incl_h2d": 1e3 * r["prepare_s"],
                                 "solve_only_value": n / r["solve_s"], "unit": UNIT,
                                 "ours_over_reference_solve": (1e3 * r["solve_s"]) / solve_ms}
        except Exception as e:  # informational only
            reference_gpu = {"unavailable": str(e)[:200]}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args),
            "clocks": clk.summary(), "e2e": e2e, "gpu_launches": KERNELS_PER_STEP * args.steps,
            "roofline": roofline, "cpu_baseline": cpu_baseline,
            "device_ms": {"grid_build": build_ms, "query": solve_ms},
            "solve_only_value": n / (solve_ms * 1e-3), "reference_gpu": reference_gpu,
            "restarts": int(timings[-1]["restarts"])}
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return main_reference(args, rank)
    main_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
