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

# torchrun exports OMP_NUM_THREADS=1; the CPU legs (reference kd-tree, the oracle that verifies sampled rows) want
# the host cores, shared between the ranks of the node.  Must happen before any OpenMP runtime starts.
if "LOCAL_RANK" in os.environ and os.environ.get("OMP_NUM_THREADS", "1") == "1":
    os.environ["OMP_NUM_THREADS"] = str(max(1, (os.cpu_count() or 1) // max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1")))))

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

SEED = 20171004
METRIC = "kNN queries/sec"
UNIT = "queries/s"
KERNELS_PER_STEP = 13  # z-band split (3: histogram, starts, scatter), count, scan, scatter, order, order_big, query (tile, select for handed-over blocks, second-chance select, general retry, short-list retry)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", "--points", dest="n", type=int, default=100_000_000)  # --points: torchrun chokes on "--n"
    ap.add_argument("--k", type=int, default=20)
    ap.add_argument("--cpu-sample", type=int, default=16_000_000,
                    help="points of the bounded CPU baseline sample timed beside the GPU arm (cpu_baseline)")
    ap.add_argument("--cpu-queries", type=int, default=1_000_000,
                    help="--impl reference: sampled queries timed on the full tree (SURVEY 8d)")
    ap.add_argument("--no-verify", action="store_true", help="skip the oracle check of sampled rows after the timed region")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--density", type=float, default=0.0,
                    help="> 0: a cubic grid with this many points per cell (3.1 = the reference's grid, "
                         "knearests.cu:249); 0 = the library's default grid (knc_fast_dims), what kn_prepare uses")
    ap.add_argument("--algorithm", type=int, default=0,
                    help="knc_options.algorithm (0 auto, 1 general kernel, 2 insertion kernel, 4 without the tile kernel)")
    ap.add_argument("--no-ref-gpu", action="store_true", help="skip timing the reference's own knearests.cu on this GPU")
    return ap.parse_args()


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

    def reset(self):
        """Forget what was sampled so far: called where the timed region starts (the sampler is started before the
        warm-up steps, because the first NVML calls of a process take ~20 ms and delay the kernel launches next to them)."""
        self.samples, self.reasons = [], set()

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


# ------------------------------------------------------------------------------------------ CPU arm
def cpu_reference_run(n: int, k: int, n_queries: int, repeats: int = 1, warmup: int = 0):
    """The reference's CPU path (kd_tree.cpp unmodified, oracle/_ref), driven as test_knearests.cu:197-212 and
    measured as SURVEY 8d prescribes: the kd-tree is built over the first `n` points of the workload's generator
    (single-threaded, as in the reference), then `n_queries` of them (the first ones: the generator's order is
    random, so they are a uniform sample) are queried on all host threads; the query time is extrapolated linearly
    to n queries when n_queries < n.  queries/s = n / (build + extrapolated query) seconds."""
    import numpy as np
    import oracle as O
    kd = O.RefKdTree()
    threads = max(1, os.cpu_count() or 1)  # explicit: torchrun's OMP_NUM_THREADS=1 must not reach the kd-tree
    pts = O.generate_uniform(n, seed=SEED)
    n_queries = min(n_queries, n)
    # ONE tree build (minutes at 100 M points), then warm-up + timed query passes over the same tree
    _, info = kd.knn_passes(pts, k, 0, n_queries, threads, warmup + repeats)
    build = float(info["build_s"])
    query_sample = float(np.mean(info["query_s"][warmup:]))
    query = query_sample * (n / n_queries)
    how = (f"kd-tree over {n} uniform points (same generator, seed {SEED}) built in {build:.2f} s single-threaded; "
           f"{n_queries} sampled queries (k={k}) in {query_sample:.2f} s on {info['threads']} threads "
           f"(mean of {repeats} passes after {warmup} warm-up)"
           + (f", extrapolated x{n / n_queries:.0f} to all {n} queries = {query:.2f} s" if n_queries < n else ""))
    return {"value": n / (build + query), "unit": UNIT, "cores": info["threads"], "kind": "reference",
            "sample": how, "build_s": build, "query_s": query, "query_only_value": n / query}


def main_reference(args, rank: int):
    """--impl reference: kd_tree.cpp on the host cores, rank 0 only.  The FULL tree of the workload and
    1 M sampled queries; the tree build is ~100 s at 100 M points, so it is timed once, whatever --steps says."""
    if rank != 0:
        return
    t0 = time.perf_counter()
    cb = cpu_reference_run(args.n, args.k, args.cpu_queries, repeats=max(1, min(args.steps, 3)), warmup=1)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * (cb["build_s"] + cb["query_s"]),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": workload_config(args),
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "timed": "tree built once (~100 s at 100 M points); query passes over the sample: 1 warm-up + min(steps, 3) timed",
            "wall_s": time.perf_counter() - t0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------ verification
def verify_against_oracle(torch, rows, idtable, ptrs, info, index_offset: int, n_total: int, k: int,
                          samples: int = 64, seed: int = 0):
    """After the timed region: sampled rows of what the engine just computed, checked by the CPU oracle
    (brute force over the regenerated cloud -- the generator is counter-based, so any rank can rebuild any part).
      rows     device view [query_count, k]: neighbour rows (sorted positions + index_offset)
      idtable  device view [points held]: original id of every sorted position held (permutation / tags)
      ptrs     device view [cells held + 1]
    The sampled queries sit in four cell planes of the handle (first, last, two in between; random rows inside), so
    that only the slices of the cloud around those planes have to be kept: the cloud is regenerated chunk by chunk
    and a point is kept if its z lies within R = 3 first-radii of a sampled query; a row is only accepted if its k-th
    brute-force distance is below R (nothing outside the kept slices can then be closer).  Checks, per query:
    (1) its row, mapped to original ids, is the brute-force k nearest (the test of test_knearests.cu:194-232 with
    brute force in place of the kd-tree); (2) the query sits in the cell the oracle's cellFromPoint gives its
    coordinates, and that cell holds exactly the cloud's points of that cell, in ascending original id (the
    permutation contract, knearests.cu:57)."""
    import math
    import numpy as np
    import oracle as O
    nq = int(info.query_count)
    if nq == 0:
        return {"queries": 0, "mismatches": 0}
    rng = np.random.default_rng(1234 + seed)
    planes = max(1, int(info.z_end) - int(info.z_begin))
    per_plane = max(1, nq // planes)
    bases = sorted({0, nq // 3, (2 * nq) // 3, max(0, nq - per_plane)})
    each = max(1, samples // len(bases))
    sel = np.unique(np.concatenate([[0, nq - 1]] + [np.minimum(nq - 1, b + rng.integers(0, per_plane, each)) for b in bases]))
    sel = sel.astype(np.int64)
    sel_t = torch.from_numpy(sel).to(rows.device)
    rows_sel = rows[sel_t].to(torch.int64) & 0xFFFFFFFF
    nb = rows_sel - int(index_offset)
    got_ids = (idtable[nb.clamp(0, idtable.shape[0] - 1)].to(torch.int64) & 0xFFFFFFFF).cpu().numpy()
    empty = (rows_sel == 0xFFFFFFFF).cpu().numpy()
    spos = sel + int(info.query_begin)                      # local sorted positions of the sampled queries
    qid = (idtable[torch.from_numpy(spos).to(rows.device)].to(torch.int64) & 0xFFFFFFFF).cpu().numpy()
    qxyz = np.stack([O.generate_uniform(1, seed=SEED, first=int(i))[0] for i in qid])
    r0 = (3.0 * (k + 2.5 * math.sqrt(k) + 2.0) / (4.0 * math.pi * (n_total / 1e9))) ** (1.0 / 3.0)
    R = np.float32(3.0 * r0)
    # the kept slices: z within R of a sampled query (merged intervals, looked up with searchsorted)
    iv = sorted((float(z - R), float(z + R)) for z in qxyz[:, 2])
    merged = [list(iv[0])]
    for a_, b_ in iv[1:]:
        if a_ <= merged[-1][1]:
            merged[-1][1] = max(merged[-1][1], b_)
        else:
            merged.append([a_, b_])
    los = np.array([m[0] for m in merged], np.float32)
    his = np.array([m[1] for m in merged], np.float32)
    keep_ids, keep_xyz = [], []
    chunk = 20_000_000
    for start in range(0, n_total, chunk):
        pts = O.generate_uniform(min(chunk, n_total - start), seed=SEED, first=start)
        z = pts[:, 2]
        j = np.searchsorted(los, z, side="right") - 1
        inside = (j >= 0) & (z <= his[np.maximum(j, 0)])
        idx = np.nonzero(inside)[0]
        keep_ids.append(idx.astype(np.int64) + start)
        keep_xyz.append(pts[idx])
    ids_sub = np.concatenate(keep_ids)
    cloud = np.ascontiguousarray(np.concatenate(keep_xyz))
    kk = min(k + 1, len(ids_sub))
    want_loc, want_d2 = O.brute_external(cloud, qxyz, kk, with_d2=True)
    mism, reordered, too_far = 0, 0, 0
    for r in range(len(sel)):
        w = [int(ids_sub[v]) for v in want_loc[r] if int(ids_sub[v]) != int(qid[r])][:k]  # the query itself is dropped (knearests.cu:123)
        g = [int(v) for v, e in zip(got_ids[r], empty[r]) if not e]
        if len(w) == k and not (want_d2[r][kk - 1] < R * R):
            too_far += 1     # the slices could hide a closer point: this row proves nothing (not expected on uniform clouds)
        if g == w:
            continue
        if sorted(g) == sorted(w):
            reordered += 1   # same set: only the order inside a run of equal d2 may differ (tie-break by sorted position)
        else:
            mism += 1
    # cell membership of the sampled queries
    dimx, dimy, dimz, zb = int(info.dimx), int(info.dimy), int(info.dimz), int(info.z_begin)
    L = O.lib()
    cells = np.array([L.kno_cell_from_point(dimx, dimy, dimz, float(x), float(y), float(z)) for x, y, z in qxyz], np.int64)
    local = cells - zb * dimx * dimy
    lo = ptrs[torch.from_numpy(local).to(rows.device)].cpu().numpy().astype(np.int64)
    hi = ptrs[torch.from_numpy(local + 1).to(rows.device)].cpu().numpy().astype(np.int64)
    cell_bad = int(((spos < lo) | (spos >= hi)).sum())
    # every kept point's cell, by the oracle's cell function (vectorised, f32 like knearests.cu:23-28)
    def axis(v, d):
        return np.clip(np.floor((v * np.float32(d)) / np.float32(1000.0)), 0, d - 1).astype(np.int64)
    cid = axis(cloud[:, 0], dimx) + dimx * (axis(cloud[:, 1], dimy) + dimy * axis(cloud[:, 2], dimz))
    for c, a_, b_ in zip(cells, lo, hi):
        members = (idtable[int(a_):int(b_)].to(torch.int64) & 0xFFFFFFFF).cpu().numpy()
        if not np.array_equal(members, ids_sub[cid == c]):   # ascending original id = stable counting sort
            cell_bad += 1
    return {"queries": int(len(sel)), "mismatches": int(mism + cell_bad + too_far), "rows_wrong": int(mism),
            "cells_wrong": int(cell_bad), "outside_kept_slices": int(too_far), "tie_reordered": int(reordered),
            "points_kept": int(len(ids_sub)),
            "checker": "oracle brute force (kno_brute_external) over the regenerated slices of the cloud around the "
                       "sampled queries + cell membership"}


def workload_config(args):
    c = {"workload": f"synthetic uniform {args.n} points in [0,1000)^3 (splitmix64 counter generator, seed {SEED}), "
                     f"k={args.k}: kn_prepare + kn_solve", "n_points": args.n, "k": args.k,
         "grid": (f"cube, {args.density} points per cell" if args.density > 0 else
                  "library default (knc_fast_dims: y/z cells as wide as the first search radius, ~1 point per cell)"),
         "l2_policy": "inputs larger than L2 (points 12 B x n, rows 4k B x n >> 126 MB); no flush needed",
         "parallelism": "single GPU" if args.gpus <= 1 else f"z-slabs x{args.gpus}"}
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
                                 ClockSampler, measured_peak_gbs,
                                 verify=None if args.no_verify else verify_against_oracle)

    pts = K.generate_uniform_device(n, seed=SEED, device=dev)
    torch.cuda.synchronize()

    def step_device():
        p = K.KnProblem(pts, k=k, density=args.density, algorithm=args.algorithm)
        p.solve()
        t = p.timings()
        p.free()
        return t

    import gc
    timings = []
    gc.collect()
    gc.disable()  # no collection inside the timed region (nor in the warm-up: the first timed step should find warm caches)
    with ClockSampler(local_rank) as clk:
        for _ in range(args.warmup):
            step_device()
        torch.cuda.synchronize()
        clk.reset()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            timings.append(step_device())
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
    gc.enable()
    ms_per_step = 1e3 * wall / args.steps
    value = n / (wall / args.steps)
    solve_ms = float(np.mean([t["solve_ms"] for t in timings]))
    build_ms = float(np.mean([t["build_ms"] for t in timings]))
    grid_dims = list(K.fast_dims(n, k)) if args.density <= 0 else [K.grid_dim(n, args.density)] * 3

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
                "traffic": traffic, "kernel": "kn_query_tile_kernel (+ select kernel for handed-over blocks and as second chance of the retry list + retry kernels)", "kernel_ms": solve_ms,
                "algorithmic_bytes": algo_bytes, "peak_kind": f"{peak_kind} copy bandwidth (burst)"}

    # ---- after the timed region: sampled rows of one more (untimed) pass, checked by the CPU oracle ------------
    verify = None
    if not args.no_verify:
        from cuda_knearests_b200.slabs import _DevView
        p = K.KnProblem(pts, k=k, density=args.density, algorithm=args.algorithm)
        p.solve()
        i = p.info
        cells = i.dimx * i.dimy * (i.z_end - i.z_begin)
        verify = verify_against_oracle(
            torch, torch.as_tensor(_DevView(i.d_knearests, (i.query_count, k)), device=dev),
            torch.as_tensor(_DevView(i.d_permutation, (i.n,)), device=dev),
            torch.as_tensor(_DevView(i.d_ptrs, (cells + 1,)), device=dev), i, 0, n, k)
        p.free()

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
        ns = min(args.cpu_sample, n)  # bounded (~15 s): a 16 M-point sub-cloud, every point queried
        cb = cpu_reference_run(ns, k, ns, repeats=1, warmup=0)
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
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args), "grid_dims": grid_dims,
            "clocks": clk.summary(), "e2e": e2e, "gpu_launches": KERNELS_PER_STEP * args.steps,
            "roofline": roofline, "cpu_baseline": cpu_baseline,
            "device_ms": {"grid_build": build_ms, "query": solve_ms},
            "solve_only_value": n / (solve_ms * 1e-3), "reference_gpu": reference_gpu, "verify": verify,
            "restarts": int(timings[-1]["restarts"]), "handed_over_blocks": int(timings[-1]["handed_over_blocks"])}
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
