"""slabs -- z-slab sharding of one kNN problem over the GPUs of a node, one process per GPU.

No reference counterpart (the reference is single-GPU); this is SURVEY 8e.  Every rank
holds an index-contiguous chunk of the cloud.  One GLOBAL grid is sized from the total point count
(knearests.cu:249); rank r owns the cell planes [zsplit[r], zsplit[r+1]) and needs `halo` more planes
on either side as read-only candidates.  One exchange step moves the points:

  knc_slab_route (CUDA, stable multi-split by destination, 16-byte records)  ->  ONE all-to-all (NCCL
  over NVLink; gloo in the CPU tests)  ->  knc_prepare_ex(slab, point_stride 4) + knc_solve per rank
  ->  rows in global sorted positions + the ids in sorted order (knc_get_tags) = the global permutation slice

Because the route is stable and ranks hold rising index ranges, a rank receives its slab in global index
order, the engine's stable cell sort reproduces the single-GPU order, and the rows are bit-identical to
the single-GPU rows (tests/test_multi_gpu.py).  The engine itself checks that the halo was wide
enough (KNC_ERR_HALO); if any rank reports a miss, all ranks widen the halo and repeat.

PyTorch is used for buffers and torch.distributed only.
"""
from __future__ import annotations

import ctypes as C
import math
import os
import sys
import time

import numpy as np


def grid_and_split(n_total: int, world: int, k: int = 0, density: float = 0.0):
    """Global grid resolution (x, y, z) and plane ownership: zsplit[r] = r * Dz // world.  density > 0: the cube
    with that many points per cell (3.1: the reference's, knearests.cu:249); else the library's own grid for k."""
    import cuda_knearests_b200 as K
    if density > 0:
        D = K.grid_dim(n_total, density)
        dims = (D, D, D)
    else:
        dx, dy, dz = K.fast_dims(n_total, k)
        if dz >= 4 * world and dz % world:
            # equal numbers of planes per rank: the slowest rank sets the pace, and 29 planes against 28 are 3.6 % more
            # queries.  Fewer (thicker) planes keep the cells wider than the first search radius, as the tile kernel wants.
            dz -= dz % world
            dx = max(1, round(n_total / (dy * dz)))
        dims = (dx, dy, dz)
    return dims, [(r * dims[2]) // world for r in range(world + 1)]


def halo_planes(k: int, n_total: int | None = None, planes: int | None = None, density: float = 3.1) -> int:
    """Planes of halo on each side: 1.75 x the first-guess search radius, in planes.  With n_total and the number
    of z planes of the grid: for that grid (2 on the library's own grid, whose planes are as thick as the radius);
    without: for the cube of `density` points per cell (k = 8 / 20 / 32 / 50 -> 2 / 3 / 3 / 4 at
    3.1).  Verified after the solve, widened if it was not enough."""
    lam = k + 2.5 * math.sqrt(k) + 2.0
    if n_total is not None and planes is not None:
        r0 = (3.0 * lam / (4.0 * math.pi * (n_total / 1e9))) ** (1.0 / 3.0)
        return max(2, math.ceil(1.75 * r0 / (1000.0 / planes)))
    r0_cells = (3.0 * lam / (4.0 * math.pi * density)) ** (1.0 / 3.0)
    return max(2, math.ceil(1.75 * r0_cells))


def held_planes(zsplit, rank: int, halo: int, D: int):
    """(z_begin, z_end, own_z_begin, own_z_end) of a rank: what knc_options wants."""
    lo, hi = zsplit[rank], zsplit[rank + 1]
    return max(0, lo - halo), min(D, hi + halo), lo, hi


def exchange(send_rec, send_counts, group=None):
    """The collective of the path: variable-size all-to-all of 16-byte (x, y, z, id) records grouped by
    destination -- one buffer, one all-to-all (plus the exchange of the group sizes).  Works on CUDA tensors
    (NCCL) and CPU tensors (gloo).  Returns (recv_rec [m,4], recv_counts): the groups of ranks 0..W-1
    concatenated in rank order."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    dev = send_rec.device
    sc = torch.tensor(list(send_counts), dtype=torch.int64, device=dev)
    rc = torch.empty(world, dtype=torch.int64, device=dev)
    dist.all_to_all_single(rc, sc, group=group)
    recv_counts = [int(v) for v in rc.tolist()]
    recv_rec = torch.empty((sum(recv_counts), 4), dtype=send_rec.dtype, device=dev)
    dist.all_to_all_single(recv_rec, send_rec[:sum(send_counts)], recv_counts, list(send_counts), group=group)
    return recv_rec, recv_counts


def route_device(local_xyz, first_id: int, D: int, zsplit, halo: int):
    """knc_slab_route on the current device.  Returns (send_rec [cap,4] float32 records, counts)."""
    import torch
    import cuda_knearests_b200 as K
    L = K.lib()
    n = local_xyz.shape[0]
    world = len(zsplit) - 1
    cap = int(n * 1.5) + 4096
    zs = (C.c_int * (world + 1))(*zsplit)
    counts = (C.c_longlong * world)()
    while True:
        out = torch.empty((cap, 4), dtype=torch.float32, device=local_xyz.device)
        torch.cuda.current_stream().synchronize()
        rc = L.knc_slab_route(C.c_void_p(local_xyz.data_ptr()), n, first_id, D, world, zs, halo,
                              C.c_void_p(out.data_ptr()), cap, counts)
        if rc == 3 and sum(counts) > cap:  # KNC_ERR_NOMEM: groups larger than guessed
            cap = int(sum(counts)) + 4096
            continue
        K._check(rc)
        return out, [int(c) for c in counts]


def record_ids(rec):
    """The id words of a record buffer [m,4] as an int32 view [m] with stride 4 (no copy)."""
    import torch
    return rec.view(torch.int32)[:, 3]


class PeerExchange:
    """Receive buffers of all ranks of the node, mapped into every process (CUDA IPC), so that the routing kernel
    writes each group straight into its destination GPU's memory over NVLink: routing + exchange in ONE kernel, no
    send buffer and no NCCL all-to-all (the only collectives left are an all-gather of W group sizes and a barrier).
    Created once per process group and reused by every step."""

    def __init__(self, capacity_records: int, group=None):
        """Collective.  Raises KnError on EVERY rank if any rank could not allocate, export or map a buffer
        (the ranks vote), so that callers can switch to the staged exchange together."""
        import torch
        import torch.distributed as dist
        import cuda_knearests_b200 as K
        self.K, self.group = K, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.capacity = int(capacity_records)
        self.bufs, self.local = [], None
        L = K.lib()
        mine = C.c_void_p()
        handle = C.create_string_buffer(64)
        err = None
        if L.knc_peer_alloc(self.capacity * 16, C.byref(mine), handle) == 0:
            self.local = mine.value
        else:
            err = L.knc_last_error().decode()
        handles = [None] * self.world
        dist.all_gather_object(handles, handle.raw if self.local else None, group=group)
        if all(h is not None for h in handles):
            for r, h in enumerate(handles):
                if r == self.rank:
                    self.bufs.append(self.local)
                    continue
                p = C.c_void_p()
                if L.knc_peer_open(h, C.byref(p)) != 0:
                    err = err or L.knc_last_error().decode()
                    break
                self.bufs.append(p.value)
        else:
            err = err or "another rank could not allocate its receive buffer"
        dev = torch.device("cuda", torch.cuda.current_device())
        flag = torch.tensor([0 if err else 1], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        if int(flag.item()) == 0:
            self.close()
            raise K.KnError(2, f"peer exchange unavailable: {err or 'another rank failed'}")
        # the whole step (routing kernels, collectives, grid build, query) is enqueued on ONE stream of this rank
        self.stream = torch.cuda.Stream(device=dev)
        self.counts3 = torch.zeros(3 * self.world, dtype=torch.int64, device=dev)
        self.allc = torch.zeros((self.world, 3 * self.world), dtype=torch.int64, device=dev)
        self.allc_host = torch.zeros((self.world, 3 * self.world), dtype=torch.int64).pin_memory()
        self.token = torch.zeros(1, dtype=torch.int32, device=dev)

    def route_async(self, local_xyz, first_id: int, D: int, zsplit, halo: int):
        """Route this rank's chunk into everybody's receive buffer with ONE host synchronisation: the count kernels,
        an all-gather of (records, own, low-halo) counts per destination, one read-back of that small matrix, the
        scatter kernel and a stream-ordered all-reduce that stands for the barrier -- all on self.stream.
        Returns (records this rank receives, query_begin, query_count, global position of its first query, sent)."""
        import torch
        import torch.distributed as dist
        K, L = self.K, self.K.lib()
        W = self.world
        zs = (C.c_int * (W + 1))(*zsplit)
        h = C.c_void_p()
        st = C.c_void_p(self.stream.cuda_stream)
        K._check(L.knc_route_begin_async(C.c_void_p(local_xyz.data_ptr()), local_xyz.shape[0], first_id, D, W, zs, halo, st,
                                         C.c_void_p(self.counts3.data_ptr()), C.byref(h)))
        try:
            dist.all_gather_into_tensor(self.allc, self.counts3, group=self.group)
            self.allc_host.copy_(self.allc, non_blocking=True)
            self.stream.synchronize()                       # the one host wait of the exchange
            a = self.allc_host.numpy()
            sent, own, low = a[:, :W], a[:, W:2 * W], a[:, 2 * W:]
            incoming = sent.sum(axis=0)
            if int(incoming.max()) > self.capacity:
                raise K.KnError(3, f"peer receive buffer holds {self.capacity} records, {int(incoming.max())} are on their way")
            first = (C.c_longlong * W)(*[int(sent[:self.rank, d].sum()) for d in range(W)])
            bufs = (C.c_void_p * W)(*self.bufs)
            K._check(L.knc_route_scatter_async(h, bufs, first, st))
        finally:
            L.knc_route_end(h)
        dist.all_reduce(self.token, group=self.group)       # in stream order: every rank's stores have landed after it
        qcount = own.sum(axis=0)
        r = self.rank
        return (int(incoming[r]), int(low[:, r].sum()), int(qcount[r]), int(qcount[:r].sum()), int(sent[r].sum()))

    def route(self, local_xyz, first_id: int, D: int, zsplit, halo: int):
        """Route this rank's chunk into everybody's receive buffer.  Returns (records_received, torch view [m,4]
        of the local receive buffer, records_sent) or raises if a receive buffer would overflow."""
        import torch
        import torch.distributed as dist
        K, L = self.K, self.K.lib()
        W = self.world
        dev = local_xyz.device
        zs = (C.c_int * (W + 1))(*zsplit)
        counts = (C.c_longlong * W)()
        h = C.c_void_p()
        K._check(L.knc_route_begin(C.c_void_p(local_xyz.data_ptr()), local_xyz.shape[0], first_id, D, W, zs, halo,
                                   C.byref(h), counts))
        try:
            allc = torch.empty((W, W), dtype=torch.int64, device=dev)  # allc[s][d] = records rank s sends to rank d
            dist.all_gather_into_tensor(allc, torch.tensor(list(counts), dtype=torch.int64, device=dev), group=self.group)
            allc = allc.tolist()
            incoming = [sum(allc[s][d] for s in range(W)) for d in range(W)]
            if max(incoming) > self.capacity:
                raise K.KnError(3, f"peer receive buffer holds {self.capacity} records, {max(incoming)} are on their way")
            first = (C.c_longlong * W)(*[sum(allc[s][d] for s in range(self.rank)) for d in range(W)])
            bufs = (C.c_void_p * W)(*self.bufs)
            K._check(L.knc_route_scatter(h, bufs, first))
        finally:
            L.knc_route_end(h)
        dist.barrier(group=self.group)  # every rank's stores have landed (each scatter call returned after its sync)
        m = incoming[self.rank]
        view = torch.as_tensor(_DevView(self.local, (m, 4), "<f4"), device=dev) if m else torch.empty((0, 4), device=dev)
        return view, int(sum(counts))

    def close(self):
        L = self.K.lib()
        for r, p in enumerate(self.bufs):
            if r != self.rank and p:
                L.knc_peer_close(C.c_void_p(p))
        if self.local:
            L.knc_peer_free(C.c_void_p(self.local))
        self.bufs, self.local = [], None


class SlabRows:
    """Result of one rank, in the single-GPU contract made global:
      rows        view of the handle's d_knearests: [query_count, k] GLOBAL sorted positions (torch, device)
      permutation [query_count] this rank's slice of the global permutation (sorted position -> original id)
      base        global sorted position of this rank's first query
    Concatenating rows / permutation of ranks 0..W-1 gives exactly what kn_get_knearests /
    kn_get_permutation return on one GPU.  The rows live in the handle: call free() when done."""

    def __init__(self, handle, rows, permutation, base, info):
        self.handle, self.rows, self.permutation, self.base, self.info = handle, rows, permutation, base, info

    def free(self):
        if self.handle is not None:
            self.handle.free()
            self.handle = None


class _DevView:
    def __init__(self, ptr, shape, typestr="<i4"):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def solve_distributed(local_xyz, first_id: int, n_total: int, k: int, group=None, halo: int | None = None,
                      density: float = 0.0, phases: dict | None = None, host_rows=None,
                      peers: PeerExchange | None = None) -> SlabRows:
    """One pass of the multi-GPU hot path on this rank's chunk (CUDA tensor [n_local, 3]).
    `phases` (optional dict) receives wall-clock milliseconds per phase, with a device sync after each
    (diagnostic runs only: the syncs are not free).  `host_rows` (optional page-locked uint32 numpy array with
    room for this rank's rows) makes the solve stream its rows to the host while computing
    (knc_solve_to_host).  `peers` (a PeerExchange) replaces route + NCCL all-to-all by the fused
    route-into-peer-memory kernel."""
    import torch
    import torch.distributed as dist
    import cuda_knearests_b200 as K
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dims, zsplit = grid_and_split(n_total, world, k, density)
    D = dims[2]  # planes
    halo = halo_planes(k, n_total, D) if halo is None else halo
    info = {"grid_dim": list(dims), "zsplit": zsplit, "attempts": 0}
    t_last = [time.perf_counter()]
    dev = local_xyz.device

    def mark(name):
        if phases is not None:
            torch.cuda.synchronize()
            now = time.perf_counter()
            phases[name] = phases.get(name, 0.0) + 1e3 * (now - t_last[0])
            t_last[0] = now

    while True:
        info["attempts"] += 1
        info["halo"] = halo
        zb, ze, ozb, oze = held_planes(zsplit, rank, halo, D)
        if peers is not None and host_rows is None:
            # ---- the fast path: everything enqueued on one stream, two host waits per step -----------------------
            with torch.cuda.stream(peers.stream):
                peers.stream.wait_stream(torch.cuda.default_stream(dev))   # the caller's chunk is ready
                m, qbegin, qcount, base, sent = peers.route_async(local_xyz, first_id, D, zsplit, halo)
                mark("route_into_peers")
                info.update(points_held=m, sent=sent, planes=(zb, ze, ozb, oze))
                p = rows = perm = None
                missed = torch.zeros(1, dtype=torch.int32, device=dev)
                if oze > ozb and m > 0:
                    recv = torch.as_tensor(_DevView(peers.local, (m, 4), "<f4"), device=dev)
                    p = K.KnProblem(recv, k=k, grid_dims=dims, slab=(zb, ze, ozb, oze), stream=peers.stream.cuda_stream,
                                    query_hint=(qbegin, qcount))
                    mark("prepare")
                    K._check(K.lib().knc_set_index_offset(p._h, base - qbegin))
                    p.solve_async()
                    missed.copy_(torch.as_tensor(_DevView(p.halo_flag_ptr(), (1,)), device=dev))
                dist.all_reduce(missed, op=dist.ReduceOp.MAX, group=group)   # the ranks' vote, still in stream order
                if p is not None:
                    try:
                        p.solve_finish()                                     # second host wait: the step is complete
                    except K.KnError as e:
                        if e.code != 5:  # KNC_ERR_HALO: the vote below says the same to everybody
                            raise
                    mark("solve")
                    info.update(p.timings(), queries=qcount)
                    perm = torch.as_tensor(_DevView(p.tags_ptr() + 4 * qbegin, (qcount,)), device=dev)
                    rows = torch.as_tensor(_DevView(p.info.d_knearests, (qcount, k)), device=dev)
                ok = int(missed.item()) == 0
                mark("halo_vote")
            if ok:
                return SlabRows(p, rows, perm, base, info)
            if p is not None:
                p.free()
            halo += 2
            continue
        if peers is not None:
            recv_xyz, sent = peers.route(local_xyz, first_id, D, zsplit, halo)  # [m,4] records: x, y, z, bits(global id)
            mark("route_into_peers")
        else:
            send_rec, counts = route_device(local_xyz, first_id, D, zsplit, halo)
            mark("route")
            recv_xyz, recv_counts = exchange(send_rec, counts, group)
            mark("exchange")
            sent = int(sum(counts))
            del send_rec
        info.update(points_held=int(recv_xyz.shape[0]), sent=sent, planes=(zb, ze, ozb, oze))
        p = None
        mine = [0, 0]
        if oze > ozb and recv_xyz.shape[0] > 0:
            p = K.KnProblem(recv_xyz, k=k, grid_dims=dims, slab=(zb, ze, ozb, oze))
            mine = [p.info.query_count, p.info.query_begin]
        mark("prepare")
        # global sorted position of a local one = local + (queries of all lower ranks - own query_begin)
        allq = torch.empty((world, 2), dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(allq, torch.tensor(mine, dtype=torch.int64, device=dev), group=group)
        allq = allq.tolist()
        base = sum(int(allq[d][0]) for d in range(rank))
        ok = 1
        perm = rows = None
        if p is not None:
            K._check(K.lib().knc_set_index_offset(p._h, base - p.info.query_begin))
            try:
                if host_rows is not None:
                    p.solve_to_host(host_rows, chunks=8)
                else:
                    p.solve()
                mark("solve")
                info.update(p.timings(), queries=p.info.query_count)
                # this rank's slice of the global permutation: the records' ids, carried through the build as tags
                perm = torch.as_tensor(_DevView(p.tags_ptr() + 4 * p.info.query_begin, (p.info.query_count,)), device=dev)
                mark("permutation_slice")
                rows = torch.as_tensor(_DevView(p.info.d_knearests, (p.info.query_count, k)), device=dev)
            except K.KnError as e:
                if e.code != 5:  # KNC_ERR_HALO
                    raise
                ok = 0
        mark("rows_view")
        flag = torch.tensor([ok], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        mark("halo_vote")
        if int(flag.item()) == 1:
            return SlabRows(p, rows, perm, base, info)
        if p is not None:
            p.free()
        halo += 2  # some query reached beyond its halo: widen everywhere and repeat (never silently wrong)


# ------------------------------------------------------------------------------------------- bench
def bench_multi(args, rank, world, local_rank, METRIC, UNIT, SEED, workload_config, ClockSampler, measured_peak_gbs,
                verify=None):
    """bench.py for N > 1: 100 M points cut into z-slabs, strong scaling.  A step is the whole multi-GPU
    hot path: route + NCCL all-to-all + kn_prepare (slab) + kn_solve (rows in global sorted positions) +
    this rank's slice of the global permutation."""
    import json
    import torch
    import torch.distributed as dist
    import cuda_knearests_b200 as K
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    n, k = args.n, args.k
    lo, hi = (rank * n) // world, ((rank + 1) * n) // world
    pts = K.generate_uniform_device(hi - lo, seed=SEED, first=lo, device=dev)
    # receive buffers of all ranks mapped into every process: 1.25x the even share + the halo, generous for uniform clouds
    peers = None
    if os.environ.get("KN_EXCHANGE", "p2p") != "nccl":
        try:
            peers = PeerExchange(int(1.35 * n / world) + (1 << 20))
        except K.KnError as e:  # raised on every rank together: all switch to the staged NCCL exchange
            if rank == 0:
                print(f"bench: {e}; using the staged NCCL exchange", file=sys.stderr)

    def step():
        r = solve_distributed(pts, lo, n, k, density=args.density, peers=peers)
        info = r.info
        r.free()
        return info

    infos, step_s = [], []
    import gc
    gc.collect()
    gc.disable()  # a collection in the middle of a 5 ms step is 1-2 ms of host time on one rank, and everybody waits for it
    with ClockSampler(local_rank) as clk:  # started before the warm-up: its first NVML calls cost a step ~20 ms (see bench.py)
        for _ in range(args.warmup):
            step()
        dist.barrier()
        torch.cuda.synchronize()
        clk.reset()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            ts = time.perf_counter()
            infos.append(step())
            step_s.append(time.perf_counter() - ts)   # every step ends with a host wait on this rank's stream
        torch.cuda.synchronize()
        dist.barrier()
        wall = time.perf_counter() - t0
    gc.enable()
    t = torch.tensor([wall], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    wall = float(t.item())
    phase_ms = {}
    rr = None
    for _ in range(2):  # untimed, instrumented (second pass: warm)
        phase_ms.clear()
        del rr  # or its permutation block is still held and the pass below pays a cudaMalloc for another
        torch.cuda.synchronize()
        tt = time.perf_counter()
        rr = solve_distributed(pts, lo, n, k, density=args.density, phases=phase_ms, peers=peers)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        rr.free()
        torch.cuda.synchronize()
        phase_ms["free"] = 1e3 * (time.perf_counter() - t1)
        phase_ms["total_with_syncs"] = 1e3 * (time.perf_counter() - tt)
    stats = torch.tensor([infos[-1].get("solve_ms", 0.0), infos[-1].get("build_ms", 0.0),
                          float(infos[-1].get("queries", 0)), float(infos[-1]["points_held"]),
                          float(infos[-1]["sent"])], dtype=torch.float64, device=dev)
    allstats = [torch.empty_like(stats) for _ in range(world)]
    dist.all_gather(allstats, stats)

    # after the timed region: every rank hands sampled rows of one more pass to the checker the caller passed in
    # (bench.py) -- rows in global sorted positions, the id table (tags) and the cell pointers of its slab
    checks = [None] * world
    if verify is not None:
        r = solve_distributed(pts, lo, n, k, density=args.density, peers=peers)
        mine = {"queries": 0, "mismatches": 0}
        if r.handle is not None:
            i = r.handle.info
            cells = i.dimx * i.dimy * (i.z_end - i.z_begin)
            mine = verify(torch, r.rows, torch.as_tensor(_DevView(r.handle.tags_ptr(), (i.n,)), device=dev),
                          torch.as_tensor(_DevView(i.d_ptrs, (cells + 1,)), device=dev), i, r.base - i.query_begin,
                          n, k, seed=rank)
        r.free()
        dist.all_gather_object(checks, mine)

    e2e = None
    if not args.no_e2e:
        host_in = torch.empty((hi - lo, 3), dtype=torch.float32, pin_memory=True)
        host_in.copy_(pts)
        nq_max = int(max(s[2].item() for s in allstats))
        host_rows = torch.empty((nq_max + 1024, k), dtype=torch.int32, pin_memory=True)
        host_qid = torch.empty((nq_max + 1024,), dtype=torch.int32, pin_memory=True)
        host_rows_np = host_rows.numpy().view(np.uint32)

        def step_e2e():
            d_in = host_in.to(dev, non_blocking=True)
            r = solve_distributed(d_in, lo, n, k, density=args.density, host_rows=host_rows_np, peers=peers)
            nq = r.rows.shape[0]
            host_qid[:nq].copy_(r.permutation, non_blocking=True)
            torch.cuda.synchronize()
            r.free()

        step_e2e()
        dist.barrier()
        torch.cuda.synchronize()
        e_steps = max(2, min(args.steps, 3))
        t0 = time.perf_counter()
        for _ in range(e_steps):
            step_e2e()
        dist.barrier()
        te = torch.tensor([(time.perf_counter() - t0) / e_steps], dtype=torch.float64, device=dev)
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e = {"value": n / float(te.item()), "unit": UNIT, "h2d_bytes_per_step": 12 * n,
               "d2h_bytes_per_step": 4 * k * n + 4 * n, "ms_per_step": 1e3 * float(te.item()), "steps": e_steps,
               "call": "per rank: pinned host chunk -> device, route + all-to-all + knc_prepare_ex(slab) + knc_solve_to_host + "
                       "knc_get_tags (permutation slice); rows (global sorted positions) stream to pinned host memory during the solve "
                       "(knc_solve_to_host), permutation slice copied after"}

    if rank == 0:
        peak, peak_kind = measured_peak_gbs()
        solve_ms = max(float(s[0].item()) for s in allstats)
        queries = [int(s[2].item()) for s in allstats]
        worst = max(range(world), key=lambda r: float(allstats[r][0].item()))
        achieved = (12 + 4 * k) * queries[worst] / (float(allstats[worst][0].item()) * 1e-3) / 1e9
        cfg = workload_config(args)
        layout = dict(halo_planes=infos[-1]["halo"], grid_dims=infos[-1]["grid_dim"],
                      points_held_per_rank=[int(s[3].item()) for s in allstats],
                      records_sent_per_rank=[int(s[4].item()) for s in allstats],
                      exchange=("routing kernel stores every group straight into the destination GPU's receive buffer over "
                                "NVLink (CUDA IPC peer memory) + all-gather of the group sizes + barrier, inside the timed step"
                                if peers is not None else "NCCL all_to_all_single of 16-byte records, inside the timed step"))
        line = {"metric": METRIC, "value": n / (wall / args.steps), "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": 1e3 * wall / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
                "layout": layout,
                "clocks": clk.summary(), "e2e": e2e, "gpu_launches": 16 * args.steps * world,  # per rank: route 3, build 5 + 3 (z-band split, slabs of 4 M points and more), query 5
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": None, "kernel": "kn_query_tile_kernel + retry kernel (slowest rank)",
                             "kernel_ms": float(allstats[worst][0].item()),
                             "algorithmic_bytes": (12 + 4 * k) * queries[worst],
                             "peak_kind": f"{peak_kind} copy bandwidth (burst), per GPU"},
                "cpu_baseline": None,
                "device_ms": {"query_max_over_ranks": solve_ms,
                              "grid_build_max_over_ranks": max(float(s[1].item()) for s in allstats)},
                "verify": None if verify is None else {
                    "queries": sum(c["queries"] for c in checks), "mismatches": sum(c["mismatches"] for c in checks),
                    "per_rank": checks},
                "step_ms_rank0": {"min": round(1e3 * min(step_s), 3), "median": round(1e3 * sorted(step_s)[len(step_s) // 2], 3),
                                  "max": round(1e3 * max(step_s), 3), "all": [round(1e3 * v, 2) for v in step_s]},
                "queries_per_rank": queries, "halo_attempts": infos[-1]["attempts"],
                "phase_ms_rank0_instrumented": {kk: round(v, 3) for kk, v in phase_ms.items()}}
        print(json.dumps(line), flush=True)
    dist.barrier()
    if peers is not None:
        peers.close()
    dist.destroy_process_group()
