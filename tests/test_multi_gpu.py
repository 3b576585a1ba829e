"""Multi-GPU path on ONE GPU: the ranks are emulated one after the other (route every rank's chunk with
knc_slab_route, hand each destination its groups in rank order -- what the all-to-all delivers -- then
slab prepare + solve + rows in global ids).  The rows must be the single-GPU rows, bit for bit.
The NCCL exchange itself is exercised by `bench.py --gpus N` under torchrun and, on the CPU, by
tests/test_multi_gpu_logic.py."""
import ctypes as C

import numpy as np
import pytest

from test_multi_gpu_logic import numpy_route

pytestmark = pytest.mark.gpu


def test_route_kernel_is_the_stable_multisplit(kn, oracle):
    import torch
    from cuda_knearests_b200 import slabs
    n, world = 300_000, 4
    pts = oracle.generate_uniform(n, seed=31)
    dims, zsplit = slabs.grid_and_split(1_200_000, world, density=3.1)
    D = dims[2]
    dev = torch.from_numpy(pts).cuda()
    for halo in (0, 3, 40):
        rec, counts = slabs.route_device(dev, 5_000_000, D, zsplit, halo)
        want, wc = numpy_route(pts, 5_000_000, D, zsplit, halo)
        assert counts == wc
        got = rec[:sum(counts)].cpu().numpy()
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))  # bit compare: ids are not floats


@pytest.mark.parametrize("world,k,n", [(2, 20, 400_000), (4, 8, 600_000), (8, 20, 1_000_000), (3, 50, 300_000),
                                       (2, 8, 42_000_000)])  # last: slabs big enough for the z-band pre-pass
def test_emulated_ranks_reproduce_single_gpu_rows(kn, oracle, world, k, n):
    """Rows in GLOBAL sorted positions + permutation slices of the ranks, concatenated, are exactly the
    single-GPU kn_get_knearests / kn_get_permutation arrays; and the same rows in original ids
    (knc_get_knearests_global) are the single-GPU rows in original ids."""
    import torch
    from cuda_knearests_b200 import slabs
    pts = oracle.generate_uniform(n, seed=20171004)
    dims, zsplit = slabs.grid_and_split(n, world, k)  # the library's own grid with planes in multiples of `world`, as the driver uses it
    with kn.KnProblem(pts, k=k, grid_dims=dims, algorithm=3 if n > 4_000_000 else 0) as p:  # big case: reference rows without the pre-pass
        p.solve()
        want_rows, want_perm, want_orig = p.get_knearests(), p.get_permutation(), p.get_knearests_original()
    D = dims[2]
    halo = slabs.halo_planes(k, n, D)
    dev = torch.from_numpy(pts).cuda()
    routed = []
    for r in range(world):
        lo, hi = r * n // world, (r + 1) * n // world
        rec, counts = slabs.route_device(dev[lo:hi].contiguous(), lo, D, zsplit, halo)
        routed.append((rec, np.concatenate([[0], np.cumsum(counts)])))
    got_rows, got_perm = [], []
    got_orig = np.full((n, k), 0xFFFFFFFF, np.uint32)
    base = 0
    for d in range(world):
        rx = torch.cat([rec[off[d]:off[d + 1]] for rec, off in routed]).contiguous()  # [m,4] records
        ids_ptr = C.c_void_p(rx.data_ptr() + 12)                                       # id words, stride 4
        zb, ze, ozb, oze = slabs.held_planes(zsplit, d, halo, D)
        with kn.KnProblem(rx, k=k, grid_dims=dims, slab=(zb, ze, ozb, oze)) as p:
            nq = p.info.query_count
            # original ids first (no offset), then the global-position form
            p.solve()
            rows = torch.empty((nq, k), dtype=torch.int32, device="cuda")
            qid = torch.empty((nq,), dtype=torch.int32, device="cuda")
            torch.cuda.synchronize()
            kn._check(kn.lib().knc_get_knearests_global(p._h, ids_ptr, 4, C.c_void_p(rows.data_ptr()),
                                                        C.c_void_p(qid.data_ptr())))
            got_orig[qid.cpu().numpy()] = rows.cpu().numpy().view(np.uint32)
            kn._check(kn.lib().knc_set_index_offset(p._h, base - p.info.query_begin))
            p.solve()
            got_rows.append(p.get_knearests())
            # the same call AFTER the offset (rows now hold global positions): must subtract it again
            rows2 = torch.full((nq, k), -1, dtype=torch.int32, device="cuda")
            torch.cuda.synchronize()
            kn._check(kn.lib().knc_get_knearests_global(p._h, ids_ptr, 4, C.c_void_p(rows2.data_ptr()), None))
            assert torch.equal(rows2, rows)
            perm = torch.empty((nq,), dtype=torch.int32, device="cuda")
            torch.cuda.synchronize()
            kn._check(kn.lib().knc_get_permutation_global(p._h, ids_ptr, 4, C.c_void_p(perm.data_ptr())))
            got_perm.append(perm.cpu().numpy().view(np.uint32))
            # the same slice without a gather: the ids travelled through the build as tags (knc_get_tags)
            tags = torch.as_tensor(slabs._DevView(p.tags_ptr() + 4 * p.info.query_begin, (nq,)), device="cuda")
            assert torch.equal(tags, perm)
            base += nq
    assert base == n                                            # every point is a query of exactly one rank
    assert np.array_equal(np.concatenate(got_perm), want_perm)  # global permutation, slice by slice
    assert np.array_equal(np.concatenate(got_rows), want_rows)  # rows in global sorted positions
    assert np.array_equal(got_orig, want_orig)                  # rows in original ids


@pytest.mark.parametrize("world,k,n", [(2, 20, 400_000), (5, 8, 700_000), (3, 50, 300_000)])
def test_emulated_ranks_through_the_asynchronous_entry_points(kn, oracle, world, k, n):
    """The calls the multi-GPU driver makes on its fast path, with the ranks emulated on one GPU: knc_route_begin_async
    (counts stay on the device: records / own / low-halo per destination), knc_route_scatter_async into the destinations'
    buffers, knc_prepare_ex on the caller's stream with the query range taken from the counts (no read-back), and
    knc_solve_async / knc_solve_finish.  Rows and permutation slices, concatenated, are the single-GPU arrays."""
    import torch
    from cuda_knearests_b200 import slabs
    L = kn.lib()
    pts = oracle.generate_uniform(n, seed=20171004)
    dims, zsplit = slabs.grid_and_split(n, world, k)
    with kn.KnProblem(pts, k=k, grid_dims=dims) as p:
        p.solve()
        want_rows, want_perm = p.get_knearests(), p.get_permutation()
    D = dims[2]
    halo = slabs.halo_planes(k, n, D)
    dev = torch.from_numpy(pts).cuda()
    side = torch.cuda.Stream()
    st = C.c_void_p(side.cuda_stream)
    zs = (C.c_int * (world + 1))(*zsplit)
    torch.cuda.synchronize()
    handles, counts = [], []
    for r in range(world):                       # step 1 of every rank: count, counts left on the device
        lo, hi = r * n // world, (r + 1) * n // world
        c3 = torch.full((3 * world,), -1, dtype=torch.int64, device="cuda")
        h = C.c_void_p()
        chunk = dev[lo:hi].contiguous()
        kn._check(L.knc_route_begin_async(C.c_void_p(chunk.data_ptr()), hi - lo, lo, D, world, zs, halo, st,
                                          C.c_void_p(c3.data_ptr()), C.byref(h)))
        handles.append((h, chunk))
        counts.append(c3)
    side.synchronize()
    a = torch.stack(counts).cpu().numpy()        # what the all-gather delivers: a[s] = rank s's 3 rows
    sent, own, low = a[:, :world], a[:, world:2 * world], a[:, 2 * world:]
    plane = np.clip(np.floor(pts[:, 2] * np.float32(D) / np.float32(1000)).astype(np.int64), 0, D - 1)
    for d in range(world):                       # the counts themselves, against numpy
        zb, ze, ozb, oze = slabs.held_planes(zsplit, d, halo, D)
        for s_ in range(world):
            pl = plane[s_ * n // world:(s_ + 1) * n // world]
            assert sent[s_, d] == ((pl >= zb) & (pl < ze)).sum()
            assert own[s_, d] == ((pl >= ozb) & (pl < oze)).sum()
            assert low[s_, d] == ((pl >= zb) & (pl < ozb)).sum()
    incoming = sent.sum(axis=0)
    recv = [torch.empty((int(incoming[d]) + 1, 4), dtype=torch.float32, device="cuda") for d in range(world)]
    bufs = (C.c_void_p * world)(*[b.data_ptr() for b in recv])
    for r in range(world):                       # step 2: every rank scatters straight into the destinations' buffers
        first = (C.c_longlong * world)(*[int(sent[:r, d].sum()) for d in range(world)])
        kn._check(L.knc_route_scatter_async(handles[r][0], bufs, first, st))
        kn._check(L.knc_route_end(handles[r][0]))
    got_rows, got_perm = [], []
    base = 0
    for d in range(world):                       # step 3: slab handles on the caller's stream, nothing read back
        zb, ze, ozb, oze = slabs.held_planes(zsplit, d, halo, D)
        qbegin, qcount = int(low[:, d].sum()), int(own[:, d].sum())
        with kn.KnProblem(recv[d][:int(incoming[d])], k=k, grid_dims=dims, slab=(zb, ze, ozb, oze),
                          stream=side.cuda_stream, query_hint=(qbegin, qcount)) as p:
            assert (p.info.query_begin, p.info.query_count) == (qbegin, qcount)
            kn._check(L.knc_set_index_offset(p._h, base - qbegin))
            p.solve_async()
            flag = torch.as_tensor(slabs._DevView(p.halo_flag_ptr(), (1,)), device="cuda")
            p.solve_finish()
            assert int(flag.item()) == 0
            got_rows.append(p.get_knearests())
            got_perm.append(torch.as_tensor(slabs._DevView(p.tags_ptr() + 4 * qbegin, (qcount,)), device="cuda").cpu().numpy().view(np.uint32))
            assert p.timings()["build_ms"] > 0
            base += qcount
    assert base == n
    assert np.array_equal(np.concatenate(got_perm), want_perm)
    assert np.array_equal(np.concatenate(got_rows), want_rows)
    # a slab on the caller's stream without the hint cannot learn its query range without waiting: refused
    with pytest.raises(kn.KnError):
        kn.KnProblem(recv[0][:int(incoming[0])], k=k, grid_dims=dims, slab=slabs.held_planes(zsplit, 0, halo, D),
                     stream=side.cuda_stream)


@pytest.mark.parametrize("ranks,k,n", [(2, 20, 300_000), (3, 50, 250_000), (5, 8, 600_000), (8, 20, 900_000)])
def test_one_process_multi_gpu_handle_returns_the_single_gpu_arrays(kn, oracle, ranks, k, n):
    """knc_prepare_multi (C ABI, one process, a host thread per rank; on a box with fewer GPUs than ranks the ranks
    share devices -- same code path, the "peer" stores are local): kn_solve / kn_get_* on that handle give exactly
    the arrays of the single-GPU handle, and the oracle's."""
    pts = oracle.generate_uniform(n, seed=20171004)
    g = oracle.Grid(pts, dims=oracle.fast_dims(n, k))
    with kn.prepare_multi(pts, k=k, ngpus=ranks) as p:
        assert (p.info.n, p.info.query_count, p.info.dimx, p.info.dimy, p.info.dimz) == (n, n, *g.dims)
        p.solve()
        rows, perm, spts, stats = p.get_knearests(), p.get_permutation(), p.get_points(), p.stats()
        t = p.timings()
        with pytest.raises(kn.KnError):
            p.get_dists()                      # per-slab extras are refused on the spanning handle, loudly
    assert np.array_equal(perm, g.permutation)
    assert np.array_equal(spts, g.sorted_xyz)
    assert np.array_equal(rows, g.knn_exact(k))
    assert stats["total"] == n and stats["histogram"][:63] == np.bincount(g.counters, minlength=64)[:63].tolist()
    assert t["solve_ms"] > 0 and t["build_ms"] > 0


def test_solve_multi_in_one_call_and_the_dropin_caller_with_kn_num_gpus(kn, oracle, pts20k, tmp_path):
    import os
    import subprocess
    from test_dropin_cpp import BIN, build_dropin
    n, k = 120_000, 20
    pts = oracle.generate_uniform(n, seed=5)
    g = oracle.Grid(pts, dims=oracle.fast_dims(n, k))
    rows = np.empty((n, k), np.uint32)
    perm = np.empty(n, np.uint32)
    kn._check(kn.lib().knc_solve_multi(C.c_void_p(pts.ctypes.data), n, k, 4, C.c_void_p(rows.ctypes.data), C.c_void_p(perm.ctypes.data)))
    assert np.array_equal(perm, g.permutation) and np.array_equal(rows, g.knn_exact(k))
    # the reference's seven entry points, unchanged caller (tests/cpp/dropin_main.cu), three ranks by environment
    if not BIN.exists():
        build_dropin()
    pts20k.tofile(tmp_path / "points.f32")
    r = subprocess.run([str(BIN), str(tmp_path / "points.f32"), str(tmp_path / "neighbors.u32")], capture_output=True,
                       text=True, env={**os.environ, "KN_NUM_GPUS": "3"})
    assert r.returncode == 0, r.stderr
    got = np.fromfile(tmp_path / "neighbors.u32", np.uint32).reshape(-1, 50)
    g = oracle.Grid(pts20k, dims=oracle.fast_dims(len(pts20k), 50))
    want = np.empty_like(got)
    want[g.permutation] = g.permutation[g.knn_exact(50)]
    assert np.array_equal(got, want)
