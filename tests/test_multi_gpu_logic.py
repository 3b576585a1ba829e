"""N > 1 host logic on the CPU: world_size 2, gloo.  Plane ownership, halo planes, the all-to-all of
(xyz, id) records in cuda_knearests_b200/slabs.py, and the claim the design rests on: a rank that
receives its slab + halo can answer its own queries exactly.  The device routing kernel is replaced here
by a numpy restatement (the checker side); the oracle does the kNN."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def numpy_route(xyz, first_id, D, zsplit, halo):
    """What knc_slab_route does, restated: stable groups by destination rank."""
    plane = np.clip(np.floor(xyz[:, 2] * np.float32(D) / np.float32(1000)).astype(np.int64), 0, D - 1)
    ids = np.arange(first_id, first_id + len(xyz), dtype=np.int32)
    groups, counts = [], []
    for d in range(len(zsplit) - 1):
        want = (plane >= zsplit[d] - halo) & (plane < zsplit[d + 1] + halo)
        groups.append((xyz[want], ids[want]))
        counts.append(int(want.sum()))
    xyz_all = np.concatenate([g[0] for g in groups])
    ids_all = np.concatenate([g[1] for g in groups])
    rec = np.empty((len(ids_all), 4), np.float32)  # the 16-byte records: x, y, z, bits(id)
    rec[:, :3] = xyz_all
    rec[:, 3] = ids_all.view(np.float32)
    return rec, counts


def _worker(rank, world, port, n, k, out_dir):
    sys.path.insert(0, str(ROOT))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist
    import oracle as O
    from cuda_knearests_b200 import slabs
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pts = O.generate_uniform(n, seed=77)
    dims, zsplit = slabs.grid_and_split(n, world, k)
    D = dims[2]
    halo = slabs.halo_planes(k, n, D)
    lo, hi = rank * n // world, (rank + 1) * n // world
    rec, counts = numpy_route(pts[lo:hi], lo, D, zsplit, halo)
    rrec, rc = slabs.exchange(torch.from_numpy(rec), counts)
    rrec = rrec.numpy()
    rx, ri = np.ascontiguousarray(rrec[:, :3]), np.ascontiguousarray(rrec[:, 3]).view(np.int32)
    zb, ze, ozb, oze = slabs.held_planes(zsplit, rank, halo, D)
    plane = np.clip(np.floor(pts[:, 2] * np.float32(D) / np.float32(1000)).astype(np.int64), 0, D - 1)
    want = np.nonzero((plane >= zb) & (plane < ze))[0]
    ok_ids = np.array_equal(ri, want.astype(np.int32))           # the slab, in global index order
    ok_xyz = np.array_equal(rx, pts[want])
    # own queries answered from the received points only == answered from the whole cloud
    own = np.nonzero((plane[want] >= ozb) & (plane[want] < oze))[0]
    sample = own[:: max(1, len(own) // 300)]
    local = np.stack([O.brute(rx, k, int(q), int(q) + 1)[0] for q in sample])
    glob = np.stack([O.brute(pts, k, int(ri[q]), int(ri[q]) + 1)[0] for q in sample])
    ok_rows = np.array_equal(ri[local], glob.astype(np.int32))
    np.save(Path(out_dir) / f"rank{rank}.npy", np.array([ok_ids, ok_xyz, ok_rows, len(sample), sum(rc)]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_exchange_and_halo_on_cpu(tmp_path):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    n, k, world = 30000, 8, 2
    mp.spawn(_worker, args=(world, port, n, k, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        ok_ids, ok_xyz, ok_rows, nsample, held = np.load(tmp_path / f"rank{r}.npy")
        assert ok_ids and ok_xyz, f"rank {r}: received set is not its slab + halo in global order"
        assert ok_rows and nsample > 100, f"rank {r}: slab + halo does not reproduce the global rows"
        assert n // 2 < held < n


def test_plan_matches_survey():
    sys.path.insert(0, str(ROOT))
    from cuda_knearests_b200 import slabs
    # halo 2 / 3 / 3 / 4 planes for k = 8 / 20 / 32 / 50
    assert [slabs.halo_planes(k) for k in (8, 20, 32, 50)] == [2, 3, 3, 4]
    dims, zs = slabs.grid_and_split(100_000_000, 8, density=3.1)
    D = dims[2]
    assert dims == (318, 318, 318) and zs[0] == 0 and zs[-1] == 318
    sizes = np.diff(zs)
    assert sizes.min() >= 39 and sizes.max() <= 40          # 318 planes over 8 ranks
    assert slabs.held_planes(zs, 0, 3, D) == (0, zs[1] + 3, 0, zs[1])
    assert slabs.held_planes(zs, 7, 3, D) == (zs[7] - 3, 318, zs[7], 318)
    # the library's own grid: planes as thick as the first search radius, so 2 planes of halo for every k
    for k in (8, 20, 32, 50):
        dims, zs = slabs.grid_and_split(100_000_000, 8, k)
        assert dims[2] % 8 == 0 and 0 <= dims[1] - dims[2] < 8 and zs[-1] == dims[2]  # equal planes per rank
        assert len(set(np.diff(zs))) == 1 and slabs.halo_planes(k, 100_000_000, dims[2]) == 2
