"""Worker for tests/test_multi_gpu_nccl.py -- run under torchrun, one rank per GPU.
Solves one cloud with the real multi-process driver twice (fused route-into-peer-memory exchange, and the NCCL
all-to-all exchange) and checks both against the whole-cloud oracle: the ranks' rows (global sorted positions) and
permutation slices, concatenated, must be the single-GPU arrays."""
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import cuda_knearests_b200 as K  # noqa: E402
import oracle as O  # noqa: E402
from cuda_knearests_b200 import slabs  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n, k = 700_000, 20
    pts = O.generate_uniform(n, seed=4242)
    g = O.Grid(pts, dims=slabs.grid_and_split(n, world, k)[0])  # the grid slabs.solve_distributed uses when no density is named
    want_rows, want_perm = g.knn_exact(k), g.permutation
    lo, hi = rank * n // world, (rank + 1) * n // world
    mine = torch.from_numpy(pts[lo:hi]).cuda()
    peers = slabs.PeerExchange(int(1.5 * n / world) + 65536)
    ok = True
    for mode, px in (("p2p", peers), ("nccl", None), ("p2p again", peers)):
        r = slabs.solve_distributed(mine, lo, n, k, peers=px)
        rows = r.rows.cpu().numpy().view(np.uint32)
        perm = r.permutation.cpu().numpy().view(np.uint32)
        good = (np.array_equal(rows, want_rows[r.base:r.base + len(rows)]) and
                np.array_equal(perm, want_perm[r.base:r.base + len(perm)]))
        print(f"rank {rank} {mode}: base {r.base} queries {len(rows)} halo {r.info['halo']} -> {'ok' if good else 'MISMATCH'}",
              flush=True)
        ok = ok and good
        r.free()
    total = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(total, op=dist.ReduceOp.MIN)
    peers.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if int(total.item()) == 1 else 1)


if __name__ == "__main__":
    main()
