"""The real multi-process path (torchrun, NCCL, CUDA IPC peer buffers) -- needs at least 2 GPUs on the box, so it
is skipped on a single-GPU box; tests/test_multi_gpu.py covers the same arithmetic with emulated ranks there."""
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
pytestmark = pytest.mark.gpu


def test_two_or_more_ranks_fused_and_nccl_exchange_match_the_oracle():
    import torch
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs >= 2 GPUs")
    world = min(ngpu, 4)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", str(ROOT / "tests" / "mp_slab_check.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("-> ok") == 3 * world
