"""NCCL transport of the sharded hot path: one process per GPU (torch.distributed.run), the library's
own communicator over NVLink.  Needs at least two GPUs; the same logic runs on one GPU through the
local transport in test_gpu_sharded.py."""
import os
import subprocess
import sys

import pytest

import common

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("transport", ["peer", "nccl_only"])
@pytest.mark.parametrize("case", ["small_default", "bundled_min", "mixed_win13", "star_tournament"])
def test_nccl_ranks_match_oracle(case, transport, oracle_bin, datasets, tmp_path):
    """transport "peer": mailbox, peer windows and the shared graph tables over CUDA IPC (NVLink stores / loads);
    "nccl_only" (HERA_NO_PEER_WINDOWS=1): the same protocol with NCCL broadcasts / all-reduce only."""
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    world = min(n, 4)
    if transport == "nccl_only" and case not in ("small_default", "star_tournament"):
        pytest.skip("the fallback transport is covered by two cases")
    golden = common.load_golden(case)
    ds = datasets[golden["dataset"]]
    dump = tmp_path / "o.bin"
    common.run_dump(oracle_bin, ds, dump, golden["options"], extra=["--rng", "philox", "--seed", "5"])
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1", "--master-port", "29517",
           str(common.ROOT / "tests" / "nccl_worker.py"), str(ds), str(dump), "5"] + list(golden["options"])
    env = dict(os.environ, OMP_NUM_THREADS="1")
    if transport == "nccl_only":
        env["HERA_NO_PEER_WINDOWS"] = "1"
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert p.stdout.count(": ok") == world
