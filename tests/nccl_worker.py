"""One rank of the NCCL-transport parity test (launched by test_gpu_nccl.py through torch.distributed.run):
runs the sharded hot path on its GPU and compares every table with the oracle dump given on the command line."""
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent))
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import common  # noqa: E402
from heradump import read_dump  # noqa: E402
from test_gpu_parity import GRAPH_SECTIONS, GROUP_SECTIONS, PATH_SECTIONS  # noqa: E402


def main():
    ds, dump, seed = Path(sys.argv[1]), Path(sys.argv[2]), int(sys.argv[3])
    opts = sys.argv[4:]
    from telomeri_b200 import capi, pipeline
    dist.init_process_group("gloo")   # only hands out the NCCL id; the data path uses the library's own communicator
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.cuda.set_device(rank)
    ids = [capi.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, 0)
    h = capi.Hera(rank)
    h.comm_init_nccl(rank, world, ids[0])
    assert (h.rank, h.size) == (rank, world)
    gpu = pipeline.run_hot_path(ds, opts, seed=seed, hera=h)
    h.close()
    ora = read_dump(dump)
    common.assert_sections_equal(ora, gpu, GRAPH_SECTIONS + ["rank_0", "rank_1"] + PATH_SECTIONS + GROUP_SECTIONS, f"nccl rank {rank}/{world}")
    dist.barrier()
    print(f"rank {rank}: ok")


if __name__ == "__main__":
    main()
