"""CPU: the N>1 host path (contiguous shards + gather of efficiencies) with world_size 2 and 3 over gloo."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, total, out_dir):
    import torch
    import torch.distributed as dist
    from physics_informed_reinforcement_learning_b200.dist import shard_range, gather_efficiencies, max_over_ranks
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = shard_range(total, rank, world)
        # stand-in for the per-rank RCWA result: a pure function of the GLOBAL structure index
        local = torch.arange(lo, hi, dtype=torch.float64) * 0.5 + 0.25
        full = gather_efficiencies(local, total)
        tmax = max_over_ranks(float(rank + 1))
        np.save(os.path.join(out_dir, f"r{rank}.npy"), np.concatenate([full.numpy(), [tmax]]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,total", [(2, 4096), (2, 7), (3, 10)])
def test_shard_and_gather(tmp_path, world, total):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(world, port, total, str(tmp_path)), nprocs=world, join=True)
    expect = np.arange(total, dtype=np.float64) * 0.5 + 0.25
    for r in range(world):
        got = np.load(os.path.join(str(tmp_path), f"r{r}.npy"))
        assert np.array_equal(got[:-1], expect)
        assert got[-1] == float(world)        # max over ranks of (rank + 1)
