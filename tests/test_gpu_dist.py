"""GPU, 2+ devices (skipped on a one-GPU box): the design's one data-path collective on NCCL -- contiguous shards of a
global batch evaluated per rank, `gather_efficiencies` (all_gather) on the device, every rank ends with the full vector,
equal to a single-GPU evaluation of the whole batch."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, total, out_dir):
    import torch
    import torch.distributed as dist
    from physics_informed_reinforcement_learning_b200 import RCWABatch, shard_range, gather_efficiencies
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        pats = (np.random.default_rng(3).integers(0, 2, (total, 256)) * 2 - 1).astype(np.int8)
        lo, hi = shard_range(total, rank, world)
        be = RCWABatch(256, 40, max_batch=max(hi - lo, 8), device=rank)
        eff = be.evaluate(torch.from_numpy(pats[lo:hi]).cuda())
        full = gather_efficiencies(eff, total)
        torch.cuda.synchronize()
        np.save(os.path.join(out_dir, f"r{rank}.npy"), full.cpu().numpy())
        be.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("total", [512, 301])
def test_nccl_gather_of_sharded_efficiencies(tmp_path, total):
    import torch
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs two GPUs")
    mp.spawn(_worker, args=(world, _free_port(), total, str(tmp_path)), nprocs=world, join=True)
    from physics_informed_reinforcement_learning_b200 import RCWABatch
    pats = (np.random.default_rng(3).integers(0, 2, (total, 256)) * 2 - 1).astype(np.int8)
    ref = RCWABatch(256, 40, max_batch=total, device=0).evaluate_host(pats)
    for r in range(world):
        got = np.load(os.path.join(str(tmp_path), f"r{r}.npy"))
        assert got.shape == (total,) and np.array_equal(got, ref)
