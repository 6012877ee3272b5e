"""Multi-GPU plumbing: candidate structures shard across ranks as contiguous blocks; the only collective on the
data path is the gather of the float64 efficiencies.  Works with NCCL (GPU) and gloo (CPU tests)."""
from __future__ import annotations


def shard_range(total: int, rank: int, world: int):
    """Contiguous shard [lo, hi) of ``total`` units for ``rank``; the first ``total % world`` ranks get one extra."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, rem = divmod(int(total), int(world))
    lo = rank * base + min(rank, rem)
    hi = lo + base + (1 if rank < rem else 0)
    return lo, hi


def gather_efficiencies(local_eff, total: int, group=None, out=None):
    """all_gather of contiguous shards -> full [total] tensor on every rank (same device / dtype).
    Equal shards (total % world == 0) take ONE ``all_gather_into_tensor`` straight into ``out`` (allocated once by the
    caller when it sits in a timed loop); ragged shards are padded to the longest one."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return local_eff
    world = dist.get_world_size(group)
    if total % world == 0 and local_eff.numel() * world == total:
        if out is None:
            out = torch.empty(total, dtype=local_eff.dtype, device=local_eff.device)
        dist.all_gather_into_tensor(out, local_eff.contiguous(), group=group)
        return out
    sizes = [shard_range(total, r, world) for r in range(world)]
    maxn = max(hi - lo for lo, hi in sizes)
    pad = torch.zeros(maxn, dtype=local_eff.dtype, device=local_eff.device)
    pad[: local_eff.numel()] = local_eff
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad, group=group)
    return torch.cat([b[: hi - lo] for b, (lo, hi) in zip(bufs, sizes)])


def max_over_ranks(value: float, device=None, group=None) -> float:
    """MAX all-reduce of one host float (bench.py: per-rank device time -> job time)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())
