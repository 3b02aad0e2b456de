"""Frame-pair sharding across GPUs: frame pairs are independent units, so ranks take disjoint
contiguous chunks of a batch / video and no data-path collective exists.  torch.distributed is used only for the
barrier and for reducing the step time to its maximum over ranks (bench.py)."""
from __future__ import annotations


def shard_range(total: int, rank: int, world: int) -> tuple[int, int]:
    """[start, stop) of the frame pairs rank `rank` of `world` processes; chunks differ by at most one pair."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, extra = divmod(max(total, 0), world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def max_over_ranks(value: float, device=None) -> float:
    """Max of a per-rank scalar (e.g. the CUDA-event time of the timed region) over all ranks; identity when
    torch.distributed is not initialised."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_per_rank(values, device=None):
    """All ranks' 1-D float lists (e.g. per-box median depths), in rank order, on every rank."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return [list(values)]
    out = [None] * dist.get_world_size()
    dist.all_gather_object(out, list(values))
    return out
