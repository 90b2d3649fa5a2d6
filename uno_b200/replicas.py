"""Host-side logic of the multi-GPU arm: independent replicas (one KKT system per rank, no data-path collective).

The only exchanges are control-plane reductions: the maximum over ranks of the timed region and the agreement on
which workload fits every device.  bench.py uses these with the NCCL backend; tests/test_replicas.py covers them
with gloo on CPU (world_size 2)."""
from __future__ import annotations


def rank_seed(rank: int) -> int:
    """Every rank factorizes its own system: same pattern family, different values."""
    return int(rank)


def max_over_ranks(seconds: float, dist=None, device="cpu") -> float:
    import torch
    t = torch.tensor([float(seconds)], dtype=torch.float64, device=device)
    if dist is not None and dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def all_ranks_ok(ok: bool, dist=None, device="cpu") -> bool:
    import torch
    t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=device)
    if dist is not None and dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(int(t.item()))


def aggregate_throughput(steps: int, world: int, max_seconds: float) -> float:
    """Whole-job units per second: every rank did `steps` units in at most `max_seconds`."""
    return world * steps / max_seconds
