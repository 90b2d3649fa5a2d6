"""Host-side glue of the subtree-partitioned arm (one process per GPU).

The data path (contribution blocks, boundary vectors, inertia and solution sums) runs inside libb200kkt.so over NCCL;
what is left to the host program is the hand-over of rank 0's NCCL unique id, for which this module uses
torch.distributed (any backend: NCCL in bench.py, gloo in the CPU tests), and the bookkeeping bench.py reports."""
from __future__ import annotations

import numpy as np


def share_unique_id(dist, make_id) -> bytes:
    """Rank 0 calls make_id() (-> 128 bytes); every rank returns the same bytes."""
    rank = dist.get_rank() if dist.is_initialized() else 0
    box = [make_id() if rank == 0 else None]
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast_object_list(box, src=0)
    uid = bytes(box[0])
    if len(uid) != 128:
        raise ValueError("NCCL unique id must be 128 bytes")
    return uid


def create_solver(dist, device: int, **options):
    """A KKTSolver handle of a partitioned solve: world = dist world size, communicator joined."""
    from .capi import KKTSolver, comm_unique_id
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    solver = KKTSolver(device=device, num_gpus=world, **options)
    if world > 1:
        solver.comm_init(rank, world, share_unique_id(dist, comm_unique_id))
    return solver


def transfer_schedule(plan: dict, part: dict, rank: int):
    """What rank `rank` sends and receives, in the order the library walks the cut fronts (level, then front index):
    list of (level, front, peer, 'send'|'recv', contribution-block doubles, boundary rows)."""
    out = []
    owner, parent, level = part["owner"], plan["parent"], plan["level"]
    for s in part["cut_fronts"]:
        s = int(s)
        a, b = int(owner[s]), int(owner[parent[s]])
        if rank not in (a, b):
            continue
        r = int(plan["f"][s] - plan["k"][s])
        out.append((int(level[s]), s, b if rank == a else a, "send" if rank == a else "recv", r * (r + 1) // 2, r))
    return out


def expected_speedup(part: dict, total_flops: float) -> float:
    """Upper bound of the factorization speed-up of this map: total work over the critical path."""
    return float(total_flops / max(part["critical_flops"], 1.0))


def summarize(plan: dict, part: dict, total_flops: float) -> dict:
    f, k = plan["f"], plan["k"]
    cut = part["cut_fronts"]
    r = (f - k)[cut].astype(np.int64)
    sh = np.flatnonzero(part["shared"])
    return {"parts": int(len(part["part_flops"])), "cut_fronts": int(len(cut)),
            "shared_fronts": [(int(f[s]), int(k[s])) for s in sh],
            "shared_flops_share": float(sum(float(k[s]) ** 3 / 3 + float(k[s]) ** 2 * float(f[s] - k[s]) + float(k[s]) * float(f[s] - k[s]) ** 2 for s in sh) / total_flops),
            "flops_share_per_rank": [float(x / total_flops) for x in part["part_flops"]],
            "critical_path_share": float(part["critical_flops"] / total_flops),
            "contribution_block_bytes_total": int((8 * r * (r + 1) // 2).sum())}
