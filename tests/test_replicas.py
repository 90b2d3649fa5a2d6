"""N > 1 arm of bench.py (independent replicas): control-plane reductions under gloo, world_size 2, on CPU."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, str(ROOT))
    from uno_b200 import KKTSolver, replicas
    from uno_b200 import generators as G
    kkt = G.ocp_distributed(8, 3, seed=replicas.rank_seed(rank))
    host = KKTSolver(device=-1)                      # symbolic phase only: no numeric path without a GPU
    host.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n)
    flops = host.stats()["flops_factor"]
    elapsed = 1.0 + rank                               # rank 1 is the slow one
    tmax = replicas.max_over_ranks(elapsed, dist)
    ok = replicas.all_ranks_ok(rank == 0, dist)        # rank 1 reports "does not fit"
    ok_all = replicas.all_ranks_ok(True, dist)
    checksum = torch.tensor([float(np.sum(kkt.val))], dtype=torch.float64)
    gathered = [torch.zeros(1, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(gathered, checksum)
    out.put((rank, tmax, ok, ok_all, flops, [float(g.item()) for g in gathered]))
    dist.destroy_process_group()


def test_replicas_world_size_two_gloo():
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(out.get(timeout=180) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    from uno_b200 import replicas
    for rank, tmax, ok, ok_all, flops, sums in res:
        assert tmax == 2.0                 # max over ranks, not the own time
        assert ok is False and ok_all is True
        assert flops > 0
        assert sums[0] != sums[1]          # the replicas hold different systems (same pattern, different values)
    assert res[0][4] == res[1][4]          # same pattern -> same plan -> same flop count
    assert replicas.aggregate_throughput(5, world, 2.0) == 5.0
