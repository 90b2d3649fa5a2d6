"""Subtree partition: the map computed by the symbolic phase, and the transfer protocol of the N > 1 path
run by two gloo ranks on CPU through the numpy emulation of the numeric phase (the CUDA path itself: tests/gpu_partition.py
under torchrun on the GPU box, and test_gpu_parity.py::test_partitioned_two_gpus where two devices are visible)."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def _systems():
    from uno_b200 import generators as G
    return {"ocp3d_12": G.ocp_distributed(12, 3, seed=0), "bnd3d_12": G.ocp_boundary_3d(12, seed=0), "ocp2d_48": G.ocp_distributed(48, 2, seed=0),
            "opf_600": G.opf_like(600, 810, 120, seed=0, delta_w=100.0)}


@pytest.mark.parametrize("parts", [1, 2, 4, 8])
@pytest.mark.parametrize("name", ["ocp3d_12", "bnd3d_12", "ocp2d_48", "opf_600"])
def test_subtree_map_invariants(lib, name, parts):
    from uno_b200 import KKTSolver, partition
    kkt = _systems()[name]
    h = KKTSolver(device=-1, num_gpus=parts)
    h.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n)
    plan, part, st = h.plan(), h.partition(), h.stats()
    owner, parent, level = part["owner"], plan["parent"], plan["level"]
    nsn = plan["num_fronts"]
    assert owner.shape == (nsn,) and owner.min() >= 0 and owner.max() < parts
    assert np.all((parent < 0) | (parent > np.arange(nsn)))               # children are numbered before their parents
    has_parent = parent >= 0
    is_cut = np.zeros(nsn, dtype=bool)
    is_cut[has_parent] = owner[has_parent] != owner[parent[has_parent]]
    assert sorted(part["cut_fronts"].tolist()) == np.flatnonzero(is_cut).tolist()
    keys = [(int(level[s]), int(s)) for s in part["cut_fronts"]]
    assert keys == sorted(keys)                                           # the order every rank walks them in
    assert st["num_cut_fronts"] == len(keys) and st["num_parts"] == parts
    total = st["flops_factor"]
    assert abs(part["part_flops"].sum() - total) <= 1e-9 * total
    assert part["critical_flops"] <= total * (1 + 1e-12)
    assert part["critical_flops"] >= part["part_flops"].max() * (1 - 1e-12) or parts == 1 or len(keys) > 0
    if parts == 1:
        assert len(keys) == 0 and part["critical_flops"] == pytest.approx(total)
    # below a cut front everything belongs to one rank: independent subtrees, no exchange inside
    sub_owner_ok = np.ones(nsn, dtype=bool)
    for s in range(nsn):                                                  # children first
        p = parent[s]
        if p >= 0 and not is_cut[s] and owner[s] != owner[p]:
            sub_owner_ok[s] = False
    assert sub_owner_ok.all()
    # every rank's send list is its peers' receive list, in the same order
    sched = [partition.transfer_schedule(plan, part, r) for r in range(parts)]
    for r in range(parts):
        for q in range(parts):
            a = [(l, s, n) for (l, s, peer, d, n, _) in sched[r] if peer == q and d == "send"]
            b = [(l, s, n) for (l, s, peer, d, n, _) in sched[q] if peer == r and d == "recv"]
            assert a == b
    summ = partition.summarize(plan, part, total)
    assert summ["cut_fronts"] == len(keys) and abs(sum(summ["flops_share_per_rank"]) - 1.0) < 1e-9
    h.close()


def test_shared_top_fronts(lib):
    """Big fronts above the cut are factorized by all ranks together: they sit above every cut front on the way to the root,
    their work is split evenly in the per-rank figures, and switching the option off gives the plain subtree map."""
    from uno_b200 import KKTSolver
    from uno_b200 import generators as G
    kkt = G.ocp_distributed(24, 3, seed=0)
    h = KKTSolver(device=-1, num_gpus=2, share_min_pivots=512, share_min_flops=1e8)
    h.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n)
    plan, part, st = h.plan(), h.partition(), h.stats()
    shared = np.flatnonzero(part["shared"])
    assert len(shared) >= 1 and np.all(plan["k"][shared] >= 512)
    below = np.zeros(plan["num_fronts"], dtype=bool)          # fronts with a cut front or a shared front at or below them
    below[part["cut_fronts"]] = True
    below[shared] = True
    for s in range(plan["num_fronts"]):
        p = plan["parent"][s]
        if p >= 0 and below[s]:
            below[p] = True
    for s in shared:                                           # a shared front never sits inside one rank's private subtree
        p = plan["parent"][s]
        assert p < 0 or below[p]
    assert abs(part["part_flops"].sum() - st["flops_factor"]) <= 1e-9 * st["flops_factor"]
    h2 = KKTSolver(device=-1, num_gpus=2, share_top_fronts=0)
    h2.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n)
    part2 = h2.partition()
    assert not part2["shared"].any() and np.array_equal(part2["owner"], part["owner"])
    assert part["critical_flops"] < part2["critical_flops"]   # sharing the top shortens the critical path
    h.close(); h2.close()


@pytest.mark.parametrize("parts", [2, 4])
def test_storage_is_partitioned_per_rank(lib, parts):
    """partition_memory (default): a rank's factor pool holds the fronts it owns and the shared ones, back to back in front order;
    the other fronts have no storage (an offset far outside the pool), and no rank's pool or arena exceeds the whole system's."""
    from uno_b200 import KKTSolver
    from uno_b200 import generators as G
    kkt = G.ocp_distributed(24, 3, seed=0)
    ref = KKTSolver(device=-1, num_gpus=parts, share_min_pivots=512, share_min_flops=1e8, partition_memory=0)
    ref.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n)
    gplan, gst = ref.plan(), ref.stats()
    assert gst["pool_doubles"] == gst["pool_doubles_global"] and gst["cb_arena_doubles"] == gst["cb_arena_doubles_global"]
    size = gplan["ld"].astype(np.int64) * gplan["k"]
    assert gst["pool_doubles"] == size.sum()
    pools, distinct = [], []
    for r in range(parts):
        h = KKTSolver(device=-1, num_gpus=parts, share_min_pivots=512, share_min_flops=1e8, debug_rank=r)
        h.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n)
        plan, part, st = h.plan(), h.partition(), h.stats()
        assert st["part_rank"] == r and np.array_equal(part["owner"], ref.partition()["owner"])
        held = (part["owner"] == r) | (part["shared"] != 0)
        off = plan["pool_offset"]
        assert np.all(off[~held] < 0)
        expect = np.concatenate([[0], np.cumsum(size[held])])
        assert np.array_equal(off[held], expect[:-1]) and st["pool_doubles"] == expect[-1]
        assert st["pool_doubles_global"] == gst["pool_doubles"] and st["cb_arena_doubles_global"] == gst["cb_arena_doubles"]
        assert 0 < st["pool_doubles"] < gst["pool_doubles"] and 0 < st["cb_arena_doubles"] <= gst["cb_arena_doubles"]
        assert st["assembly_distinct"] <= gst["assembly_distinct"]
        pools.append(st["pool_doubles"]); distinct.append(st["assembly_distinct"])
        h.close()
    shared_doubles = size[ref.partition()["shared"] != 0].sum()
    assert sum(pools) == gst["pool_doubles"] + (parts - 1) * shared_doubles       # owned parts are disjoint, shared fronts are on every rank
    assert sum(distinct) >= gst["assembly_distinct"]
    ref.close()


def test_map_is_deterministic(lib):
    from uno_b200 import KKTSolver
    kkt = _systems()["ocp3d_12"]
    maps = []
    for _ in range(2):
        h = KKTSolver(device=-1, num_gpus=4)
        h.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n)
        maps.append(h.partition()["owner"].copy())
        h.close()
    assert np.array_equal(maps[0], maps[1])


def test_multi_gpu_handle_needs_communicator_and_device(lib):
    """No CPU numeric path and no silent single-GPU fallback: a host-only multi-GPU handle analyses but does not factorize."""
    from uno_b200 import B200KKTError, KKTSolver
    kkt = _systems()["ocp3d_12"]
    h = KKTSolver(device=-1, num_gpus=2)
    h.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n)
    with pytest.raises(B200KKTError):
        h.do_numerical_factorization(kkt.val)
    with pytest.raises(B200KKTError):
        h.comm_init(0, 2, b"\0" * 128)
    h.close()


# ---- world_size 2, gloo: the transfer protocol of the partitioned numeric phase ----------------------------------------
class GlooTransport:
    def exchange(self, sends, recvs):
        bufs = [torch.empty(shape, dtype=torch.float64) for _, shape in recvs]
        ops = [dist.P2POp(dist.isend, torch.from_numpy(np.array(a, dtype=np.float64)), peer) for peer, a in sends]
        ops += [dist.P2POp(dist.irecv, b, peer) for (peer, _), b in zip(recvs, bufs)]
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        return [b.numpy() for b in bufs]

    def allreduce_sum(self, a):
        t = torch.from_numpy(np.array(a, dtype=np.float64))
        dist.all_reduce(t)
        return t.numpy()

    def bcast(self, a, root):
        t = torch.from_numpy(np.array(a, dtype=np.float64))
        dist.broadcast(t, src=root)
        return t.numpy()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, str(ROOT))
    sys.path.insert(0, str(ROOT / "tests"))
    import emulator
    from helpers import backward_error
    from uno_b200 import KKTSolver, partition
    from uno_b200 import generators as G
    uid = partition.share_unique_id(dist, lambda: bytes(range(128)))        # the id hand-over bench.py uses (NCCL id there)
    res = {"rank": rank, "uid_ok": uid == bytes(range(128))}
    k = 8 if world == 2 else 12
    for name, kkt in (("ocp3d", G.ocp_distributed(k, 3, seed=1)), ("bnd3d", G.ocp_boundary_3d(k, seed=2))):
        host = KKTSolver(device=-1, num_gpus=world, partition_memory=0)      # the emulation keeps the whole system's pool on every rank
        host.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n)
        plan, part = host.plan(), host.partition()
        perm, scatter = host.permutation()
        emu = emulator.PartitionedEmulation(plan, part, rank, GlooTransport(), perm, scatter, kkt, kkt.val)
        ine = emu.inertia()
        x = emu.solve(kkt.rhs)
        whole = emulator.EmulatedFactorization(plan, perm, scatter, kkt, kkt.val)   # the unpartitioned run of the same plan
        x1 = whole.solve(kkt.rhs)
        res[name] = {"inertia": ine, "inertia_whole": whole.inertia(), "expected": (kkt.n, kkt.m, 0), "berr": float(backward_error(kkt, kkt.val, x)),
                     "diff": float(np.max(np.abs(x - x1)) / np.max(np.abs(x1))), "cuts": len(part["cut_fronts"]),
                     "checksum": float(np.sum(x)), "own_fronts": int((part["owner"] == rank).sum())}
        host.close()
    out.put(res)
    dist.destroy_process_group()


def _worker_shared(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, str(ROOT))
    sys.path.insert(0, str(ROOT / "tests"))
    import emulator
    from helpers import backward_error
    from uno_b200 import KKTSolver
    from uno_b200 import generators as G
    kkt = G.ocp_boundary_3d(26, seed=4)     # shared fronts (2001, 649) - contribution block, 649 is no multiple of the panel width - and the root
    host = KKTSolver(device=-1, num_gpus=world, share_min_pivots=512, share_min_flops=1e8, partition_memory=0)
    host.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n)
    plan, part = host.plan(), host.partition()
    perm, scatter = host.permutation()
    emu = emulator.PartitionedEmulation(plan, part, rank, GlooTransport(), perm, scatter, kkt, kkt.val)
    ine = emu.inertia()
    x = emu.solve(kkt.rhs)
    shared = np.flatnonzero(part["shared"])
    out.put({"rank": rank, "inertia": ine, "expected": (kkt.n, kkt.m, 0), "berr": float(backward_error(kkt, kkt.val, x)), "checksum": float(np.sum(x)),
             "shared": [(int(plan["f"][s]), int(plan["k"][s])) for s in shared], "levels_ok": _shared_levels_above_private(plan, part)})
    host.close()
    dist.destroy_process_group()


def _shared_levels_above_private(plan, part):
    """Every front with a shared front at or below it sits above all purely private fronts (partitioned level schedule)."""
    nsn = plan["num_fronts"]
    top = np.array(part["shared"], dtype=bool).copy()
    for s in range(nsn):                       # children are numbered before their parents
        if top[s] and plan["parent"][s] >= 0:
            top[plan["parent"][s]] = True
    if not top.any() or top.all():
        return True
    return bool(plan["level"][top].min() > plan["level"][~top].max())


def test_shared_front_protocol_world_size_two_gloo(lib):
    """The shared-front factorization (column chunks per rank, rotated ownership, panels factored by the chunk owner and broadcast,
    straggler columns of a last panel, contribution block gathered on every rank) run by two gloo ranks through the numpy emulation:
    exact inertia, backward error, the same solution on both ranks."""
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_shared, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted((out.get(timeout=600) for _ in range(world)), key=lambda r: r["rank"])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for r in res:
        assert len(r["shared"]) >= 1 and any(k < f for f, k in r["shared"])      # at least one shared front with a contribution block
        assert tuple(r["inertia"]) == tuple(r["expected"])
        assert r["berr"] <= 1e-12
        assert r["levels_ok"]
    assert res[0]["checksum"] == res[1]["checksum"]


@pytest.mark.parametrize("world", [2, 4])
def test_partitioned_protocol_gloo(lib, world):
    """world size 2 and 4: blocks, boundary sums and boundary solutions between arbitrary pairs of ranks."""
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted((out.get(timeout=300) for _ in range(world)), key=lambda r: r["rank"])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for name in ("ocp3d", "bnd3d"):
        per_rank = [r[name] for r in res]
        assert per_rank[0]["cuts"] >= world - 1 and all(r["own_fronts"] > 0 for r in per_rank)
        for r in per_rank:
            assert tuple(r["inertia"]) == tuple(r["inertia_whole"]) == tuple(r["expected"])
            assert r["berr"] <= 1e-12
            assert r["diff"] <= 1e-9
        assert len({r["checksum"] for r in per_rank}) == 1      # every rank ends up with the same solution vector
    assert all(r["uid_ok"] for r in res)
