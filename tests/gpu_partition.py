"""Subtree-partitioned solve on N GPUs against the single-GPU solve of the same system (run under torchrun on the GPU box):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/gpu_partition.py [k]

Every rank must return the same inertia and the same x as the one-GPU handle (to rounding), backward error <= 1e-12."""
import json
import os
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from uno_b200 import KKTSolver, partition  # noqa: E402
from uno_b200 import generators as G  # noqa: E402


def emit(record):
    """One write per line: the ranks share the launcher's stdout, a separate newline write can land behind another rank's line."""
    sys.stdout.write(json.dumps(record) + "\n")
    sys.stdout.flush()


def main():
    names = [a for a in sys.argv[1:] if "=" not in a] or ["ocp3d:16", "bnd3d:20", "opf", "ocp3d:32"]
    options = {a.split("=")[0]: float(a.split("=")[1]) for a in sys.argv[1:] if "=" in a}   # e.g. share_min_pivots=512
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ok_all = True
    for name in names:
        kind, _, arg = name.partition(":")
        if kind == "ocp3d":
            kkt = G.ocp_distributed(int(arg), 3, seed=0)
        elif kind == "bnd3d":
            kkt = G.ocp_boundary_3d(int(arg), seed=0)
        elif kind == "ocp2d":
            kkt = G.ocp_distributed(int(arg), 2, seed=0)
        else:
            kkt = G.opf_like(3000, 4050, 600, seed=0, delta_w=100.0)
        multi = partition.create_solver(dist, local, **options)
        multi.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n, values=kkt.val)
        st = multi.stats()
        times = []
        for it in range(3):
            dist.barrier(); torch.cuda.synchronize()
            t0 = time.perf_counter()
            multi.do_numerical_factorization(kkt.val)
            ine = multi.get_inertia()
            x, berr = multi.solve_indefinite_system(kkt.rhs, return_backward_error=True)
            torch.cuda.synchronize()
            times.append(time.perf_counter() - t0)
        stm = multi.stats()
        # inertia-loop step: only the regularization slots change
        if kkt.reg_size > 0:
            multi.refactorize_shifted(kkt.val[:kkt.reg_size])
            assert multi.get_inertia() == ine
        single = KKTSolver(device=local)
        single.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n, values=kkt.val)
        for it in range(2):
            single.do_numerical_factorization(kkt.val)
            ine1 = single.get_inertia()
            x1, berr1 = single.solve_indefinite_system(kkt.rhs, return_backward_error=True)
        st1 = single.stats()
        diff = float(np.max(np.abs(x - x1)) / max(np.max(np.abs(x1)), 1e-300))
        # all ranks hold the same x bit for bit
        xs = torch.from_numpy(x).cuda()
        x0 = xs.clone()
        dist.broadcast(x0, src=0)
        same = bool(torch.equal(xs, x0))
        ok = ine == ine1 and berr <= 1e-12 and diff <= 1e-6 and same
        ok_all &= ok
        part = multi.partition()
        emit({"rank": rank, "case": kkt.name, "N": kkt.dim, "ok": ok, "inertia": ine, "inertia_1gpu": ine1, "berr": berr, "berr_1gpu": berr1,
                          "x_rel_diff_vs_1gpu": diff, "x_identical_on_all_ranks": same, "cut_fronts": int(st["num_cut_fronts"]),
                          "share": st["flops_own"] / st["flops_factor"], "critical_share": st["flops_critical"] / st["flops_factor"],
                          "factor_ms": 1e3 * stm["t_factor_last"], "solve_ms": 1e3 * stm["t_solve_last"], "factor_ms_1gpu": 1e3 * st1["t_factor_last"],
                          "solve_ms_1gpu": 1e3 * st1["t_solve_last"], "wall_ms": 1e3 * min(times), "comm_MB_factor": stm["comm_bytes_factor"] / 1e6,
                          "owners": np.bincount(part["owner"], minlength=world).tolist(), "shared_fronts": int(part["shared"].sum()),
                          "speedup_factor": st1["t_factor_last"] / stm["t_factor_last"],
                          "pool_MB": 8e-6 * st["pool_doubles"], "pool_MB_whole_system": 8e-6 * st["pool_doubles_global"],
                          "arena_MB": 8e-6 * st["cb_arena_doubles"], "arena_MB_whole_system": 8e-6 * st["cb_arena_doubles_global"]})
        multi.close(); single.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok_all else 1)


if __name__ == "__main__":
    main()
