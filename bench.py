#!/usr/bin/env python
"""bench.py — KKT LDL^T numeric factorize + inertia + solve(with refinement) per second.

A "step" is one pass of the hot path over one synthetic KKT system with a fixed sparsity pattern:
numeric factorization (K1' scaling, K1 assembly, K2-K5 multifrontal LDL^T), inertia (K6) and one
solve with iterative refinement (K7/K8).  Symbolic analysis is done once before the timed region,
as the reference does (PrimalDualRegularization.hpp:145-149), and reported separately.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl b200|reference]

`value`   : device-resident inputs, timed with CUDA events on the solver's own stream.
`e2e`     : same metric through the C ABI with HOST buffers (H2D of values+rhs, D2H of x inside).
`roofline`: K4 trailing-update kernel (FP64 tensor pipe) — algorithmic flops / CUDA-event time of
            its launches on fronts of order >= 1024, against cuBLAS DGEMM measured in this run.
`cpu_baseline` / `--impl reference`: the CPU oracle (oracle/kkt_oracle.cpp, a restatement of the
            MUMPS algorithm; MUMPS itself is not installable here) on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os

# before anything loads libgomp: the CPU arm's OpenMP workers spin for about a millisecond after a region, then sleep - the pivot
# loop of a big front opens thousands of short regions back to back; every BLAS call of the oracle runs inside its OpenMP threads
os.environ.setdefault("GOMP_SPINCOUNT", "3000000")
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "kkt_ldlt_factorize_solve_per_sec"
UNIT = "factorize+solve/s"

WORKLOADS = {
    # name: (description, generator, cpu-sample generator)
    "ocp3d_k128": ("BASELINE configs[4]: synthetic 3D PDE-constrained optimal control KKT, 128^3 grid, boundary control "
                   "(2.2 M primal rows + 2.1 M multipliers), whole system on ONE B200",
                   lambda G, seed: G.ocp_boundary_3d(128, seed=seed), lambda G, seed: G.ocp_boundary_3d(64, seed=seed)),
    "ocp3d_k64": ("BASELINE configs[3]: synthetic 3D PDE-constrained optimal control KKT, 64^3 grid, 7-point, distributed control",
                  lambda G, seed: G.ocp_distributed(64, 3, seed=seed), lambda G, seed: G.ocp_distributed(48, 3, seed=seed)),
    "ocp2d_k256": ("BASELINE configs[1]: synthetic 2D PDE-constrained optimal control KKT, 256x256 grid",
                   lambda G, seed: G.ocp_distributed(256, 2, seed=seed), lambda G, seed: G.ocp_distributed(256, 2, seed=seed)),
    "ocp3d_k32": ("3D optimal control KKT, 32^3 grid (reduced; smoke-sized)",
                  lambda G, seed: G.ocp_distributed(32, 3, seed=seed), lambda G, seed: G.ocp_distributed(20, 3, seed=seed)),
    "opf_10k": ("BASELINE configs[2]: AC-OPF-like KKT on a random 10k-bus network (delta_w = 100 regularization)",
                lambda G, seed: G.opf_like(seed=seed, delta_w=100.0), lambda G, seed: G.opf_like(seed=seed, delta_w=100.0)),
}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for name, flag in zip(names, parts[3:7]):
                if flag.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d.get("hbm_gbs", 6650.0)), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def fp64_dgemm_peak(torch):
    """FP64 tensor-pipe denominator: cuBLAS DGEMM 8192^3 through torch.matmul, best of 6 (MEASURED_PEAKS.json has no FP64 entry)."""
    n = 8192
    a = torch.randn(n, n, dtype=torch.float64, device="cuda")
    b = torch.randn(n, n, dtype=torch.float64, device="cuda")
    torch.matmul(a, b)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e-3)
    del a, b
    torch.cuda.empty_cache()
    return 2.0 * n ** 3 / best / 1e12


def cpu_threads():
    cores = os.cpu_count() or 1
    # "all the host threads it can use": torchrun exports OMP_NUM_THREADS=1, which OpenBLAS and the oracle's OpenMP loops would obey
    if os.environ.get("OMP_NUM_THREADS", "1") == "1":
        os.environ["OMP_NUM_THREADS"] = str(cores)
    os.environ.setdefault("OPENBLAS_NUM_THREADS", str(cores))
    return cores


def measure_cpu(G, name, slot, steps, warmup, memory_fraction=None):
    """One MEASURED CPU run (oracle/cpu_arm.py): workload `name` itself (slot 1) or its bounded sample (slot 2), same ordering as the
    GPU run (computed by a child process), all host cores.  Nothing is projected.  memory_fraction: refuse (MemoryError) when the
    oracle's storage, estimated from the plan, exceeds that share of the available host memory."""
    from oracle import cpu_arm as CA
    cores = cpu_threads()
    kkt = WORKLOADS[name][slot](G, 0)
    perm, info = CA.ordering_from_child(name, slot)
    if memory_fraction is not None:
        need, have = CA.oracle_memory_estimate_gb(info), CA.host_memory_available_gb()
        if need > memory_fraction * have:
            raise MemoryError(f"the CPU oracle needs about {need:.0f} GB for {name}, {have:.0f} GB of host memory available, limit {100 * memory_fraction:.0f} %")
    res = CA.run_oracle(kkt, perm, steps, warmup, cores)
    res.update({"workload": kkt.name, "N": kkt.dim, "factor_flops": info["flops"], "factor_gflops": info["flops"] / res["factor_seconds"] / 1e9,
                "value": 1.0 / res["seconds_per_step"], "steps": steps, "warmup": warmup, "cores": cores, "cpu": CA.cpu_model(),
                "expected_inertia": [kkt.n, kkt.m, 0], "plan": info})
    res["host_dgemm_peak_gflops"] = CA.host_dgemm_peak_gflops()
    res["peak_rss_gb"] = CA.peak_rss_gb()
    mp = CA.mumps_probe()
    res["mumps"] = mp["note"] if not mp["available"] else CA.run_mumps(mp["exe"], kkt, perm, cores)
    return res


def cpu_baseline_object(res, what):
    return {"value": res["value"], "unit": UNIT, "cores": res["cores"], "kind": "port",
            "sample": f"{what}: {res['workload']} (N={res['N']}, {res['factor_flops']:.3g} factor flops) measured in full, {res['steps']} step(s) after {res['warmup']} warm-up: "
                      f"{res['seconds_per_step']:.2f} s per factorize+inertia+solve, factorization {res['factor_seconds']:.2f} s = {res['factor_gflops']:.0f} GFLOP/s on "
                      f"{res['cores']} x {res['cpu']}; CPU oracle = multifrontal LDL^T, threshold pivoting u=0.01 over the fully-summed rows, delayed pivots, "
                      f"OpenBLAS dgemm/dtrsm {'bound' if res['openblas'] else 'NOT bound (plain loops)'}, same ordering as the GPU run; inertia {res['inertia']}",
            "seconds_per_step": res["seconds_per_step"], "factor_seconds": res["factor_seconds"], "factor_gflops": res["factor_gflops"],
            "steps": res["steps"], "warmup": res["warmup"], "inertia": res["inertia"], "backward_error_first_sweep": res["backward_error_first_sweep"],
            "host_dgemm_peak_gflops": res["host_dgemm_peak_gflops"],
            "dgemm_peak_note": "numpy (OpenBLAS) DGEMM 4096^3 on all host cores, best of 3: no CPU factorization of F flops on this box can take less than F / this",
            "mumps": res["mumps"]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=os.environ.get("B200KKT_WORKLOAD", "auto"), choices=sorted(WORKLOADS) + ["auto"],
                    help="auto = BASELINE configs[4] (128^3, needs ~150 GB of HBM), falling back to configs[3] (64^3) if it does not fit")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--multi", default=os.environ.get("B200KKT_MULTI", "partition"), choices=["partition", "replicas"],
                    help="N > 1: 'partition' = ONE system, elimination subtrees mapped to the GPUs, top fronts shared, NCCL over NVLink (strong "
                         "scaling, BASELINE configs[4]); 'replicas' = one independent system per GPU (weak scaling)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    from uno_b200 import generators as G

    auto = args.workload == "auto"
    if auto:
        args.workload = "ocp3d_k128"
    if args.impl == "reference":
        if rank != 0:
            return 0
        # The CPU arm MEASURES: it runs the workload its line names, in full.  The b200 arm's default workload (128^3, 2e14 factor
        # flops, ~110 GB of factors and workspace on the host) is run - one step - when its bounded sample (the same generator at
        # 64^3), measured first, predicts that the step fits the time budget and the host memory.  Otherwise the line reports
        # BASELINE configs[3] (64^3, distributed control) - config.workload says so - and the driver's same_config check reads
        # false, which is the truth.
        from oracle import cpu_arm as CA
        requested = args.workload
        budget = float(os.environ.get("B200KKT_REFERENCE_BUDGET_S", "900"))
        t_start = time.perf_counter()
        runs = []
        if requested == "ocp3d_k128":
            sample = measure_cpu(G, "ocp3d_k128", 2, 1, 1)
            runs.append(sample)
            full_flops_estimate = 2.014e14                               # the b200 arm's plan (profiles/r02_bench_k128_v3.json)
            # the oracle's rate grows with the front sizes, so the sample's rate is a cautious one; + generator, ordering, analysis
            predicted = full_flops_estimate / (sample["factor_gflops"] * 1e9) + 150.0
            # memory: estimated from the plan before the run (80 % of what the host has), and watched by the oracle during it (it
            # stops with an error while the host can still take it): a host driven out of memory is worse than a line that names
            # the smaller configuration
            res, note = None, None
            if predicted <= budget:
                try:
                    res = measure_cpu(G, "ocp3d_k128", 1, 1, 0, memory_fraction=0.8)
                    runs.append(res)
                except (MemoryError, RuntimeError) as exc:
                    note = f"ocp3d_k128 not measured: {exc}"
            else:
                note = (f"ocp3d_k128 not run: predicted {predicted:.0f} s for one step (sample: {sample['factor_gflops']:.0f} GFLOP/s) against a budget of "
                        f"{budget:.0f} s (B200KKT_REFERENCE_BUDGET_S)")
            if res is None:
                one = measure_cpu(G, "ocp3d_k64", 1, 1, 0)
                runs.append(one)
                more = int(min(max(args.steps, 1), max(1.0, 0.25 * budget / max(one["seconds_per_step"], 1e-3))))
                res = measure_cpu(G, "ocp3d_k64", 1, more, 1) if more > 1 else one
                if more > 1:
                    runs.append(res)
                res["note"] = note
                requested = "ocp3d_k64"
        else:
            one = measure_cpu(G, requested, 1, 1, 0)
            runs.append(one)
            more = int(min(max(args.steps, 1), max(1.0, 0.25 * budget / max(one["seconds_per_step"], 1e-3))))
            res = measure_cpu(G, requested, 1, more, 1) if more > 1 else one
            if more > 1:
                runs.append(res)
        # not a measurement of the workload, a floor: the requested workload's factor flops at 100 % of the host's measured DGEMM peak
        floor = None
        if args.workload == "ocp3d_k128" and requested != args.workload:
            floor = {"workload": args.workload, "factor_flops_estimate": full_flops_estimate, "host_dgemm_peak_gflops": res["host_dgemm_peak_gflops"],
                     "seconds_at_dgemm_peak": full_flops_estimate / (res["host_dgemm_peak_gflops"] * 1e9),
                     "note": "lower bound on ANY CPU factorization of the b200 arm's workload on this box; the b200 arm's ms_per_step divides into it"}
        cb = cpu_baseline_object(res, "reference arm")
        line = {"impl": "reference", "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": res["steps"], "warmup": res["warmup"],
                "steps_requested": args.steps, "warmup_requested": args.warmup,
                "ms_per_step": 1e3 * res["seconds_per_step"], "higher_is_better": True, "scaling": "strong" if (args.gpus > 1 and args.multi == "partition") else "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": f"{requested}: {WORKLOADS[requested][0]}", "N": res["N"], "factor_flops": res["factor_flops"],
                           "parallelism": f"host cores ({res['cores']} threads)", "note": res.get("note")},
                "cpu_baseline": cb, "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0,
                "cpu_time_floor_for_requested_workload": None if floor is None else floor,
                "measured_runs": [{k: r[k] for k in ("workload", "N", "factor_flops", "seconds_per_step", "factor_seconds", "factor_gflops", "steps", "warmup", "inertia")} for r in runs],
                "wall_seconds_total": time.perf_counter() - t_start}
        assert res["inertia"] == res["expected_inertia"], (res["inertia"], res["expected_inertia"])
        print(json.dumps(line))
        return 0

    import torch
    import torch.distributed as dist
    from uno_b200 import KKTSolver, partition, replicas
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device: the numeric phase has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    fallback_note = None
    parted = world > 1 and args.multi == "partition"
    while True:
        desc = WORKLOADS[args.workload][0]
        # partition: every rank holds the same system (SPMD, like the reference's MUMPS over MPI); replicas: one system per rank
        kkt = WORKLOADS[args.workload][1](G, 0 if parted else replicas.rank_seed(rank))
        # B200KKT_OPTIONS="name=value,name=value": solver options for experiments (recorded in config.options)
        extra_options = {kv.split("=")[0]: float(kv.split("=")[1]) for kv in os.environ.get("B200KKT_OPTIONS", "").split(",") if "=" in kv}
        solver = partition.create_solver(dist, local_rank, **extra_options) if parted else KKTSolver(device=local_rank, **extra_options)
        t0 = time.perf_counter()
        ok = True
        try:
            solver.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n, values=kkt.val)   # Uno analyses a filled matrix
        except Exception as exc:   # device pools do not fit
            ok = False
            fallback_note = f"{args.workload} did not fit this device ({exc}); fell back to ocp3d_k64"
        ok = replicas.all_ranks_ok(ok, dist, "cuda")
        if ok or not auto or args.workload == "ocp3d_k64":
            if not ok:
                raise SystemExit(fallback_note or "analysis failed on another rank")
            break
        solver.close()
        del solver
        args.workload = "ocp3d_k64"
    t_analyse = time.perf_counter() - t0
    st0 = solver.stats()
    d_val = torch.from_numpy(kkt.val).cuda()
    d_rhs = torch.from_numpy(kkt.rhs).cuda()
    d_x = torch.empty_like(d_rhs)
    h_val = torch.from_numpy(kkt.val).pin_memory()
    h_rhs = torch.from_numpy(kkt.rhs).pin_memory()
    h_x = torch.empty_like(h_rhs).pin_memory()
    stream = torch.cuda.ExternalStream(solver.stream())
    torch.cuda.synchronize()

    def step_device():
        solver.factorize_device(d_val.data_ptr())
        ine = solver.get_inertia()
        berr = solver.solve_device(d_rhs.data_ptr(), d_x.data_ptr())
        return ine, berr

    def step_host():
        solver._check(solver._lib.b200kkt_factorize(solver._h, h_val.data_ptr()), "factorize")
        ine = solver.get_inertia()
        import ctypes
        berr = ctypes.c_double()
        solver._check(solver._lib.b200kkt_solve(solver._h, h_rhs.data_ptr(), h_x.data_ptr(), ctypes.byref(berr)), "solve")
        return ine, berr.value

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        ine, berr = step_device()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    launches0 = solver.stats()["kernel_launches"]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    t_fact = t_solve = 0.0
    refine = 0
    for _ in range(args.steps):
        ine, berr = step_device()
        s_ = solver.stats()
        t_fact += s_["t_factor_last"]; t_solve += s_["t_solve_last"]; refine += s_["refine_steps_last"]
    e1.record(stream)
    barrier()
    clocks = sampler.stop()
    final_stats = solver.stats()
    launches = final_stats["kernel_launches"] - launches0
    # parity gate of the bench itself: a fast wrong answer is not a result
    if tuple(ine) != (kkt.n, kkt.m, 0) or not (berr <= 1e-12) or not (final_stats["backward_error_componentwise_last"] <= 1e-12):
        raise SystemExit(f"bench: parity check failed: inertia {ine} (expected {(kkt.n, kkt.m, 0)}), backward error {berr}, "
                         f"componentwise {final_stats['backward_error_componentwise_last']}")
    elapsed = replicas.max_over_ranks(e0.elapsed_time(e1) * 1e-3, dist, "cuda")
    units = 1 if parted else world     # systems factorized per step by the whole job
    value = replicas.aggregate_throughput(args.steps, units, elapsed)

    # end to end through the C ABI with host buffers
    step_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ine_h, berr_h = step_host()
    torch.cuda.synchronize()
    e2e_value = replicas.aggregate_throughput(args.steps, units, replicas.max_over_ranks(time.perf_counter() - t0, dist, "cuda"))
    x_ok = bool(np.isfinite(h_x.numpy()).all())

    # secondary metric of SURVEY 8(d): one step of the caller's inertia-correction loop = refactorization after only the leading
    # regularization slots changed (uploads reg_size doubles) + inertia.  One GPU only (on several it is a collective).
    inertia_loop = None
    if world == 1 and kkt.reg_size > 0:
        try:
            reg = np.ascontiguousarray(kkt.val[:kkt.reg_size])
            solver.refactorize_shifted(reg)
            reps = max(1, min(args.steps, 2))
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(reps):
                solver.refactorize_shifted(reg)
                ine_r = solver.get_inertia()
            torch.cuda.synchronize()
            inertia_loop = {"value": reps / (time.perf_counter() - t0), "unit": "refactorize_shifted+inertia/s", "h2d_bytes_per_step": int(8 * kkt.reg_size),
                            "d2h_bytes_per_step": 32, "inertia": list(ine_r), "steps": reps}
        except Exception as exc:   # a reported extra, never the reason for a missing bench line
            inertia_loop = {"value": None, "error": str(exc)}

    def profile_k4():
        # roofline of the dominant kernel (K4): one extra, un-graphed factorization with event pairs around its big-front launches
        solver.set_option("use_graph", 0)
        solver.set_option("profile_updates", 1)
        solver.factorize_device(d_val.data_ptr())
        sp_ = solver.stats()
        solver.set_option("profile_updates", 0)
        solver.set_option("use_graph", 1)
        return sp_

    part_info = None
    if parted:
        sp = profile_k4()        # a partitioned factorization is a collective: every rank runs it
        part_info = partition.summarize(solver.plan(), solver.partition(), st0["flops_factor"])
        part_info["comm_bytes_per_factorization_rank0"] = int(solver.stats()["comm_bytes_factor"])
        part_info["comm_bytes_per_solve_sweep_rank0"] = int(solver.stats()["comm_bytes_sweep"])
        # one profiled factorization (rank 0's main stream): time inside the shared-front batches, of which waiting for a panel
        mine = {"rank": rank, "factor_s": sp["t_factor_last"], "before_shared_s": sp["t_before_shared_profiled"], "transfer_wait_s": sp["t_transfer_wait_profiled"],
                "shared_s": sp["t_shared_profiled"], "panel_wait_s": sp["t_shared_stall_profiled"], "k4_local_s": sp["t_update_profiled"],
                "k4_local_tflops": sp["flops_update_profiled"] / max(sp["t_update_profiled"], 1e-12) / 1e12}
        per_rank = [None] * world
        dist.all_gather_object(per_rank, mine)
        part_info["profiled_factorization_per_rank"] = per_rank
        part_info["shared_phase"] = {"seconds": sp["t_shared_profiled"], "stall_seconds": sp["t_shared_stall_profiled"],
                                     "aggregate_tflops": sp["flops_shared_profiled"] / max(sp["t_shared_profiled"], 1e-12) / 1e12,
                                     "factor_seconds_profiled_run": sp["t_factor_last"]}
    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return 0
    if not parted:
        sp = profile_k4()
    peak64 = fp64_dgemm_peak(torch)
    # DRAM traffic of the dominant kernel from the committed ncu --set full capture of one representative launch
    k4_traffic = {}
    tp = ROOT / "profiles" / "r01_k4_traffic.json"
    if tp.exists():
        t_ = json.loads(tp.read_text())
        k4_traffic = {"dram_bytes_per_launch": t_["dram_bytes_read"] + t_["dram_bytes_write"], "algorithmic_bytes_per_launch": t_["algorithmic_bytes"],
                      "launch": f"rows={t_['rows']} K={t_['K']}", "source": t_["source"]}
    hbm_peak, hbm_src = measured_peaks()
    asm_bytes = 12 * st0["nnz"] + 8 * st0["assembly_distinct"] + 8 * st0["pool_doubles"]
    ach = sp["flops_update_profiled"] / sp["t_update_profiled"] / 1e12 if sp["t_update_profiled"] > 0 else 0.0
    roofline = {"bound": "tensor", "kernel": "k_update<128,64,4,2,16,2> (K4 Schur update, DMMA.8x8x4, 8 warps, two CTAs per SM)", "achieved": ach, "peak": peak64, "unit": "TFLOP/s",
                "frac": ach / peak64 if peak64 > 0 else None, "traffic": k4_traffic.get("dram_bytes_per_launch"), "traffic_detail": k4_traffic,
                "peak_source": "cuBLAS DGEMM 8192^3 via torch.matmul(float64), best of 6, measured in this run (MEASURED_PEAKS.json has no FP64 entry)",
                "launches_profiled": sp["n_update_profiled"], "flops_profiled": sp["flops_update_profiled"], "seconds_profiled": sp["t_update_profiled"],
                "share_of_factor_time": sp["t_update_profiled"] / sp["t_factor_last"] if sp["t_factor_last"] > 0 else None,
                "whole_factorization_tflops": st0["flops_factor"] / (t_fact / args.steps) / 1e12,
                "assembly_hbm": {"kernel": "k_assemble_fill (K1 fused with the zero-fill of the factor pool: every pool tile written once)",
                                 "seconds": sp["t_assemble_profiled"], "algorithmic_bytes": asm_bytes,
                                 "algorithmic_bytes_formula": "12 nnz + 8 distinct destinations (SURVEY 8(d) K1) + 8 pool doubles (zero-fill, part of the same pass)",
                                 "achieved_gbs": asm_bytes / max(sp["t_assemble_profiled"], 1e-12) / 1e9,
                                 "frac": asm_bytes / max(sp["t_assemble_profiled"], 1e-12) / 1e9 / hbm_peak,
                                 "peak_gbs": hbm_peak, "peak_source": hbm_src}}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": 1e3 * elapsed / args.steps, "higher_is_better": True, "scaling": "strong" if parted else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {desc}", "N": kkt.dim, "n": kkt.n, "m": kkt.m, "nnz": kkt.nnz, "fronts": st0["num_fronts"],
                   "max_front": st0["max_front"], "nnz_factor": st0["nnz_factor"], "factor_flops": st0["flops_factor"],
                   "parallelism": "single GPU" if world == 1 else
                   (f"ONE system over {world} GPUs: elimination subtrees mapped to the ranks, big top fronts factorized by all ranks together "
                    f"(column chunks, panels broadcast), contribution blocks and boundary vectors over NCCL/NVLink" if parted else
                    f"{world} independent replicas (one KKT system per GPU, no data-path collective)"),
                   "partition": part_info,
                   "l2": f"front pool {st0['pool_doubles'] * 8 / 1e9:.1f} GB per step >> 126 MB L2 (inputs larger than L2, no explicit flush)",
                   "options": extra_options or None, "analysis_seconds_untimed": t_analyse, "ordering": "METIS nested dissection on the KKT-pair-compressed graph (graphs of 150 000 nodes and more: top three levels split by the library, "
                               "halves ordered by parallel processes; smaller ones: METIS_NodeND)", "fallback": fallback_note},
        "clocks": clocks, "gpu_launches": int(launches),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(8 * kkt.nnz + 8 * kkt.dim), "d2h_bytes_per_step": int(8 * kkt.dim + 32 + 32),
                "finite": x_ok},
        "roofline": roofline,
        "inertia_loop_step": inertia_loop,
        "breakdown": {"factor_ms": 1e3 * t_fact / args.steps, "solve_ms": 1e3 * t_solve / args.steps, "refine_steps": refine / args.steps,
                      "inertia": list(ine), "expected_inertia": [kkt.n, kkt.m, 0], "backward_error": berr,
                      "backward_error_componentwise": final_stats["backward_error_componentwise_last"],
                      "delayed_pivots_since_analysis": final_stats["num_delayed_total"], "replans_since_analysis": final_stats["num_replans_total"]},
    }
    if world == 1 and not args.no_cpu_baseline:
        try:
            # bounded sample of the workload (slot 2 of WORKLOADS), measured; value = the sample's own factorize+solve/s
            res_cpu = measure_cpu(G, args.workload, 2, 2, 1)
            line["cpu_baseline"] = cpu_baseline_object(res_cpu, "bounded sample of the workload")
            # the SAME sample through this arm's host-buffer path (factorize + inertia + solve, copies inside): a measured ratio on one
            # identical system, next to the two arms' lines
            solver.close()
            sk = WORKLOADS[args.workload][2](G, 0)
            s2 = KKTSolver(device=local_rank, **extra_options)
            s2.do_symbolic_analysis(sk.dim, sk.row, sk.col, sk.reg_size, sk.n, values=sk.val)
            t_s2 = []
            for it in range(args.warmup + args.steps):
                torch.cuda.synchronize()
                t0s = time.perf_counter()
                s2.do_numerical_factorization(sk.val)
                ine2 = s2.get_inertia()
                x2 = s2.solve_indefinite_system(sk.rhs)
                t_s2.append(time.perf_counter() - t0s)
            t_gpu = float(np.mean(t_s2[args.warmup:]))
            if list(ine2) != list(res_cpu["inertia"]):
                raise SystemExit(f"bench: inertia of the sample differs between the arms: {ine2} against {res_cpu['inertia']}")
            line["cpu_baseline"]["b200_on_the_same_sample"] = {
                "value": 1.0 / t_gpu, "unit": UNIT, "ms_per_step": 1e3 * t_gpu, "steps": args.steps, "warmup": args.warmup, "inertia": list(ine2),
                "backward_error": s2.stats()["backward_error_last"], "speedup_measured": res_cpu["seconds_per_step"] / t_gpu,
                "note": "host buffers in and out, wall clock around the three calls; same system, same ordering, same inertia as cpu_baseline.value"}
            s2.close()
            floor_s = st0["flops_factor"] / (res_cpu["host_dgemm_peak_gflops"] * 1e9)
            line["cpu_baseline"]["time_floor_this_workload"] = {
                "seconds_at_host_dgemm_peak": floor_s, "b200_seconds_per_step": elapsed / args.steps, "speedup_lower_bound": floor_s / (elapsed / args.steps),
                "note": "this workload's exact factor flops at 100 % of the host's measured DGEMM peak: a floor for any CPU solver on this box, MUMPS included"}
        except Exception as exc:   # the oracle is only a reported baseline
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {exc}"}
    print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
