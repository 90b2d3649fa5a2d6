"""bench.py --impl reference (the CPU arm) on CPU: the line it prints, and the guards around the big workload."""
import json
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def test_reference_arm_prints_a_measured_line_for_the_workload_it_names():
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--workload", "ocp3d_k32", "--steps", "2", "--warmup", "1"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line["impl"] == "reference" and line["metric"] == "kkt_ldlt_factorize_solve_per_sec" and line["unit"] == "factorize+solve/s"
    assert line["higher_is_better"] is True and line["dtype"] == "f64" and line["gpu_launches"] == 0
    assert line["config"]["workload"].startswith("ocp3d_k32")
    assert line["value"] > 0 and abs(line["value"] * line["ms_per_step"] - 1e3) < 1e-6 * 1e3     # value = 1 / seconds per step, measured
    assert line["e2e"] == {"value": line["value"], "unit": line["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = line["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == line["value"] and "measured in full" in cb["sample"]
    assert cb["inertia"] == [65536, 32768, 0] and cb["host_dgemm_peak_gflops"] > 0
    assert line["steps"] >= 1 and line["measured_runs"][-1]["workload"].startswith("ocp3d_distributed_k32")


def test_memory_estimate_refuses_what_the_host_cannot_hold(monkeypatch):
    """The 128^3 step is attempted only when the oracle's storage, estimated from the plan, fits a share of the host memory."""
    sys.path.insert(0, str(ROOT))
    import bench
    from oracle import cpu_arm as CA
    info_128 = {"nnz_factor": 9.16e9, "max_front": 49367, "cb_arena_doubles": 6.1e9, "N": 4292608}     # profiles/r02_bench_k128_v4.json
    need = CA.oracle_memory_estimate_gb(info_128)
    assert 120.0 < need < 150.0                       # measured on the pool's hosts: ~95 GB resident (the estimate errs on the safe side)
    monkeypatch.setattr(CA, "ordering_from_child", lambda name, slot, seed=0: (None, info_128))
    monkeypatch.setattr(CA, "host_memory_available_gb", lambda: 64.0)
    from uno_b200 import generators as G
    monkeypatch.setitem(bench.WORKLOADS, "ocp3d_k128", (bench.WORKLOADS["ocp3d_k128"][0], lambda G_, seed: G.ocp_distributed(4, 2, seed=seed), None))
    with pytest.raises(MemoryError):
        bench.measure_cpu(G, "ocp3d_k128", 1, 1, 0, memory_fraction=0.8)
