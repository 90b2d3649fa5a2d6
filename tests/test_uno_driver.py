"""End-to-end parity through the UNCHANGED Uno host code (last row): the bundled examples must
reach the same iteration count, termination status and per-iteration regularization with the B200 solver as the
reference run (Uno + CPU oracle, golden in tests/golden/uno_examples.json).  The driver binary is built by
oracle/build_ref.py where /root/reference exists and travels to the GPU box prebuilt."""
import json
from pathlib import Path

import pytest

from uno_driver_util import DRIVER, run_driver

GOLDEN = json.loads((Path(__file__).resolve().parent / "golden" / "uno_examples.json").read_text())
MODELS = ["hs015", "polak5", "control2d:16", "control3d:6"]
needs_driver = pytest.mark.skipif(not DRIVER.exists(), reason="oracle/_ref/uno_driver not built (needs /root/reference at build time)")


def same_run(result, golden):
    assert result["iterations"] == golden["iterations"]
    assert result["optimization_status"] == golden["optimization_status"]
    assert result["iterate_status"] == golden["iterate_status"]
    assert result["regularization_per_iteration"] == golden["regularization_per_iteration"]
    assert abs(result["objective"] - golden["objective"]) <= 1e-6 * max(1.0, abs(golden["objective"]))   # reference CI uses 1e-4


@needs_driver
@pytest.mark.parametrize("model", MODELS)
def test_reference_run_reproduces_golden(model):
    same_run(run_driver(model, "ORACLE"), GOLDEN[model])


def test_known_optima():
    # hs015: f* = 306.5 (Hock & Schittkowski), polak5: f* = 50
    assert abs(GOLDEN["hs015"]["objective"] - 306.5) <= 1e-4
    assert abs(GOLDEN["polak5"]["objective"] - 50.0) <= 1e-4
    assert GOLDEN["hs015"]["iterate_status"] == GOLDEN["polak5"]["iterate_status"] == "Feasible KKT point"


@pytest.mark.gpu
@needs_driver
@pytest.mark.parametrize("model", MODELS)
def test_b200_solver_matches_reference_run(model):
    same_run(run_driver(model, "B200"), GOLDEN[model])


@pytest.mark.gpu
@needs_driver
def test_b200_is_default_linear_solver():
    # first (here: only product) entry of available_solvers() becomes the default (DefaultOptions.cpp:203-206)
    import subprocess
    out = subprocess.run([str(DRIVER)], capture_output=True, text=True).stdout
    assert "Linear solvers: B200" in out


def _expected_primal_regularization(k, shift, scale=1.0, initial=1e-4, increase=2.0):
    """PrimalRegularization.hpp:90-129 with dense eigenvalues as the inertia: delta = 0 when the smallest diagonal entry is positive,
    else initial - smallest; doubled (or set to `initial` when it was 0) until H + delta I is positive definite."""
    import numpy as np
    import scipy.sparse as sp
    t = sp.diags([-np.ones(k - 1), np.zeros(k), -np.ones(k - 1)], [-1, 0, 1])
    e = sp.identity(k)
    H = scale * ((sp.kron(e, t) + sp.kron(t, e)).toarray() + (4.0 - shift) * np.eye(k * k))
    smallest = scale * (4.0 - shift)
    delta = 0.0 if smallest > 0.0 else initial - smallest
    while np.linalg.eigvalsh(H + delta * np.eye(k * k)).min() <= 0.0:
        delta = initial if delta == 0.0 else increase * delta
    return delta


@needs_driver
@pytest.mark.parametrize("shift", [1.5, 0.1, 4.5])
def test_primal_regularization_strategy_drives_the_oracle(shift):
    """The reference's PrimalRegularization::regularize_hessian (byrd cadence, Hessian only, m = 0, lazy create + initialize_memory
    of the solver by name: PrimalRegularization.hpp:79-129) run unchanged on an indefinite grid Hessian."""
    r = run_driver("hessian:12", "ORACLE", extra=(f"shift={shift}",))
    assert r["regularization_factors"] == [_expected_primal_regularization(12, shift), _expected_primal_regularization(12, shift, scale=2.0)]
    assert r["inertia"] == [144, 0, 0] and r["residual"] <= 1e-10 * max(1.0, r["solution_norm"])


@pytest.mark.gpu
@needs_driver
@pytest.mark.parametrize("shift", [1.5, 0.1, 4.5])
def test_primal_regularization_strategy_drives_the_b200_solver(shift):
    """the same strategy object, the B200 solver behind it - same shifts, inertia (n, 0, 0)."""
    ref = run_driver("hessian:40", "ORACLE", extra=(f"shift={shift}",))
    r = run_driver("hessian:40", "B200", extra=(f"shift={shift}",))
    assert r["regularization_factors"] == ref["regularization_factors"]
    assert r["inertia"] == ref["inertia"] == [1600, 0, 0]
    assert r["residual"] <= 1e-10 * max(1.0, r["solution_norm"])


@pytest.mark.gpu
@needs_driver
@pytest.mark.parametrize("model", ["hs015", "control3d:6"])
def test_b200_solver_spmd_on_two_gpus(model, tmp_path):
    """two copies of the Uno process, one GPU each (B200KKT_WORLD = 2), run the same iteration; every
    solver instance joins its own NCCL communicator through the hand-over file.  Both ranks must reproduce the golden run."""
    import os
    import subprocess
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices (run with gpurun --gpus 2)")
    procs = []
    for rank in range(2):
        env = dict(os.environ, B200KKT_WORLD="2", B200KKT_RANK=str(rank), B200KKT_COMM_FILE=str(tmp_path / "nccl_id"), B200KKT_RUN_ID=f"t{os.getpid()}")
        procs.append(subprocess.Popen([str(DRIVER), model, "preset=ipopt", "linear_solver=B200"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env))
    outs = [p.communicate(timeout=600)[0] for p in procs]
    for out in outs:
        lines = [l for l in out.splitlines() if l.startswith("UNO_DRIVER_RESULT")]
        assert lines, out[-1500:]
        r = json.loads(lines[-1][len("UNO_DRIVER_RESULT"):])
        g = GOLDEN[model]
        assert r["iterations"] == g["iterations"] and r["optimization_status"] == g["optimization_status"] and r["iterate_status"] == g["iterate_status"]
        assert abs(r["objective"] - g["objective"]) <= 1e-6 * max(1.0, abs(g["objective"]))
    assert not list(tmp_path.glob("nccl_id*"))          # rank 0 removes every hand-over file once the ranks have joined
