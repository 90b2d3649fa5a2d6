"""the reference's bundled `.nl` examples (BASELINE.json configs[0]: examples/hs015.nl with preset=ipopt) run
through the UNCHANGED Uno host code without the AMPL Solver Library: uno_plugin/models/NLModel reads the text .nl file, keeps the
expressions as a DAG and differentiates them exactly (reverse sweeps, forward-over-reverse for the Hessian).  Checked against the
golden runs of the hand-written models, closed-form derivatives and finite differences.  (Named test_zz_* so that it runs after
the solver suites.)"""
import json
import re
import subprocess
from pathlib import Path

import numpy as np
import pytest

from uno_driver_util import DRIVER, run_driver

HERE = Path(__file__).resolve().parent
NL = HERE / "golden" / "nl"
GOLDEN = json.loads((HERE / "golden" / "uno_examples.json").read_text())
needs_driver = pytest.mark.skipif(not DRIVER.exists(), reason="oracle/_ref/uno_driver not built (needs /root/reference at build time)")


def derivatives(model, multipliers):
    out = subprocess.run([str(DRIVER), str(model), "derivatives", *[repr(float(y)) for y in multipliers]], capture_output=True, text=True, timeout=120).stdout
    line = next(l for l in out.splitlines() if l.startswith("UNO_DRIVER_DERIVATIVES"))
    d = json.loads(re.sub(r"(?<![\w.])(-?)inf\b", r"\1Infinity", line[len("UNO_DRIVER_DERIVATIVES"):]))
    n, m = d["n"], d["m"]
    J = np.zeros((m, n))
    for j, i, v in d["jacobian"]:
        J[int(j), int(i)] = v
    H = np.zeros((n, n))
    for i, j, v in d["hessian"]:
        assert i <= j                      # upper triangle (AMPLModel.cpp:170-178)
        H[int(i), int(j)] += v
        if i != j:
            H[int(j), int(i)] += v
    d["J"], d["H"] = J, H
    return d


def same_run(result, golden):
    assert result["iterations"] == golden["iterations"]
    assert result["optimization_status"] == golden["optimization_status"]
    assert result["iterate_status"] == golden["iterate_status"]
    assert result["regularization_per_iteration"] == golden["regularization_per_iteration"]
    assert abs(result["objective"] - golden["objective"]) <= 1e-9 * max(1.0, abs(golden["objective"]))


@needs_driver
@pytest.mark.parametrize("name", ["hs015", "polak5"])
def test_nl_file_reproduces_the_golden_run(name):
    """Same iterates as the hand-written model of the same problem (reference run: Uno + CPU oracle as linear solver)."""
    same_run(run_driver(str(NL / f"{name}.nl"), "ORACLE"), GOLDEN[name])


@needs_driver
def test_hs015_closed_form_derivatives():
    y = np.array([0.3, -0.7])
    d = derivatives(NL / "hs015.nl", y)
    x0, x1 = d["x"]
    assert (x0, x1) == (-2.0, 1.0)                                     # start point of examples/hs015.mod
    t = x1 - x0 * x0
    assert d["objective"] == 100 * t * t + (1 - x0) ** 2
    assert np.allclose(d["gradient"], [-400 * x0 * t - 2 * (1 - x0), 200 * t], rtol=1e-15)
    assert np.allclose(d["constraints"], [x0 * x1, x0 + x1 * x1], rtol=1e-15)
    assert np.allclose(d["J"], [[x1, x0], [1.0, 2 * x1]], rtol=1e-15)
    H = np.array([[1200 * x0 * x0 - 400 * x1 + 2, -400 * x0 - y[0]], [-400 * x0 - y[0], 200 - 2 * y[1]]])   # Hess f - sum y_j Hess c_j
    assert np.allclose(d["H"], H, rtol=1e-14)
    assert d["hessian_nonzeros"] == 3 and d["jacobian_nonzeros"] == 4
    assert d["upper"][0] == 0.5 and d["lower"] == [-np.inf, -np.inf] and d["constraint_lower"] == [1.0, 0.0]


def ops_functions(x):
    """tests/golden/nl/ops.nl restated: objective (as Uno sees it: sign * f for a maximization), constraints."""
    x0, x1, x2 = x
    v3 = 2 * x2 + x0 * x1
    f = -((x0 - 1) ** 2) - np.tanh(x1) * x2 - np.arctan(v3) + 0.125 * x2
    c0 = np.exp(x0) + v3 ** 2 / x2 + 0.25 * x2
    c1 = np.sin(x0) + np.cos(x1) + np.sqrt(x2) - np.log(v3) - 0.5 * x1
    return -f, np.array([c0, c1])


def central(fun, x, h=1e-6):
    x = np.asarray(x, dtype=float)
    cols = []
    for i in range(x.size):
        e = np.zeros_like(x); e[i] = h
        cols.append((np.asarray(fun(x + e)) - np.asarray(fun(x - e))) / (2 * h))
    return np.array(cols).T


@needs_driver
def test_operator_coverage_against_finite_differences():
    """Defined variable, maximization, n-ary sum, exp log sin cos sqrt tanh atan, division, powers, linear parts in J and G."""
    y = np.array([0.3, -0.7])
    d = derivatives(NL / "ops.nl", y)
    x = np.array(d["x"])
    assert d["objective_sign"] == -1.0 and np.allclose(x, [0.5, 0.7, 1.3])
    f, c = ops_functions(x)
    assert abs(d["objective"] - f) <= 1e-14 * abs(f)
    assert np.allclose(d["constraints"], c, rtol=1e-14)
    assert np.allclose(d["gradient"], central(lambda z: ops_functions(z)[0], x).ravel(), atol=1e-8)
    assert np.allclose(d["J"], central(lambda z: ops_functions(z)[1], x), atol=1e-8)
    # Hessian of the Lagrangian f - y.c from differences of its exact gradient
    def lagrangian_gradient(z):
        dz = derivatives_at(z)
        return np.array(dz[0]) - y @ dz[1]
    def derivatives_at(z):
        g = central(lambda w: ops_functions(w)[0], z, 1e-5).ravel()
        J = central(lambda w: ops_functions(w)[1], z, 1e-5)
        return g, J
    H = central(lagrangian_gradient, x, 1e-4)
    assert np.allclose(d["H"], 0.5 * (H + H.T), atol=2e-5)
    assert np.allclose(d["H"], d["H"].T)
    assert d["hessian_nonzeros"] == 6          # dense 3 x 3: every pair meets through v3


@needs_driver
def test_unsupported_nl_content_is_rejected(tmp_path):
    text = (NL / "hs015.nl").read_text().replace("o5\nv1\nn2\nO0 0", "o13\nv1\nO0 0", 1)     # floor(): not differentiable, not supported
    bad = tmp_path / "bad.nl"
    bad.write_text(text)
    out = subprocess.run([str(DRIVER), str(bad), "derivatives"], capture_output=True, text=True, timeout=60)
    assert out.returncode != 0 and "not supported" in out.stdout


@pytest.mark.gpu
@needs_driver
@pytest.mark.parametrize("name", ["hs015", "polak5"])
def test_nl_file_with_b200_solver(name):
    """BASELINE.json configs[0] through the drop-in: examples/<name>.nl, preset=ipopt, linear_solver=B200."""
    result = run_driver(str(NL / f"{name}.nl"), "B200")
    golden = GOLDEN[name]
    assert result["iterations"] == golden["iterations"]
    assert result["optimization_status"] == golden["optimization_status"] and result["iterate_status"] == golden["iterate_status"]
    assert abs(result["objective"] - golden["objective"]) <= 1e-6 * max(1.0, abs(golden["objective"]))
