"""Run oracle/_ref/uno_driver (unchanged Uno host code + this repo's adapter) and parse its output."""
import json
import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
DRIVER = ROOT / "oracle" / "_ref" / "uno_driver"


def run_driver(model, linear_solver, preset="ipopt", extra=(), timeout=600):
    cmd = [str(DRIVER), model, f"preset={preset}", f"linear_solver={linear_solver}", *extra]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout).stdout
    result = None
    regularization = []
    for line in out.splitlines():
        if line.startswith("UNO_DRIVER_RESULT"):
            result = json.loads(line[len("UNO_DRIVER_RESULT"):])
        parts = line.split()
        # iteration rows of Uno's statistics table: iter, LS iter, barrier, step length, phase, regulariz, ...
        if len(parts) >= 7 and parts[0].isdigit() and parts[4] in ("OPT", "FEAS"):
            regularization.append(parts[5])
    if result is None:
        raise RuntimeError("uno_driver produced no result line:\n" + out[-2000:])
    result["regularization_per_iteration"] = regularization
    return result
