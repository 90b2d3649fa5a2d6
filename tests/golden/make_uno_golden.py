"""Generates tests/golden/uno_examples.json: the reference run of the bundled examples = UNCHANGED Uno host code
(compiled from /root/reference by oracle/build_ref.py) + the CPU oracle as linear solver (MUMPS is not installable
here).  Run in the authoring container:  python tests/golden/make_uno_golden.py"""
import json
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parent.parent))
sys.path.insert(0, str(HERE.parent))
from oracle import build_ref  # noqa: E402
from uno_driver_util import run_driver  # noqa: E402

build_ref.build()
golden = {"_how": "uno_driver <model> preset=ipopt linear_solver=ORACLE (unchanged Uno 2.0.0 host code + oracle/kkt_oracle.cpp); "
                  "hs015 / polak5 = examples/hs015.mod, examples/polak5.mod with hand-written derivatives (no ASL here)"}
for model in ("hs015", "polak5", "control2d:16", "control3d:6"):
    golden[model] = run_driver(model, "ORACLE")
    print(model, golden[model]["iterations"], golden[model]["iterate_status"], golden[model]["objective"])
(HERE / "uno_examples.json").write_text(json.dumps(golden, indent=1) + "\n")
