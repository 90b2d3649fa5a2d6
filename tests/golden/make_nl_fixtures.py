"""Copies the two bundled example problems of the reference (examples/hs015.nl, examples/polak5.nl - problem DATA in AMPL's text
.nl format, BASELINE.json configs[0]) into tests/golden/nl/ so that the GPU box, which has no /root/reference, can run them.
tests/golden/nl/ops.nl is written by hand in this repo (operator coverage of uno_plugin/models/NLModel)."""
import shutil
from pathlib import Path

REF = Path("/root/reference/examples")
OUT = Path(__file__).resolve().parent / "nl"
OUT.mkdir(exist_ok=True)
for name in ("hs015.nl", "polak5.nl"):
    shutil.copyfile(REF / name, OUT / name)
    print("copied", name)
