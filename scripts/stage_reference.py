"""Stages the reference's two ENTRY POINTS, byte for byte, under baseline/_ref/ (git-ignored, travels to the
GPU box with the snapshot) so that `tests/test_gpu_dropin.py` can run them UNMODIFIED against this repo's
`nerf/` package on a B200: `/root/reference` does not exist there and this container has no GPU.

  baseline/_ref/train_conditional_nerf.py   copy of /root/reference/train_conditional_nerf.py
  baseline/_ref/example_cells.json          the code cells of examples/example.ipynb (sources only, no outputs)

Nothing under baseline/_ref is product source or committed; the tests pin the SHA-256 of what they run.
Run by `__graft_entry__.build()` whenever /root/reference is present."""
import hashlib
import json
import os
import shutil

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
OUT = os.path.join(ROOT, "baseline", "_ref")


def stage():
    if not os.path.isdir(REF):
        return None
    os.makedirs(OUT, exist_ok=True)
    shutil.copyfile(os.path.join(REF, "train_conditional_nerf.py"), os.path.join(OUT, "train_conditional_nerf.py"))
    nb = json.load(open(os.path.join(REF, "examples", "example.ipynb")))
    cells = ["".join(c["source"]) for c in nb["cells"] if c["cell_type"] == "code"]
    with open(os.path.join(OUT, "example_cells.json"), "w") as f:
        json.dump(cells, f)
    return {"train_conditional_nerf.py": hashlib.sha256(open(os.path.join(OUT, "train_conditional_nerf.py"), "rb").read()).hexdigest(),
            "cells": [hashlib.sha256(c.encode()).hexdigest() for c in cells]}


if __name__ == "__main__":
    print(json.dumps(stage(), indent=1))
