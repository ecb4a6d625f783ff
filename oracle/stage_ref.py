"""Stage the UNMODIFIED reference model file for the reference arm.  TEST / BASELINE INFRASTRUCTURE ONLY.

    python oracle/stage_ref.py        (also run by __graft_entry__.build() when /root/reference exists)

The reference is a flat folder of Python scripts with no package or build system; the only file of it that the hot
path needs is ``bsrnn.py`` (it imports nothing but torch).  ``/root/reference`` does not exist on the GPU box, so the
file is copied byte for byte to ``oracle/_ref/bsrnn.py`` -- a git-ignored, gpurun-shipped build output, exactly like
the compiled ``oracle/_ref`` library of a C reference would be -- and never into the tracked tree.  ``bench.py --impl
reference`` and its ``cpu_baseline`` / ``gpu_comparator`` legs import it from there (``load_reference_module``) and
report ``kind: "reference"``; when it is absent they fall back to the oracle port (``kind: "port"``).
The callers around the model (m_dataset.py, infer*.py) cannot be staged: they need torchmetrics / soundfile /
torchaudio.io, which are not installed; their ten wrap lines are restated in oracle/bsrnn_oracle.py.
"""
from __future__ import annotations

import hashlib
import importlib.util
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/bsrnn.py"
DST_DIR = os.path.join(HERE, "_ref")
DST = os.path.join(DST_DIR, "bsrnn.py")


def stage() -> bool:
    """Copy the reference model file when the reference tree is present. True if oracle/_ref/bsrnn.py exists."""
    if os.path.exists(SRC):
        os.makedirs(DST_DIR, exist_ok=True)
        shutil.copyfile(SRC, DST)
        with open(os.path.join(DST_DIR, "STAGED.json"), "w") as f:
            json.dump({"source": SRC, "sha256": hashlib.sha256(open(SRC, "rb").read()).hexdigest(),
                       "note": "unmodified copy made by oracle/stage_ref.py; not tracked by git"}, f)
    return os.path.exists(DST)


def load_reference_module():
    """The reference's bsrnn module (from /root/reference if present, else the staged copy), or None.
    Loaded by file path under a private name: the repo root ships its own `bsrnn` drop-in shim."""
    path = SRC if os.path.exists(SRC) else DST
    if not os.path.exists(path):
        return None
    import contextlib
    import io
    spec = importlib.util.spec_from_file_location("reference_bsrnn", path)
    mod = importlib.util.module_from_spec(spec)
    with contextlib.redirect_stdout(io.StringIO()):
        spec.loader.exec_module(mod)
    return mod


def reference_model(seed: int = 0, device="cpu"):
    """(model, kind): the reference's own BSRNN with `torch.manual_seed(seed)` default init in eval mode on
    `device`, kind = "reference"; or (None, "port") when no reference file is available."""
    import contextlib
    import io
    import torch
    mod = load_reference_module()
    if mod is None:
        return None, "port"
    torch.manual_seed(seed)
    with contextlib.redirect_stdout(io.StringIO()):      # the ctor prints the band table (bsrnn.py:351)
        m = mod.BSRNN()
    return m.to(device).eval(), "reference"


if __name__ == "__main__":
    ok = stage()
    print("staged" if ok else "reference not present and nothing staged", DST)
    sys.exit(0 if ok else 1)
