"""Recipe for oracle/_ref: the reference's OWN implementation of the hot path, staged for the GPU box.

The reference is pure Python (no build system, nothing to compile): its hot path is `attention.py` + `decoder.py`.
`/root/reference` does not exist on the GPU box, and reference sources must never enter this repo's history, so this
recipe -- run by `__graft_entry__.build()` in the build container whenever /root/reference is present -- copies those two
files, byte for byte, into `oracle/_ref/` (git-ignored, NOT gpurun-ignored: it travels with the snapshot like the built
.so).  `oracle/ref_runner.py` then imports them from there; the only touch is the injection of the module global
`mps_device`, which upstream defines on Apple-Silicon only.

TEST INFRASTRUCTURE ONLY: `bench.py --impl reference`, bench.py's cpu_baseline leg and tests/ use it as the timed CPU
baseline / checker.  The product package never imports it.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.environ.get("SAT_REFERENCE_DIR", "/root/reference")
DST = os.path.join(HERE, "_ref")
FILES = ("attention.py", "decoder.py")


def build(verbose: bool = True) -> bool:
    """Stage the reference modules; returns True when oracle/_ref is usable afterwards."""
    if not all(os.path.exists(os.path.join(REF_SRC, f)) for f in FILES):
        ok = all(os.path.exists(os.path.join(DST, f)) for f in FILES)
        if verbose:
            print(f"[oracle/_ref] {REF_SRC} not present; using the staged copy" if ok else
                  f"[oracle/_ref] {REF_SRC} not present and nothing staged")
        return ok
    os.makedirs(DST, exist_ok=True)
    manifest = {"source": REF_SRC, "files": {}}
    for f in FILES:
        shutil.copyfile(os.path.join(REF_SRC, f), os.path.join(DST, f))
        with open(os.path.join(DST, f), "rb") as fh:
            manifest["files"][f] = hashlib.sha256(fh.read()).hexdigest()
    with open(os.path.join(DST, "MANIFEST.json"), "w") as fh:
        json.dump(manifest, fh, indent=1)
    if verbose:
        print(f"[oracle/_ref] staged {', '.join(FILES)} from {REF_SRC}")
    return True


if __name__ == "__main__":
    build()
