"""Recipe that places the UNMODIFIED reference implementation of the hot path under oracle/_ref/ -- TEST INFRASTRUCTURE.

The reference path is pure Python: the vendored `segment_anything` package
(/root/reference/InkLayer/third_party/segment-anything/segment_anything: build_sam.py, predictor.py, modeling/*.py,
utils/transforms.py).  There is nothing to compile; "building" the reference means making that package importable on the
GPU box, where /root/reference does not exist.  This script copies the package directory byte for byte into
oracle/_ref/segment_anything (git-ignored, so no reference source ever enters the history; not gpurun-ignored, so it
travels with the snapshot like libinksam.so does) and records the sha256 of every file in oracle/_ref/MANIFEST.json.

    python oracle/install_ref.py          # no-op with a message when /root/reference is absent

`__graft_entry__.build()` runs it when the reference tree is present.  `oracle/ref_loader.py` imports the copy;
`bench.py --impl reference` / `cpu_baseline` time it (`kind: "reference"`) and fall back to the port
(oracle/sam_oracle.py, `kind: "port"`) only when the copy is absent.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/InkLayer/third_party/segment-anything/segment_anything"
DST = os.path.join(HERE, "_ref", "segment_anything")


def install(verbose: bool = True) -> bool:
    if not os.path.isdir(SRC):
        if verbose:
            print(f"install_ref: {SRC} not present (GPU box?) -- keeping whatever oracle/_ref already holds")
        return os.path.isdir(DST)
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    shutil.copytree(SRC, DST, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    manifest = {}
    for root, _, files in os.walk(DST):
        for f in sorted(files):
            p = os.path.join(root, f)
            manifest[os.path.relpath(p, DST)] = hashlib.sha256(open(p, "rb").read()).hexdigest()
    with open(os.path.join(HERE, "_ref", "MANIFEST.json"), "w") as fh:
        json.dump({"source": SRC, "files": manifest}, fh, indent=1, sort_keys=True)
    if verbose:
        print(f"install_ref: copied {len(manifest)} files of the vendored segment_anything to {DST}")
    return True


if __name__ == "__main__":
    sys.exit(0 if install() else 1)
