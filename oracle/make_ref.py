"""Recipe for oracle/_ref: the UNMODIFIED reference modules of the loss-head path.

TEST / BENCH INFRASTRUCTURE ONLY (see oracle/ref_loader.py for who may use the result).

The reference is pure Python, so "building" it is placing its own files where the GPU box can
import them: /root/reference does not exist there, oracle/_ref/ (git-ignored, NOT gpurun-ignored)
travels with the snapshot like the built .so files do.  Nothing under oracle/_ref/ is ever
committed, and nothing is edited: the files are byte-for-byte those of /root/reference (a
SHA-256 manifest is written next to them so that a run can say exactly what it timed).

    python oracle/make_ref.py            # idempotent; a no-op when /root/reference is absent

Files (reference path -> role):
    loss.py           MILNCELoss (loss.py:6-18) and the composite soft-DTW losses (:20-134)
    soft_dtw_cuda.py  SoftDTW with its numba-CPU and numba-CUDA paths
    dtw.py            imported by loss.py at module level
    utils.py          AllGather (utils.py:8-24)
    s3dg.py           the S3D-G towers for the BASELINE configs[4] harness
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.environ.get("B200HEAD_REFERENCE_DIR", "/root/reference")
REF_DST = os.path.join(HERE, "_ref")
FILES = ("loss.py", "soft_dtw_cuda.py", "dtw.py", "utils.py", "s3dg.py")


def make(verbose: bool = True) -> bool:
    """Returns True when oracle/_ref is complete afterwards."""
    if not os.path.isdir(REF_SRC):
        if verbose:
            print(f"make_ref: {REF_SRC} not present; leaving oracle/_ref as it is")
        return all(os.path.exists(os.path.join(REF_DST, f)) for f in FILES)
    os.makedirs(REF_DST, exist_ok=True)
    manifest = {}
    for f in FILES:
        src, dst = os.path.join(REF_SRC, f), os.path.join(REF_DST, f)
        shutil.copyfile(src, dst)
        manifest[f] = hashlib.sha256(open(dst, "rb").read()).hexdigest()
    with open(os.path.join(REF_DST, "MANIFEST.json"), "w") as fh:
        json.dump({"source": REF_SRC, "sha256": manifest}, fh, indent=1)
    if verbose:
        print(f"make_ref: {len(FILES)} reference modules placed under {REF_DST}")
    return True


if __name__ == "__main__":
    sys.exit(0 if make() else 1)
