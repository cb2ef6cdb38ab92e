#!/usr/bin/env python3
"""Stage the reference's own Python for the correlation path under oracle/_ref/refcore/ (test infrastructure).

    python oracle/stage_ref_python.py [--if-missing]

Runs only where the reference checkout is mounted (the build container).  It copies, byte for byte, the files
that hold the path -- core/corr.py (CorrBlock, AlternateCorrBlock) and core/utils/utils.py (bilinear_sampler,
coords_grid) plus the package marker -- into oracle/_ref/refcore/, which is git-ignored (never part of this
repository's history) and travels to the GPU box with gpurun like the other reference-derived checkers there.
bench.py's CPU arm (`--impl reference` and `cpu_baseline`) then times the reference's UNMODIFIED CorrBlock on the
box's host cores (kind "reference"); without the staged files it falls back to oracle/torch_port.py (kind "port").
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("RAFT_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref", "refcore")
FILES = ["corr.py", os.path.join("utils", "utils.py"), os.path.join("utils", "__init__.py")]


def main() -> int:
    if "--if-missing" in sys.argv and all(os.path.exists(os.path.join(OUT, f)) for f in FILES):
        return 0
    src = os.path.join(REF, "core")
    if not os.path.isdir(src):
        print(f"reference checkout not found at {REF}; nothing to do", file=sys.stderr)
        return 0
    for f in FILES:
        dst = os.path.join(OUT, f)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(src, f), dst)
    print(f"staged {len(FILES)} reference files under {OUT}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
