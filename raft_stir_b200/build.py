"""Compile libraftcorr.so (hand-written sm_100a CUDA behind a C ABI) in-tree.

    python -m raft_stir_b200.build [--force]

nvcc cross-compiles without a GPU.  The explicit -gencode form is required:
plain -arch=sm_100a also emits a compute_100 PTX pass that rejects tcgen05.
No torch headers are involved, so a full build takes seconds.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "_build")
LIB = os.path.join(HERE, "libraftcorr.so")
SOURCES = ["api.cu", "pack.cu", "build_simt.cu", "build_tc.cu", "build_bwd.cu", "lookup.cu", "convc1_tc.cu", "alt.cu", "alt_tc.cu", "upsample.cu"]
HEADERS = [os.path.join(CSRC, "common.cuh"), os.path.join(CSRC, "tc_ptx.cuh"), os.path.join(HERE, "..", "include", "raft_corr.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--expt-relaxed-constexpr",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libraftcorr cannot be built")


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, attrib: bool = False) -> str:
    """Build (if stale) and return the path of libraftcorr.so.

    attrib=True builds a second, separate library under _build/attrib/ with -DRC_TC_ATTRIB: the build kernel then
    obeys the RC_TC_DEBUG switches that turn parts of its epilogue off (timing / power attribution only, results are
    wrong by construction).  Load it with RAFTCORR_LIB=<path>; the shipped library contains none of it."""
    nvcc = _nvcc()
    obj_dir = os.path.join(OBJ, "attrib") if attrib else OBJ
    lib_path = os.path.join(obj_dir, "libraftcorr.so") if attrib else LIB
    extra = ["-DRC_TC_ATTRIB"] if attrib else []
    os.makedirs(obj_dir, exist_ok=True)
    jobs = []
    for src in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(obj_dir, src.replace(".cu", ".o"))
        if force or _stale(o, [s] + HEADERS):
            jobs.append((s, o))

    def compile_one(job):
        s, o = job
        cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", s, "-o", o]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {s}:\n{r.stdout}\n{r.stderr}")
        return s, r.stderr

    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for s, log in ex.map(compile_one, jobs):
                if verbose and log:
                    print(f"== {os.path.basename(s)}\n{log}", file=sys.stderr)
    objs = [os.path.join(obj_dir, s.replace(".cu", ".o")) for s in SOURCES]
    if force or jobs or _stale(lib_path, objs):
        cuda_lib = os.path.join(os.path.dirname(os.path.dirname(nvcc)), "lib64")
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", lib_path] + objs + [
            "-L" + cuda_lib, "-Xlinker", "-rpath=" + cuda_lib]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return lib_path


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, attrib="--attrib" in sys.argv))
