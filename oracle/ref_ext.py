"""Loader of the reference's own alt_cuda_corr extension, compiled unmodified by
oracle/build_ref_ext.sh into oracle/_ref/ (test infrastructure: checker and timed reference arm,
never a code path of the library).  CUDA-only, like the reference: usable on the GPU box."""
from __future__ import annotations

import glob
import importlib.util
import os

HERE = os.path.dirname(os.path.abspath(__file__))
_mod = None


def path():
    c = sorted(glob.glob(os.path.join(HERE, "_ref", "alt_cuda_corr_ref.cpython-*.so")))
    return c[0] if c else None


def available() -> bool:
    return path() is not None


def load():
    """module with forward(fmap1, fmap2, coords, radius) / backward(..., corr_grad, radius)
    (alt_cuda_corr/correlation.cpp:51-54)"""
    global _mod
    if _mod is None:
        import torch  # noqa: F401  (the extension links against libtorch)
        spec = importlib.util.spec_from_file_location("alt_cuda_corr_ref", path())
        _mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(_mod)
    return _mod
