"""raft_stir_b200 -- RAFT's correlation hot path on B200 (sm_100a).

Drop-in replacements for the reference's `CorrBlock` / `AlternateCorrBlock`
(core/corr.py) and its optional `alt_cuda_corr` extension, backed by
hand-written CUDA kernels behind the C ABI declared in include/raft_corr.h.
There is no CPU path: constructing a block without the compiled library or
without a CUDA tensor raises.
"""
from .corr import (AlternateCorrBlock, CorrBlock, FusedConvc1, LazyCorr, coords_grid, fuse_motion_encoder,  # noqa: F401
                   install_into, pack_convc1_weights, pack_features)
from . import alt_cuda_corr  # noqa: F401
from .upsample import upflow8, upsample_flow  # noqa: F401
from .graph import GraphedCallable  # noqa: F401

__all__ = ["CorrBlock", "AlternateCorrBlock", "alt_cuda_corr", "coords_grid", "install_into", "pack_features", "GraphedCallable", "FusedConvc1", "LazyCorr",
           "fuse_motion_encoder", "pack_convc1_weights", "upsample_flow",
           "upflow8"]
