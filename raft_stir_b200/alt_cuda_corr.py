"""Module-level stand-in for the reference's optional `alt_cuda_corr` extension
(alt_cuda_corr/correlation.cpp:51-54): `forward` and `backward` with the same
argument order, shapes, return lists and error type (RuntimeError).

    sys.modules["alt_cuda_corr"] = raft_stir_b200.alt_cuda_corr

lets the reference's own AlternateCorrBlock (core/corr.py:5-9, :86) run on it.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import RaftCorrError, check
from .corr import _guard


def _check_input(t, name):
    # CHECK_CUDA / CHECK_CONTIGUOUS of correlation.cpp:19-21
    if not t.is_cuda:
        raise RaftCorrError(f"{name} must be a CUDA tensor")
    if not t.is_contiguous():
        raise RaftCorrError(f"{name} must be contiguous")
    if t.dtype != torch.float32:
        raise RaftCorrError(f"{name} must be float32")


def _dims(fmap1, fmap2, coords):
    if fmap1.dim() != 4 or fmap2.dim() != 4 or coords.dim() != 5 or coords.shape[-1] != 2:
        raise RaftCorrError("expected fmap1 [B,H1,W1,C], fmap2 [B,H2,W2,C], coords [B,N,H1,W1,2]")
    B, H1, W1, Cc = (int(v) for v in fmap1.shape)
    B2, H2, W2, C2 = (int(v) for v in fmap2.shape)
    Bc, Nc, Hc, Wc, _ = (int(v) for v in coords.shape)
    if B2 != B or C2 != Cc or Bc != B or Hc != H1 or Wc != W1:
        raise RaftCorrError("fmap1 / fmap2 / coords shapes disagree")
    return B, H1, W1, H2, W2, Cc, Nc


def forward(fmap1, fmap2, coords, radius):
    """-> [corr [B, N, (2r+1)^2, H1, W1]]; not scaled by 1/sqrt(C) (core/corr.py:91 does that)."""
    for t, n in ((fmap1, "fmap1"), (fmap2, "fmap2"), (coords, "coords")):
        _check_input(t, n)
    B, H1, W1, H2, W2, Cc, Nc = _dims(fmap1, fmap2, coords)
    rd = 2 * int(radius) + 1
    corr = torch.empty(B, Nc, rd * rd, H1, W1, dtype=torch.float32, device=fmap1.device)
    with _guard(fmap1.device):
        check(_lib.lib().rc_alt_fwd(fmap1.data_ptr(), fmap2.data_ptr(), coords.data_ptr(), B, H1, W1, H2, W2,
                                    Cc, Nc, int(radius), corr.data_ptr(),
                                    torch.cuda.current_stream(fmap1.device).cuda_stream), "rc_alt_fwd")
    return [corr]


def backward(fmap1, fmap2, coords, corr_grad, radius):
    """-> [fmap1_grad, fmap2_grad, coords_grad]; coords_grad is zeros, as in the
    reference (correlation_kernel.cu:307: allocated, never written)."""
    for t, n in ((fmap1, "fmap1"), (fmap2, "fmap2"), (coords, "coords"), (corr_grad, "corr_grad")):
        _check_input(t, n)
    B, H1, W1, H2, W2, Cc, Nc = _dims(fmap1, fmap2, coords)
    g1 = torch.empty_like(fmap1)
    g2 = torch.empty_like(fmap2)
    gc = torch.empty_like(coords)
    with _guard(fmap1.device):
        check(_lib.lib().rc_alt_bwd(fmap1.data_ptr(), fmap2.data_ptr(), coords.data_ptr(), corr_grad.data_ptr(),
                                    B, H1, W1, H2, W2, Cc, Nc, int(radius), g1.data_ptr(), g2.data_ptr(),
                                    gc.data_ptr(), torch.cuda.current_stream(fmap1.device).cuda_stream),
              "rc_alt_bwd")
    return [g1, g2, gc]
