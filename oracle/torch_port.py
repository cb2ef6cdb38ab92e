"""CPU port of the reference CorrBlock that issues the same torch library calls.

TEST INFRASTRUCTURE ONLY (see oracle/corr_oracle.py for the rule).  This port
exists for one purpose: the timed CPU baseline.  The reference's default path
spends its time inside torch.matmul, F.avg_pool2d and F.grid_sample
(core/corr.py:58, :26; core/utils/utils.py:65); timing those same calls with
all host threads is the honest "reference on the host cores" number when
/root/reference itself is not on the machine (it is absent on the GPU box).
It is also a second, independent checker for tests that want torch tensors.

The structure is deliberately not the reference's: one function per stage,
precomputed window offsets, no class.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def build_pyramid(fmap1: torch.Tensor, fmap2: torch.Tensor, num_levels: int = 4):
    """core/corr.py:13-27 and :52-60 -> list of [B*N,1,H_l,W_l]."""
    B, D, H, W = fmap1.shape
    N = H * W
    vol = torch.bmm(fmap1.reshape(B, D, N).transpose(1, 2), fmap2.reshape(B, D, N))
    vol = vol / torch.sqrt(torch.tensor(D).float())
    level = vol.reshape(B * N, 1, H, W)
    out = [level]
    for _ in range(1, num_levels):
        level = F.avg_pool2d(level, 2, stride=2)
        out.append(level)
    return out


def _window_offsets(radius: int, device, dtype):
    n = 2 * radius + 1
    d = torch.linspace(-radius, radius, n, device=device, dtype=dtype)
    # [n, n, 2]: entry [i, j] = (d[i], d[j]) is added to (x, y)  (core/corr.py:37-43)
    return torch.stack(torch.meshgrid(d, d, indexing="ij"), dim=-1)


def lookup(pyramid, coords: torch.Tensor, radius: int = 4) -> torch.Tensor:
    """core/corr.py:29-50 with core/utils/utils.py:57-65 inlined."""
    B, _, H, W = coords.shape
    n = 2 * radius + 1
    centre = coords.permute(0, 2, 3, 1).reshape(B * H * W, 1, 1, 2)
    offs = _window_offsets(radius, coords.device, coords.dtype).view(1, n, n, 2)
    feats = []
    for l, level in enumerate(pyramid):
        Hl, Wl = level.shape[-2:]
        pix = centre / 2 ** l + offs
        gx = 2 * pix[..., 0:1] / (Wl - 1) - 1
        gy = 2 * pix[..., 1:2] / (Hl - 1) - 1
        s = F.grid_sample(level, torch.cat([gx, gy], dim=-1), align_corners=True)
        feats.append(s.view(B, H, W, n * n))
    return torch.cat(feats, dim=-1).permute(0, 3, 1, 2).contiguous().float()


def hot_path(fmap1, fmap2, coords_seq, radius: int = 4, num_levels: int = 4):
    """One build followed by one lookup per entry of coords_seq."""
    pyr = build_pyramid(fmap1, fmap2, num_levels)
    return [lookup(pyr, c, radius) for c in coords_seq]
