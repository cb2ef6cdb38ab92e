"""CPU oracle for RAFT's flow upsampling (numpy, fp64-capable).

TEST INFRASTRUCTURE ONLY, like corr_oracle.py: nothing under raft_stir_b200/ may import this file.

Restates with explicit index arithmetic what the reference obtains from torch.softmax + F.unfold +
a broadcast multiply-sum (core/raft.py:72-83, `RAFT.upsample_flow`) and from F.interpolate
(core/utils/utils.py:80-82, `upflow8`).  Pinned against the reference itself:
oracle/make_golden_upsample.py freezes outputs and autograd gradients of those two functions under
tests/golden/upsample.npz; tests/test_oracle_golden.py checks this file against them.
"""
from __future__ import annotations

import numpy as np

__all__ = ["upsample_flow", "upsample_flow_backward", "upflow8", "upflow8_backward"]


def _softmax9(mask: np.ndarray) -> np.ndarray:
    """mask [B, 576, H, W] -> weights [B, 9, 8, 8, H, W], softmax over the 9 neighbours
    (core/raft.py:75-76: channel = k*64 + sy*8 + sx)."""
    B, C, H, W = mask.shape
    assert C == 576
    m = mask.reshape(B, 9, 8, 8, H, W)
    m = m - m.max(axis=1, keepdims=True)
    e = np.exp(m)
    return e / e.sum(axis=1, keepdims=True)


def _neighbours(flow: np.ndarray) -> np.ndarray:
    """8*flow [B, 2, H, W] -> [B, 2, 9, H, W]: neighbour k = ky*3 + kx is the value at (y+ky-1, x+kx-1), zero
    outside the map (core/raft.py:78-79: F.unfold(8 * flow, [3, 3], padding=1))."""
    B, _, H, W = flow.shape
    pad = np.zeros((B, 2, H + 2, W + 2), dtype=flow.dtype)
    pad[:, :, 1:-1, 1:-1] = 8 * flow
    out = np.empty((B, 2, 9, H, W), dtype=flow.dtype)
    for ky in range(3):
        for kx in range(3):
            out[:, :, ky * 3 + kx] = pad[:, :, ky:ky + H, kx:kx + W]
    return out


def upsample_flow(flow: np.ndarray, mask: np.ndarray) -> np.ndarray:
    """core/raft.py:72-83.  out[b, c, 8y+sy, 8x+sx] = sum_k w[b, k, sy, sx, y, x] * 8 * flow[b, c, y+ky-1, x+kx-1]."""
    B, _, H, W = flow.shape
    w = _softmax9(mask)                                  # [B, 9, 8, 8, H, W]
    nb = _neighbours(flow)                               # [B, 2, 9, H, W]
    up = np.einsum("bkstyx,bckyx->bcysxt", w, nb)        # [B, 2, H, 8, W, 8]  (s = sy, t = sx)
    return np.ascontiguousarray(up).reshape(B, 2, 8 * H, 8 * W)


def upsample_flow_backward(flow: np.ndarray, mask: np.ndarray, gout: np.ndarray):
    """Gradients of upsample_flow w.r.t. flow and mask for an output gradient gout [B, 2, 8H, 8W]
    (what autograd derives for core/raft.py:72-83)."""
    B, _, H, W = flow.shape
    w = _softmax9(mask)
    nb = _neighbours(flow)
    g = gout.reshape(B, 2, H, 8, W, 8)                   # [b, c, y, sy, x, sx]
    dw = np.einsum("bcysxt,bckyx->bkstyx", g, nb)        # d/dw
    dmask = w * (dw - (w * dw).sum(axis=1, keepdims=True))
    dnb = np.einsum("bcysxt,bkstyx->bckyx", g, w)        # d/d(neighbour values)
    dpad = np.zeros((B, 2, H + 2, W + 2), dtype=flow.dtype)
    for ky in range(3):
        for kx in range(3):
            dpad[:, :, ky:ky + H, kx:kx + W] += dnb[:, :, ky * 3 + kx]
    dflow = 8 * dpad[:, :, 1:-1, 1:-1]
    return dflow, dmask.reshape(B, 576, H, W)


def _lerp_axis(n_in: int, n_out: int):
    """align_corners=True source positions (core/utils/utils.py:82): i0, i1, fraction for every output index."""
    if n_out == 1 or n_in == 1:
        pos = np.zeros(n_out)
    else:
        pos = np.arange(n_out) * ((n_in - 1) / (n_out - 1))
    i0 = np.minimum(np.floor(pos).astype(np.int64), n_in - 1)
    i1 = np.minimum(i0 + 1, n_in - 1)
    return i0, i1, pos - i0


def upflow8(flow: np.ndarray) -> np.ndarray:
    """core/utils/utils.py:80-82: 8 * bilinear interpolation to (8H, 8W), align_corners=True."""
    B, C, H, W = flow.shape
    y0, y1, fy = _lerp_axis(H, 8 * H)
    x0, x1, fx = _lerp_axis(W, 8 * W)
    fy = fy.astype(flow.dtype)[None, None, :, None]
    fx = fx.astype(flow.dtype)[None, None, None, :]
    rows = flow[:, :, y0] * (1 - fy) + flow[:, :, y1] * fy           # [B, C, 8H, W]
    return 8 * (rows[:, :, :, x0] * (1 - fx) + rows[:, :, :, x1] * fx)


def upflow8_backward(shape, gout: np.ndarray) -> np.ndarray:
    """Gradient of upflow8 w.r.t. flow (shape [B, C, H, W]) for gout [B, C, 8H, 8W]."""
    B, C, H, W = shape
    y0, y1, fy = _lerp_axis(H, 8 * H)
    x0, x1, fx = _lerp_axis(W, 8 * W)
    g = 8 * gout
    grow = np.zeros((B, C, 8 * H, W), dtype=gout.dtype)
    np.add.at(grow, (slice(None), slice(None), slice(None), x0), g * (1 - fx).astype(g.dtype))
    np.add.at(grow, (slice(None), slice(None), slice(None), x1), g * fx.astype(g.dtype))
    dflow = np.zeros((B, C, H, W), dtype=gout.dtype)
    np.add.at(dflow, (slice(None), slice(None), y0), grow * (1 - fy).astype(g.dtype)[None, None, :, None])
    np.add.at(dflow, (slice(None), slice(None), y1), grow * fy.astype(g.dtype)[None, None, :, None])
    return dflow
