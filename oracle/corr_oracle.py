"""CPU oracle for RAFT's correlation hot path (numpy, fp64-capable).

TEST INFRASTRUCTURE ONLY.  Nothing in the product path (raft_stir_b200/) may
import this file; only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference leg use it, and there only as the checker.

It restates, in plain numpy with explicit index arithmetic, what the reference
obtains from three torch library calls (matmul, avg_pool2d, grid_sample) and
from its optional CUDA extension.  Each function cites the reference lines it
follows (paths relative to the reference checkout).

Parity status: the reference ships no golden vectors or tests for this path
(SURVEY.md section 8c), so the oracle is pinned against outputs of the
reference itself: oracle/make_golden.py imports the reference's core/corr.py
in the build container and freezes seeded input/output tensors (values and
autograd gradients) under tests/golden/; tests/test_oracle_golden.py checks
this file against them.
"""
from __future__ import annotations

import math

import numpy as np

__all__ = [
    "pyramid_dims", "corr_volume", "corr_pyramid", "window_lookup",
    "window_lookup_backward", "pyramid_backward", "alt_forward", "alt_backward",
    "coords_grid", "alt_block_lookup",
]


def pyramid_dims(H: int, W: int, num_levels: int = 4):
    """(H_l, W_l) per level: avg_pool2d(2, stride=2) floors odd sizes at every
    level (core/corr.py:25-27)."""
    return [(H >> l, W >> l) for l in range(num_levels)]


def coords_grid(B: int, H: int, W: int, dtype=np.float32):
    """[B,2,H,W] with channel 0 = x, channel 1 = y (core/utils/utils.py:74-77)."""
    ys, xs = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
    g = np.stack([xs, ys], axis=0).astype(dtype)
    return np.broadcast_to(g[None], (B, 2, H, W)).copy()


def corr_volume(fmap1: np.ndarray, fmap2: np.ndarray) -> np.ndarray:
    """All-pairs volume C[b,q,t] = <f1[b,:,q], f2[b,:,t]> / sqrt(D), returned as
    [B*N, H, W] (core/corr.py:52-60 followed by the reshape at :21-22)."""
    B, D, H, W = fmap1.shape
    a = fmap1.reshape(B, D, H * W)
    b = fmap2.reshape(B, D, H * W)
    vol = np.einsum("bdq,bdt->bqt", a, b, optimize=True)
    vol = vol / np.sqrt(np.asarray(D, dtype=np.float32)).astype(vol.dtype)
    return vol.reshape(B * H * W, H, W)


def _pool2(x: np.ndarray) -> np.ndarray:
    """2x2 mean, stride 2, trailing odd row/column dropped (F.avg_pool2d
    semantics used at core/corr.py:26 and :70-71).  x: [..., H, W]."""
    H, W = x.shape[-2:]
    h, w = H // 2, W // 2
    x = x[..., : 2 * h, : 2 * w]
    return 0.25 * (x[..., 0::2, 0::2] + x[..., 0::2, 1::2]
                   + x[..., 1::2, 0::2] + x[..., 1::2, 1::2])


def corr_pyramid(fmap1, fmap2, num_levels: int = 4):
    """List of num_levels arrays [B*N, H_l, W_l] (core/corr.py:13-27)."""
    pyr = [corr_volume(fmap1, fmap2)]
    for _ in range(num_levels - 1):
        pyr.append(_pool2(pyr[-1]))
    return pyr


def _taps(img, X, Y):
    """Zero-padded bilinear sampling in pixel coordinates.

    img: [M, Hl, Wl]; X, Y: [M, K] -> [M, K].  This is what
    bilinear_sampler (core/utils/utils.py:57-65) computes: it maps pixel
    coordinates to [-1,1] and calls grid_sample(align_corners=True), which maps
    them straight back, with the default zeros padding."""
    M, Hl, Wl = img.shape
    x0 = np.floor(X)
    y0 = np.floor(Y)
    ax = (X - x0).astype(img.dtype)
    ay = (Y - y0).astype(img.dtype)
    x0 = x0.astype(np.int64)
    y0 = y0.astype(np.int64)
    rows = np.arange(M)[:, None]

    def fetch(yy, xx):
        ok = (yy >= 0) & (yy < Hl) & (xx >= 0) & (xx < Wl)
        v = img[rows, np.clip(yy, 0, Hl - 1), np.clip(xx, 0, Wl - 1)]
        return np.where(ok, v, 0)

    v00 = fetch(y0, x0)
    v01 = fetch(y0, x0 + 1)
    v10 = fetch(y0 + 1, x0)
    v11 = fetch(y0 + 1, x0 + 1)
    return ((1 - ay) * ((1 - ax) * v00 + ax * v01)
            + ay * ((1 - ax) * v10 + ax * v11))


def window_lookup(pyramid, coords: np.ndarray, radius: int = 4) -> np.ndarray:
    """CorrBlock.__call__ (core/corr.py:29-50).

    coords: [B,2,H,W] (x,y).  Returns [B, L*(2r+1)^2, H, W].  Channel
    l*(2r+1)^2 + i*(2r+1) + j samples level l at
    (x/2^l + (i-r), y/2^l + (j-r)): the slow window index moves x, the fast one
    moves y (consequence of meshgrid(dy,dx) 'ij' added to an (x,y) centroid,
    core/corr.py:37-43)."""
    B, _, H, W = coords.shape
    r = radius
    n = 2 * r + 1
    cx = coords[:, 0].reshape(-1, 1)
    cy = coords[:, 1].reshape(-1, 1)
    d = np.arange(-r, r + 1, dtype=coords.dtype)
    off_x = np.repeat(d, n)[None, :]      # i (slow) -> x
    off_y = np.tile(d, n)[None, :]        # j (fast) -> y
    outs = []
    for l, img in enumerate(pyramid):
        s = coords.dtype.type(2 ** l)
        outs.append(_taps(img, cx / s + off_x, cy / s + off_y))
    out = np.concatenate(outs, axis=1)                      # [B*N, L*n*n]
    return out.reshape(B, H, W, -1).transpose(0, 3, 1, 2).astype(np.float32)


def window_lookup_backward(pyramid, coords, radius, grad_out):
    """Gradients of window_lookup as torch autograd yields them for
    core/corr.py:29-50: d/dpyramid[l] (same shapes as pyramid) and d/dcoords
    [B,2,H,W] (picks up 1/2^l per level)."""
    B, _, H, W = coords.shape
    r = radius
    n = 2 * r + 1
    M = B * H * W
    g = grad_out.transpose(0, 2, 3, 1).reshape(M, len(pyramid), n * n)
    cx = coords[:, 0].reshape(-1, 1).astype(np.float64)
    cy = coords[:, 1].reshape(-1, 1).astype(np.float64)
    d = np.arange(-r, r + 1, dtype=np.float64)
    off_x = np.repeat(d, n)[None, :]
    off_y = np.tile(d, n)[None, :]
    rows = np.broadcast_to(np.arange(M)[:, None], (M, n * n))
    dpyr = []
    dcx = np.zeros((M,), np.float64)
    dcy = np.zeros((M,), np.float64)
    for l, img in enumerate(pyramid):
        _, Hl, Wl = img.shape
        s = float(2 ** l)
        # the reference divides in fp32 (core/corr.py:41); do the same so that
        # floor() lands on the same cell
        X = (coords[:, 0].reshape(-1, 1) / coords.dtype.type(s)).astype(np.float64) + off_x
        Y = (coords[:, 1].reshape(-1, 1) / coords.dtype.type(s)).astype(np.float64) + off_y
        x0 = np.floor(X).astype(np.int64)
        y0 = np.floor(Y).astype(np.int64)
        ax = X - x0
        ay = Y - y0
        gl = g[:, l, :].astype(np.float64)
        dimg = np.zeros(img.shape, np.float64)
        vals = {}
        for dy_, wy in ((0, 1 - ay), (1, ay)):
            for dx_, wx in ((0, 1 - ax), (1, ax)):
                yy = y0 + dy_
                xx = x0 + dx_
                ok = (yy >= 0) & (yy < Hl) & (xx >= 0) & (xx < Wl)
                yc = np.clip(yy, 0, Hl - 1)
                xc = np.clip(xx, 0, Wl - 1)
                np.add.at(dimg, (rows[ok], yc[ok], xc[ok]), (gl * wy * wx)[ok])
                vals[(dy_, dx_)] = np.where(ok, img[rows, yc, xc], 0).astype(np.float64)
        ddx = ((1 - ay) * (vals[(0, 1)] - vals[(0, 0)]) + ay * (vals[(1, 1)] - vals[(1, 0)]))
        ddy = ((1 - ax) * (vals[(1, 0)] - vals[(0, 0)]) + ax * (vals[(1, 1)] - vals[(0, 1)]))
        dcx += (gl * ddx).sum(1) / s
        dcy += (gl * ddy).sum(1) / s
        dpyr.append(dimg)
    dcoords = np.stack([dcx.reshape(B, H, W), dcy.reshape(B, H, W)], axis=1)
    return dpyr, dcoords


def pyramid_backward(dpyr, fmap1, fmap2):
    """Back-propagate pyramid gradients to the feature maps: transpose of
    avg_pool2d at each level, then the two matmul gradients of
    core/corr.py:58-60.  Returns (dfmap1, dfmap2), each [B,D,H,W]."""
    B, D, H, W = fmap1.shape
    N = H * W
    dvol = np.zeros((B * N, H, W), np.float64)
    for l, gl in enumerate(dpyr):
        k = 2 ** l
        Hl, Wl = gl.shape[-2:]
        up = np.repeat(np.repeat(gl, k, axis=-2), k, axis=-1) / float(k * k)
        dvol[:, : Hl * k, : Wl * k] += up
    dvol = dvol.reshape(B, N, N) / math.sqrt(D)
    a = fmap1.reshape(B, D, N).astype(np.float64)
    b = fmap2.reshape(B, D, N).astype(np.float64)
    d1 = np.einsum("bqt,bdt->bdq", dvol, b, optimize=True)
    d2 = np.einsum("bqt,bdq->bdt", dvol, a, optimize=True)
    return d1.reshape(B, D, H, W), d2.reshape(B, D, H, W)


def alt_forward(fmap1, fmap2, coords, radius):
    """alt_cuda_corr.forward (alt_cuda_corr/correlation_kernel.cu:18-119,
    260-286) restated without its tiling.

    fmap1: [B,H1,W1,C]; fmap2: [B,H2,W2,C]; coords: [B,Nc,H1,W1,2] (x,y) already
    in fmap2's pixel units.  Returns corr [B,Nc,(2r+1)^2,H1,W1], NOT scaled by
    1/sqrt(C) (the caller does that, core/corr.py:91).

    The kernel visits the (2r+2)^2 integer taps around floor(coords), takes the
    feature dot product at each in-bounds tap (:76-90) and splats it with the
    four bilinear weights into output index iy + rd*ix (:92-114); that equals
    sampling the dot-product image at (x-r+ix, y-r+iy) with zero padding, which
    is how it is written here."""
    B, H1, W1, C = fmap1.shape
    _, H2, W2, _ = fmap2.shape
    Nc = coords.shape[1]
    r = radius
    rd = 2 * r + 1
    out = np.zeros((B, Nc, rd * rd, H1, W1), np.float64)
    f1 = fmap1.astype(np.float64)
    f2 = fmap2.astype(np.float64)
    for b in range(B):
        dots = np.einsum("yxc,vuc->yxvu", f1[b], f2[b], optimize=True)   # [H1,W1,H2,W2]
        img = dots.reshape(H1 * W1, H2, W2)
        for n in range(Nc):
            x = coords[b, n, :, :, 0].reshape(-1, 1).astype(np.float64)
            y = coords[b, n, :, :, 1].reshape(-1, 1).astype(np.float64)
            d = np.arange(-r, r + 1, dtype=np.float64)
            off_x = np.repeat(d, rd)[None, :]     # ix slow
            off_y = np.tile(d, rd)[None, :]       # iy fast
            v = _taps(img, x + off_x, y + off_y)  # [H1*W1, rd*rd]
            out[b, n] = v.T.reshape(rd * rd, H1, W1)
    return out.astype(np.float32)


def alt_backward(fmap1, fmap2, coords, corr_grad, radius):
    """alt_cuda_corr.backward (correlation_kernel.cu:122-256, 288-324).

    Returns (fmap1_grad [B,H1,W1,C], fmap2_grad [B,H2,W2,C], coords_grad
    [B,Nc,H1,W1,2]).  coords_grad is all zeros: the reference allocates it
    (:307) and its kernel never writes it."""
    B, H1, W1, C = fmap1.shape
    _, H2, W2, _ = fmap2.shape
    Nc = coords.shape[1]
    r = radius
    rd = 2 * r + 1
    f1 = fmap1.astype(np.float64)
    f2 = fmap2.astype(np.float64)
    g1 = np.zeros_like(f1)
    g2 = np.zeros_like(f2)
    M = H1 * W1
    rows = np.broadcast_to(np.arange(M)[:, None], (M, rd * rd))
    for b in range(B):
        ddots = np.zeros((M, H2, W2), np.float64)
        for n in range(Nc):
            x = coords[b, n, :, :, 0].reshape(-1, 1).astype(np.float64)
            y = coords[b, n, :, :, 1].reshape(-1, 1).astype(np.float64)
            d = np.arange(-r, r + 1, dtype=np.float64)
            X = x + np.repeat(d, rd)[None, :]
            Y = y + np.tile(d, rd)[None, :]
            x0 = np.floor(X).astype(np.int64)
            y0 = np.floor(Y).astype(np.int64)
            ax = X - x0
            ay = Y - y0
            g = corr_grad[b, n].reshape(rd * rd, M).T.astype(np.float64)
            for dy_, wy in ((0, 1 - ay), (1, ay)):
                for dx_, wx in ((0, 1 - ax), (1, ax)):
                    yy = y0 + dy_
                    xx = x0 + dx_
                    ok = (yy >= 0) & (yy < H2) & (xx >= 0) & (xx < W2)
                    np.add.at(ddots, (rows[ok], yy[ok], xx[ok]), (g * wy * wx)[ok])
        dd = ddots.reshape(H1, W1, H2, W2)
        g1[b] = np.einsum("yxvu,vuc->yxc", dd, f2[b], optimize=True)
        g2[b] = np.einsum("yxvu,yxc->vuc", dd, f1[b], optimize=True)
    cg = np.zeros((B, Nc, H1, W1, 2), np.float32)
    return g1.astype(np.float32), g2.astype(np.float32), cg


def alt_block_lookup(fmap1, fmap2, coords, num_levels=4, radius=4):
    """AlternateCorrBlock end to end (core/corr.py:63-91): pool the features,
    call the native op per level on coords/2^l, stack, scale by 1/sqrt(D).
    fmap1, fmap2: [B,D,H,W]; coords [B,2,H,W] -> [B, L*(2r+1)^2, H, W]."""
    B, D, H, W = fmap1.shape
    f1 = np.ascontiguousarray(fmap1.transpose(0, 2, 3, 1))
    outs = []
    f2l = fmap2
    c = coords.transpose(0, 2, 3, 1)
    for l in range(num_levels):
        f2 = np.ascontiguousarray(f2l.transpose(0, 2, 3, 1))
        ci = (c / coords.dtype.type(2 ** l)).reshape(B, 1, H, W, 2)
        outs.append(alt_forward(f1, f2, ci, radius)[:, 0])
        f2l = _pool2(f2l)
    out = np.stack(outs, axis=1).reshape(B, -1, H, W)
    return (out / np.sqrt(np.float32(D))).astype(np.float32)
