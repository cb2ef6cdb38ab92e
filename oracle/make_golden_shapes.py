"""Freeze a SUBSAMPLE of the reference CorrBlock's outputs at the three BASELINE feature shapes.

Run in the build container (where /root/reference is mounted):

    python oracle/make_golden_shapes.py

tests/golden/corrblock_{basic,small}.npz hold complete tensors of tiny cases.  The shapes BASELINE.json names --
Sintel 55x128, FlyingChairs 46x62, 1080p video 135x240, all D = 256, r = 4, 4 levels -- are too large to commit
whole (a 1080p volume is 4.2 GB), so this script runs the reference's own core/corr.py (unmodified, CPU, fp32) on
seeded inputs and keeps, per shape: the lookup output at 256 random queries for two coordinate fields (grid + N(0,5^2)
flow, and grid + integer flow) and, for the two training-size shapes, reference autograd's d/dcoords (whole) and
d/dfmap1, d/dfmap2 at 4096 random positions.  The inputs are NOT stored: make_inputs() below regenerates them from
the seed (same torch CPU generator on the same image); float64 checksums are stored so a test can tell if that ever
stops holding.  A few hundred KB per shape.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

REF = os.environ.get("RAFT_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")
SHAPES = {"sintel": (55, 128, True), "chairs": (46, 62, True), "video1080": (135, 240, False)}   # H, W, with gradients
D, RADIUS, LEVELS, NQ, NG = 256, 4, 4, 256, 4096


def make_inputs(name):
    """Seeded inputs of one shape (also imported by tests/test_gpu_golden_shapes.py)."""
    H, W, _ = SHAPES[name]
    g = torch.Generator().manual_seed(1000 + H * W)
    f1 = torch.randn(1, D, H, W, generator=g)
    f2 = torch.randn(1, D, H, W, generator=g)
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    grid = torch.stack([xs, ys], 0).float()[None]
    coords_rand = grid + 5.0 * torch.randn(1, 2, H, W, generator=g)
    coords_int = grid + torch.randint(-6, 7, (1, 2, H, W), generator=g).float()
    grad_out = torch.randn(1, LEVELS * (2 * RADIUS + 1) ** 2, H, W, generator=g)
    queries = torch.randperm(H * W, generator=g)[:NQ].sort().values
    gpos = torch.randperm(D * H * W, generator=g)[:NG].sort().values
    return dict(f1=f1, f2=f2, coords_rand=coords_rand, coords_int=coords_int, grad_out=grad_out, queries=queries, gpos=gpos)


def checksums(inp):
    return np.array([inp[k].double().sum().item() for k in ("f1", "f2", "coords_rand", "coords_int", "grad_out")])


def main():
    core = os.path.join(REF, "core")
    if not os.path.isdir(core):
        raise SystemExit(f"reference not found at {REF}")
    sys.path.insert(0, core)
    import warnings
    warnings.filterwarnings("ignore")
    import corr as ref_corr   # the reference's core/corr.py, unmodified
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    for name, (H, W, with_grad) in SHAPES.items():
        inp = make_inputs(name)
        q = inp["queries"]
        data = dict(checksums=checksums(inp), queries=q.numpy(), gpos=inp["gpos"].numpy(),
                    torch_version=np.array(torch.__version__))
        with torch.no_grad():
            blk = ref_corr.CorrBlock(inp["f1"], inp["f2"], num_levels=LEVELS, radius=RADIUS)
            for kind in ("rand", "int"):
                out = blk(inp[f"coords_{kind}"])                       # [1, 324, H, W]
                data[f"out_{kind}"] = out.reshape(out.shape[1], H * W)[:, q].numpy().copy()
            # a few complete level rows of the pyramid itself (query -> its maps at levels 1..3)
            for l in range(1, LEVELS):
                data[f"pyr{l}"] = blk.corr_pyramid[l][q[:16], 0].numpy().copy()
            del blk
        if with_grad:
            f1 = inp["f1"].clone().requires_grad_(True)
            f2 = inp["f2"].clone().requires_grad_(True)
            c = inp["coords_rand"].clone().requires_grad_(True)
            blk = ref_corr.CorrBlock(f1, f2, num_levels=LEVELS, radius=RADIUS)
            blk(c).backward(inp["grad_out"])
            data["dcoords"] = c.grad.numpy().copy()
            data["dfmap1"] = f1.grad.reshape(-1)[inp["gpos"]].numpy().copy()
            data["dfmap2"] = f2.grad.reshape(-1)[inp["gpos"]].numpy().copy()
            data["dfmap1_rms"] = np.float64(f1.grad.double().pow(2).mean().sqrt().item())
            data["dfmap2_rms"] = np.float64(f2.grad.double().pow(2).mean().sqrt().item())
            del blk
        path = os.path.join(OUT, f"shape_{name}.npz")
        np.savez_compressed(path, **data)
        print(f"wrote {path}: {os.path.getsize(path) / 1e3:.0f} KB")


if __name__ == "__main__":
    main()
