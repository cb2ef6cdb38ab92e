"""Freeze outputs of the reference's own CorrBlock into tests/golden/*.npz.

Run in the build container (where /root/reference is mounted):

    python oracle/make_golden.py

The reference (athaddius/RAFT_STIR) has no golden vectors of its own for the
correlation path, so this script imports core/corr.py unmodified, feeds it
seeded inputs and stores inputs, pyramid levels, lookup outputs and autograd
gradients.  The .npz files are committed; this script is what made them.
torch version used is stored in each file.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

REF = os.environ.get("RAFT_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def _import_reference():
    core = os.path.join(REF, "core")
    if not os.path.isdir(core):
        raise SystemExit(f"reference not found at {REF}")
    sys.path.insert(0, core)
    import corr as ref_corr  # noqa: E402  (the reference's core/corr.py)
    from utils.utils import coords_grid  # noqa: E402
    return ref_corr, coords_grid


def make_case(name, B, D, H, W, radius, seed, flow_sigma):
    ref_corr, coords_grid = _import_reference()
    g = torch.Generator().manual_seed(seed)
    fmap1 = torch.randn(B, D, H, W, generator=g)
    fmap2 = torch.randn(B, D, H, W, generator=g)
    grid = coords_grid(B, H, W, "cpu")
    coords_rand = grid + flow_sigma * torch.randn(B, 2, H, W, generator=g)
    coords_int = grid + torch.randint(-3, 4, (B, 2, H, W), generator=g).float()
    coords_far = grid + 1000.0
    coords_edge = grid.clone()
    # half-integer and negative positions that straddle the border
    coords_edge[:, 0] = coords_edge[:, 0] * 1.37 - 3.5
    coords_edge[:, 1] = coords_edge[:, 1] * 0.61 + (H - 2.25)

    with torch.no_grad():
        blk = ref_corr.CorrBlock(fmap1, fmap2, num_levels=4, radius=radius)
        pyr = [p[:, 0].numpy().copy() for p in blk.corr_pyramid]
        outs = {k: blk(c).numpy().copy() for k, c in
                (("rand", coords_rand), ("int", coords_int), ("far", coords_far), ("edge", coords_edge))}

    # gradients through build + lookup, as autograd computes them for the reference
    f1 = fmap1.clone().requires_grad_(True)
    f2 = fmap2.clone().requires_grad_(True)
    cr = coords_rand.clone().requires_grad_(True)
    blk = ref_corr.CorrBlock(f1, f2, num_levels=4, radius=radius)
    out = blk(cr)
    grad_out = torch.randn(out.shape, generator=g)
    out.backward(grad_out)

    # gradient of the lookup alone w.r.t. each pyramid level: make the levels leaves
    leaf = ref_corr.CorrBlock(fmap1, fmap2, num_levels=4, radius=radius)
    leaf.corr_pyramid = [p.detach().clone().requires_grad_(True) for p in leaf.corr_pyramid]
    leaf(coords_rand).backward(grad_out)
    dpyr = [p.grad[:, 0].numpy() for p in leaf.corr_pyramid]

    data = dict(
        fmap1=fmap1.numpy(), fmap2=fmap2.numpy(), radius=np.int32(radius),
        coords_rand=coords_rand.numpy(), coords_int=coords_int.numpy(),
        coords_far=coords_far.numpy(), coords_edge=coords_edge.numpy(),
        out_rand=outs["rand"], out_int=outs["int"], out_far=outs["far"], out_edge=outs["edge"],
        grad_out=grad_out.numpy(), dfmap1=f1.grad.numpy(), dfmap2=f2.grad.numpy(),
        dcoords=cr.grad.numpy(), torch_version=np.array(torch.__version__),
    )
    for l, p in enumerate(pyr):
        data[f"pyr{l}"] = p
    for l, gl in enumerate(dpyr):
        if l == 0:   # level 0 is as large as the volume: keep per-query row sums only
            data["dpyr0_rowsum"] = gl.sum(axis=2)
        else:
            data[f"dpyr{l}"] = gl
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, f"{name}.npz")
    np.savez_compressed(path, **data)
    print(f"wrote {path}: {os.path.getsize(path) / 1e6:.2f} MB")


def main():
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    torch.manual_seed(0)
    # RAFT-basic style: r = 4; odd sizes at levels 0 and 1 (17->8->4->2, 22->11->5->2)
    make_case("corrblock_basic", B=2, D=32, H=17, W=22, radius=4, seed=11, flow_sigma=2.0)
    # RAFT-small style: r = 3, narrower features
    make_case("corrblock_small", B=1, D=16, H=16, W=24, radius=3, seed=12, flow_sigma=3.0)


if __name__ == "__main__":
    main()
