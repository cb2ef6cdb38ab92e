#!/usr/bin/env python3
"""Randomised check of the training direction: d/dfmap1, d/dfmap2 and d/dcoords of build + several lookups through this
library's autograd-enabled CorrBlock against torch autograd of the reference's own torch calls (oracle/torch_port.py) on
the same GPU, in fp32 (TF32 off).  Random batch, channels, map size, levels, radius, number of lookups.  One JSON line;
exits 1 on a mismatch.  (Lives under tests/ because it uses the oracle as the checker.)"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import raft_stir_b200 as R            # noqa: E402
from oracle import torch_port as P    # noqa: E402

torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
dev = torch.device("cuda:0")
iters = int(os.environ.get("ITER", "100"))
rng = np.random.default_rng(int(os.environ.get("SEED", "0")))
bad, worst, boundary = [], 0.0, 0
for it in range(iters):
    B = int(rng.integers(1, 4)); D = int(rng.choice([16, 32, 64, 128])); H = int(rng.integers(8, 40)); W = int(rng.integers(8, 56))
    L = int(rng.choice([2, 3, 4])) if min(H, W) >= 16 else 2
    r = int(rng.choice([3, 4])); nlook = int(rng.integers(1, 4))
    g = torch.Generator().manual_seed(500 + it)
    f1 = torch.randn(B, D, H, W, generator=g).to(dev); f2 = torch.randn(B, D, H, W, generator=g).to(dev)
    cs = [R.coords_grid(B, H, W, dev) + 3 * torch.randn(B, 2, H, W, generator=g).to(dev) for _ in range(nlook)]
    gos = [torch.randn(B, L * (2 * r + 1) ** 2, H, W, generator=g).to(dev) for _ in range(nlook)]

    def run(make, look):
        a1, a2 = f1.clone().requires_grad_(True), f2.clone().requires_grad_(True)
        cc = [c.clone().requires_grad_(True) for c in cs]
        blk = make(a1, a2)
        loss = sum((look(blk, c) * go).sum() for c, go in zip(cc, gos))
        grads = torch.autograd.grad(loss, [a1, a2] + cc)
        return grads

    ref = run(lambda a, b: P.build_pyramid(a, b, L), lambda pyr, c: P.lookup(pyr, c, r))
    got = run(lambda a, b: R.CorrBlock(a, b, L, r, storage="fp32", math="fp32"), lambda blk, c: blk(c))
    for name, a, b in zip(["dfmap1", "dfmap2"] + [f"dcoords{i}" for i in range(nlook)], got, ref):
        rms = b.pow(2).mean().sqrt().item() + 1e-12
        e = ((a - b).abs() / (b.abs() + rms)).max().item()
        worst = max(worst, e)
        if e > 2e-4 or not torch.isfinite(a).all():
            off = ((a - b).abs() / (b.abs() + rms)) > 2e-4
            if name.startswith("dcoords"):
                # d/dcoords jumps where a coordinate crosses a cell boundary of some level; the reference reaches the
                # sample position through a normalise / denormalise round trip (core/utils/utils.py:61-62), so a
                # coordinate within fp32 rounding of a boundary may sit on the other side there.  Not a mismatch.
                c = cs[int(name[7:])]
                near = torch.zeros_like(c, dtype=torch.bool)
                for l in range(L):
                    v = c / 2 ** l
                    near |= (v - v.round()).abs() < 2e-5 * (1 + v.abs())
                pix = off.any(dim=1, keepdim=True)
                if bool((pix & ~near.any(dim=1, keepdim=True)).any()) is False:
                    boundary += int(pix.sum())
                    continue
            bad.append({"case": it, "B": B, "D": D, "H": H, "W": W, "L": L, "r": r, "grad": name, "rel": e,
                        "elements": int(off.sum())})
torch.cuda.synchronize()
print(json.dumps({"cases": iters, "worst_rel_including_boundary_pixels": worst, "pixels_on_a_cell_boundary": boundary,
                  "n_mismatch": len(bad), "mismatches": bad[:8]}))
sys.exit(1 if bad else 0)
