#!/usr/bin/env python3
"""Short driver for ncu captures of the kernels bench.py's timed step does not launch: the lookup
backward, the build backward helpers and the on-the-fly kernels (Chairs shape for the backward,
Sintel shape for the on-the-fly lookup).  No timing here; see bench.py extras for numbers."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R   # noqa: E402

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
B, D, H, W = 10, 256, 46, 62
f1 = torch.randn(B, D, H, W, generator=g).to(dev).requires_grad_(True)
f2 = torch.randn(B, D, H, W, generator=g).to(dev).requires_grad_(True)
cs = [R.coords_grid(B, H, W, dev) + 3 * torch.randn(B, 2, H, W, generator=g).to(dev) for _ in range(3)]
for rep in range(2):
    f1.grad = None; f2.grad = None
    blk = R.CorrBlock(f1, f2, 4, 4)
    loss = sum(blk(c).sum() for c in cs)
    loss.backward()
with torch.no_grad():
    a1 = torch.randn(1, D, 55, 128, generator=g).to(dev)
    a2 = torch.randn(1, D, 55, 128, generator=g).to(dev)
    c = R.coords_grid(1, 55, 128, dev) + 5 * torch.randn(1, 2, 55, 128, generator=g).to(dev)
    alt = R.AlternateCorrBlock(a1, a2, 4, 4, math="fp32")
    for _ in range(2):
        alt(c)
torch.cuda.synchronize()
print("ok")
