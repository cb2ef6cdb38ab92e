#!/usr/bin/env python3
"""Driver for an ncu capture of the persistent tensor-core on-the-fly kernel at a batch of Sintel-shape pairs:
three launches with a smooth flow, three with a random one.  BATCH env (default 16)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R   # noqa: E402

dev = torch.device("cuda:0")
B = int(os.environ.get("BATCH", "16"))
H, W = 55, 128
g = torch.Generator().manual_seed(0)
with torch.no_grad():
    f1 = torch.randn(B, 256, H, W, generator=g).to(dev)
    f2 = torch.randn(B, 256, H, W, generator=g).to(dev)
    ys, xs = torch.meshgrid(torch.arange(H, device=dev), torch.arange(W, device=dev), indexing="ij")
    smooth = R.coords_grid(B, H, W, dev) + torch.stack([3.3 + 0.02 * xs, -2.7 + 0.03 * ys])[None].float()
    rnd = R.coords_grid(B, H, W, dev) + 5 * torch.randn(B, 2, H, W, generator=g).to(dev)
    blk = R.AlternateCorrBlock(f1, f2, 4, 4, math="bf16")
    for c in (smooth, rnd):
        for _ in range(3):
            blk(c)
torch.cuda.synchronize()
print("ok")
