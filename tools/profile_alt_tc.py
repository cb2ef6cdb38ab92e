#!/usr/bin/env python3
"""Driver for an ncu capture of the tensor-core on-the-fly kernel: one Sintel-shape pair and one 1080p pair,
smooth flow, a few launches each."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R   # noqa: E402

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
with torch.no_grad():
    for H, W in ((55, 128), (135, 240)):
        f1 = torch.randn(1, 256, H, W, generator=g).to(dev)
        f2 = torch.randn(1, 256, H, W, generator=g).to(dev)
        ys, xs = torch.meshgrid(torch.arange(H, device=dev), torch.arange(W, device=dev), indexing="ij")
        c = R.coords_grid(1, H, W, dev) + torch.stack([3.3 + 0.02 * xs, -2.7 + 0.03 * ys])[None].float()
        blk = R.AlternateCorrBlock(f1, f2, 4, 4, math="bf16")
        for _ in range(3):
            blk(c)
torch.cuda.synchronize()
print("ok")
