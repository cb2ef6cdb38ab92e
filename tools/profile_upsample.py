#!/usr/bin/env python3
"""Driver for an ncu capture of the flow-upsampling kernels: 64 Sintel-size maps, forward x3, forward+backward x1."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R   # noqa: E402

dev = torch.device("cuda:0")
B, H, W = int(os.environ.get("BATCH", "64")), 55, 128
g = torch.Generator().manual_seed(0)
flow = (5 * torch.randn(B, 2, H, W, generator=g)).to(dev)
mask = (2 * torch.randn(B, 576, H, W, generator=g)).to(dev)
with torch.no_grad():
    for _ in range(3):
        R.upsample_flow(flow, mask)
fl, mk = flow.clone().requires_grad_(True), mask.clone().requires_grad_(True)
up = R.upsample_flow(fl, mk)
torch.autograd.grad(up, (fl, mk), torch.randn_like(up))
torch.cuda.synchronize()
print("ok")
