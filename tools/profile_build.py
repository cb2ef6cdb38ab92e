#!/usr/bin/env python3
"""Two tensor-core builds of 16 (BATCH) Sintel-shape pairs from packed NHWC bf16 features (no pack kernel in the way):
the driver for an ncu capture of build_pyramid_tc_kernel."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R   # noqa: E402

dev = torch.device("cuda:0")
B = int(os.environ.get("BATCH", "16"))
f1 = torch.randn(B, 256, 55, 128, device=dev)
f2 = torch.randn(B, 256, 55, 128, device=dev)
p1, p2 = R.pack_features(f1, f2)
with torch.no_grad():
    for _ in range(3):
        blk = R.CorrBlock(p1, p2, 4, 4)
        del blk
torch.cuda.synchronize()
print("ok")
