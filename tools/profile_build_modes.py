#!/usr/bin/env python3
"""Driver for an ncu capture of the tensor-core build in its three operand/storage modes: 16 Sintel-shape pairs,
bf16 math -> bf16 storage (the default), bf16 math -> fp32 storage, tf32 math -> fp32 storage; two builds each."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R   # noqa: E402

dev = torch.device("cuda:0")
B = int(os.environ.get("BATCH", "16"))
g = torch.Generator().manual_seed(0)
f1 = torch.randn(B, 256, 55, 128, generator=g).to(dev)
f2 = torch.randn(B, 256, 55, 128, generator=g).to(dev)
with torch.no_grad():
    for storage, math in (("bf16", "bf16"), ("fp32", "bf16"), ("fp32", "tf32")):
        for _ in range(2):
            blk = R.CorrBlock(f1, f2, 4, 4, storage=storage, math=math)
            del blk
torch.cuda.synchronize()
print("ok")
