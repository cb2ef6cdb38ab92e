#!/usr/bin/env python3
"""Randomised stress of the tensor-core volume build (bf16, fp16 and tf32 operands, all three storage types) against the
CUDA-core fp32 build of the same library on operand-rounded features, plus a lookup on each: random batch, channel
count, map size and number of levels, ITER cases (default 200).  One JSON line; exits 1 on any mismatch."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R   # noqa: E402

dev = torch.device("cuda:0")
iters = int(os.environ.get("ITER", "200"))
rng = np.random.default_rng(int(os.environ.get("SEED", "0")))


def tf32_round(x):
    u = x.contiguous().view(torch.int32)
    return ((u + 0x1000) & ~0x1FFF).view(torch.float32)


bad, worst = [], 0.0
with torch.no_grad():
    for it in range(iters):
        B = int(rng.integers(1, 5))
        D = int(rng.choice([16, 24, 64, 96, 128, 256]))
        H = int(rng.integers(8, 64))
        W = int(rng.integers(8, 90))
        L = int(rng.choice([1, 2, 3, 4]))
        math = ["bf16", "tf32", "fp16"][it % 3]
        storage = ["fp32", "bf16", "fp16"][(it // 3) % 3]
        g = torch.Generator().manual_seed(1000 + it)
        f1 = torch.randn(B, D, H, W, generator=g).to(dev)
        f2 = torch.randn(B, D, H, W, generator=g).to(dev)
        r1, r2 = {"bf16": (f1.bfloat16().float(), f2.bfloat16().float()), "fp16": (f1.half().float(), f2.half().float()),
                  "tf32": (tf32_round(f1), tf32_round(f2))}[math]
        ref = R.CorrBlock(r1, r2, L, 4, storage="fp32", math="fp32")
        blk = R.CorrBlock(f1, f2, L, 4, storage=storage, math=math)
        tol = {"fp32": 2e-5, "bf16": 2.0 ** -8, "fp16": 2.0 ** -11 * 1.01}[storage]
        for l, (a, b) in enumerate(zip(blk.corr_pyramid, ref.corr_pyramid)):
            a, b = a.float(), b.float()
            rms = b.pow(2).mean().sqrt().item() + 1e-12
            viol = ((a - b).abs() - tol * (b.abs() + rms)).max().item()
            worst = max(worst, ((a - b).abs() / (b.abs() + rms)).max().item() / tol)
            if viol > 0 or not torch.isfinite(a).all():
                bad.append({"case": it, "B": B, "D": D, "H": H, "W": W, "L": L, "math": math, "storage": storage, "level": l,
                            "excess": viol})
        coords = R.coords_grid(B, H, W, dev) + 4 * torch.randn(B, 2, H, W, generator=g).to(dev)
        lo, lr = blk(coords), ref(coords)
        rms = ref.corr_pyramid[0].float().pow(2).mean().sqrt().item() + 1e-12     # of the volume (windows outside the map read zeros)
        # an interpolated value carries the storage rounding of its four taps, each of which can be several rms large
        if ((lo - lr).abs() - tol * (lr.abs() + 6 * rms)).max().item() > 0 or not torch.isfinite(lo).all():
            bad.append({"case": it, "lookup": True, "B": B, "D": D, "H": H, "W": W, "L": L, "math": math, "storage": storage})
torch.cuda.synchronize()
print(json.dumps({"cases": iters, "worst_error_over_tolerance": worst, "n_mismatch": len(bad), "mismatches": bad[:10]}))
sys.exit(1 if bad else 0)
