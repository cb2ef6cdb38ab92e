#!/usr/bin/env python3
"""Throughput of the tensor-core on-the-fly lookup at batch sizes 1..64 (Sintel shape), smooth and random flow,
next to the materialised path (build once + lookup) for the same batch."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R

dev = torch.device("cuda:0")
H, W, D = 55, 128, 256
def timeit(fn, reps):
    fn(); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
res = {}
with torch.no_grad():
    for B in (1, 4, 16, 64):
        g = torch.Generator().manual_seed(B)
        f1 = torch.randn(B, D, H, W, generator=g).to(dev); f2 = torch.randn(B, D, H, W, generator=g).to(dev)
        ys, xs = torch.meshgrid(torch.arange(H, device=dev), torch.arange(W, device=dev), indexing="ij")
        smooth = R.coords_grid(B, H, W, dev) + torch.stack([3.3 + 0.02 * xs, -2.7 + 0.03 * ys])[None].float()
        rnd = R.coords_grid(B, H, W, dev) + 5 * torch.randn(B, 2, H, W, generator=g).to(dev)
        tc = R.AlternateCorrBlock(f1, f2, 4, 4, math="bf16")
        blk = R.CorrBlock(f1, f2, 4, 4)
        res[f"B{B}"] = {"alt_tc_smooth_us": 1e3 * timeit(lambda: tc(smooth), 20), "alt_tc_random_us": 1e3 * timeit(lambda: tc(rnd), 20),
                        "default_lookup_us": 1e3 * timeit(lambda: blk(rnd), 20),
                        "default_build_us": 1e3 * timeit(lambda: R.CorrBlock(f1, f2, 4, 4), 5)}
print(json.dumps(res))
