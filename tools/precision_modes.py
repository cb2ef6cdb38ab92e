#!/usr/bin/env python3
"""Build and lookup time of every (storage, math) setting at BATCH (default 64) Sintel-shape pairs, and the largest
deviation of the looked-up values from the fp32/fp32 setting.  One JSON line."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R   # noqa: E402

dev = torch.device("cuda:0")
B = int(os.environ.get("BATCH", "64"))
g = torch.Generator().manual_seed(0)
f1 = torch.randn(B, 256, 55, 128, generator=g).to(dev)
f2 = torch.randn(B, 256, 55, 128, generator=g).to(dev)
coords = R.coords_grid(B, 55, 128, dev) + 5 * torch.randn(B, 2, 55, 128, generator=g).to(dev)


def timeit(fn, reps):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


res = {"pairs": B}
with torch.no_grad():
    for _ in range(2):                       # the process's first launches (module load, allocator) are not any setting's
        R.CorrBlock(f1, f2, 4, 4)(coords)
    torch.cuda.synchronize()
    ref = None
    if B <= 16:
        ref = R.CorrBlock(f1, f2, 4, 4, storage="fp32", math="fp32")(coords)
    for storage, math in (("bf16", "bf16"), ("fp16", "bf16"), ("fp16", "fp16"), ("fp16", "tf32"), ("fp32", "bf16"),
                          ("fp32", "fp16"), ("fp32", "tf32")):
        bms, blk = timeit(lambda: R.CorrBlock(f1, f2, 4, 4, storage=storage, math=math), 5)
        lms, out = timeit(lambda: blk(coords), 10)
        e = {"build_ms": bms, "lookup_ms": lms, "step_ms_build_plus_12_lookups": bms + 12 * lms}
        if ref is not None:
            e["max_abs_diff_vs_fp32"] = (out - ref).abs().max().item()
            e["rms_diff_vs_fp32"] = (out - ref).pow(2).mean().sqrt().item()
        res[f"{storage}_storage_{math}_math"] = e
        del blk, out
print(json.dumps(res))
