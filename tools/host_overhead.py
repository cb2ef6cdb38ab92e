#!/usr/bin/env python3
"""Host time per call (no synchronisation inside the loop) next to the GPU time per call, one Sintel pair:
is a call's host side (argument marshalling, TMA descriptor encoding, launch) shorter than its kernel?"""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R   # noqa: E402

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
H, W = 55, 128
f1 = torch.randn(1, 256, H, W, generator=g).to(dev)
f2 = torch.randn(1, 256, H, W, generator=g).to(dev)
c = R.coords_grid(1, H, W, dev) + 3 * torch.randn(1, 2, H, W, generator=g).to(dev)
fl = torch.randn(1, 2, H, W, generator=g).to(dev)
mk = torch.randn(1, 576, H, W, generator=g).to(dev)


def probe(fn, n=300):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    host = (time.perf_counter() - t0) / n * 1e6
    torch.cuda.synchronize()
    return {"host_us": round(host, 1), "gpu_us": round(e0.elapsed_time(e1) / n * 1e3, 1)}


with torch.no_grad():
    blk = R.CorrBlock(f1, f2, 4, 4)
    alt = R.AlternateCorrBlock(f1, f2, 4, 4, math="bf16")
    res = {"build": probe(lambda: R.CorrBlock(f1, f2, 4, 4), 100), "lookup": probe(lambda: blk(c)),
           "alt_tc_lookup": probe(lambda: alt(c)), "upsample": probe(lambda: R.upsample_flow(fl, mk))}
print(json.dumps(res))
