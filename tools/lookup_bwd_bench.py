#!/usr/bin/env python3
"""Times the lookup backward alone (d/dpyramid accumulated into one fp32 gradient pyramid, with and without d/dcoords)
at the FlyingChairs training batch (10 pairs of 46x62) and at 16 / 64 Sintel pairs.  One JSON line per case."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R   # noqa: E402

dev = torch.device("cuda:0")
L, RAD, D = 4, 4, 256
for name, (B, H, W) in {"chairs_b10": (10, 46, 62), "sintel_b16": (16, 55, 128), "sintel_b64": (64, 55, 128)}.items():
    with torch.no_grad():
        f1 = torch.randn(B, D, H, W, device=dev)
        f2 = torch.randn(B, D, H, W, device=dev)
        blk = R.CorrBlock(f1, f2, L, RAD)
        c = R.coords_grid(B, H, W, dev) + 3 * torch.randn(B, 2, H, W, device=dev)
        go = torch.randn(B, L * (2 * RAD + 1) ** 2, H, W, device=dev)
        for want_dc in (False, True):
            fn = lambda: blk._lookup_backward(c, go, True, want_dc)
            fn(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 30
            e0.record()
            for _ in range(reps):
                fn()
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            nq = B * H * W
            byts = nq * (L * (2 * RAD + 1) ** 2 * 4 + 8 + L * (2 * RAD + 2) ** 2 * 4 * 2)
            print(json.dumps({"case": name, "dcoords": want_dc, "ms": round(ms, 4), "GB/s": round(byts / ms / 1e6, 1)}), flush=True)
        del blk, f1, f2
