#!/usr/bin/env python3
"""SURVEY 8(f) row 1 in numbers: lookup -> convc1 -> ReLU at 64 (BATCH) Sintel-shape pairs, (a) as the reference
composes it on this library's lookup (fp32 [B,324,H,W] out, cuDNN 1x1 convolution, ReLU) and (b) fused
(rc_lookup_convc1_fwd: fp16 samples, tcgen05 convolution).  One JSON line.  PROFILE=1: one pass of each (for ncu)."""
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R   # noqa: E402

dev = torch.device("cuda:0")
B = int(os.environ.get("BATCH", "64"))
H, W, D = 55, 128, 256
profile = os.environ.get("PROFILE") == "1"
torch.backends.cudnn.allow_tf32 = False
f1 = torch.randn(B, D, H, W, device=dev)
f2 = torch.randn(B, D, H, W, device=dev)
coords = R.coords_grid(B, H, W, dev) + 5 * torch.randn(B, 2, H, W, device=dev)
w = torch.randn(256, 324, 1, 1, device=dev) / 18
b = torch.randn(256, device=dev) * 0.1
with torch.no_grad():
    blk = R.CorrBlock(f1, f2, 4, 4)
    del f1, f2
    wp = R.pack_convc1_weights(w)

    def unfused():
        return F.relu(F.conv2d(blk(coords), w, b))

    def fused():
        return blk.lookup_convc1(coords, w, b, packed_weight=wp)

    def fused16():
        return blk.lookup_convc1(coords, w, b, out_dtype=torch.float16, packed_weight=wp)

    def lookup_only():
        return blk(coords)

    res = {"pairs": B}
    for name, fn in (("lookup_fp32_out", lookup_only), ("lookup+cudnn_conv1x1+relu", unfused), ("fused_fp32_out", fused),
                     ("fused_fp16_out", fused16)):
        for _ in range(1 if profile else 3):
            o = fn()
        torch.cuda.synchronize()
        if profile:
            continue
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 20
        e0.record()
        for _ in range(reps):
            o = fn()
        e1.record(); torch.cuda.synchronize()
        res[name + "_ms"] = e0.elapsed_time(e1) / reps
    if not profile:
        ref = unfused(); got = fused()
        res["max_abs_diff_fused_vs_unfused"] = (ref - got).abs().max().item()
        N = B * H * W
        res["algorithmic_bytes_unfused"] = N * (2104 + 1296 + 1024)
        res["algorithmic_bytes_fused_fp32_out"] = N * (808 + 648 + 648 + 1024)
    print(json.dumps(res))
