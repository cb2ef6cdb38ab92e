#!/usr/bin/env python3
"""Flow upsampling: this library's fused kernel against the reference's torch ops (core/raft.py:72-83) on the same GPU,
forward and forward+backward, at the Sintel feature-map size for 1 and 64 pairs and at 1080p.  One JSON line."""
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R   # noqa: E402

dev = torch.device("cuda:0")


def ref_upsample(flow, mask):
    N, _, H, W = flow.shape
    mask = torch.softmax(mask.view(N, 1, 9, 8, 8, H, W), dim=2)
    up = F.unfold(8 * flow, [3, 3], padding=1).view(N, 2, 9, 1, 1, H, W)
    up = torch.sum(mask * up, dim=2).permute(0, 1, 4, 2, 5, 3)
    return up.reshape(N, 2, 8 * H, 8 * W)


def timeit(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


res = {}
for name, (B, H, W) in {"sintel_b1": (1, 55, 128), "sintel_b64": (64, 55, 128), "v1080_b1": (1, 135, 240)}.items():
    g = torch.Generator().manual_seed(0)
    flow = (5 * torch.randn(B, 2, H, W, generator=g)).to(dev)
    mask = (2 * torch.randn(B, 576, H, W, generator=g)).to(dev)
    gout = torch.randn(B, 2, 8 * H, 8 * W, generator=g).to(dev)
    with torch.no_grad():
        a, b = R.upsample_flow(flow, mask), ref_upsample(flow, mask)
        err = (a - b).abs().max().item()
        ours = timeit(lambda: R.upsample_flow(flow, mask), 50)
        ref = timeit(lambda: ref_upsample(flow, mask), 20)
        m16 = mask.half()
        ours16 = timeit(lambda: R.upsample_flow(flow, m16), 50)
        ref16 = timeit(lambda: ref_upsample(flow, m16), 20)       # what the reference does with autocast's fp16 mask
    fl, mk = flow.clone().requires_grad_(True), mask.clone().requires_grad_(True)

    def fb(fn):
        up = fn(fl, mk)
        torch.autograd.grad(up, (fl, mk), gout)
    ours_fb, ref_fb = timeit(lambda: fb(R.upsample_flow), 20), timeit(lambda: fb(ref_upsample), 10)
    bytes_fwd = B * H * W * (576 * 4 + 2 * 4 + 128 * 4)
    res[name] = {"ours_ms": ours, "reference_torch_ms": ref, "speedup": ref / ours, "max_abs_diff": err,
                 "ours_GB/s": bytes_fwd / ours / 1e6, "ours_fp16_mask_ms": ours16, "reference_torch_fp16_mask_ms": ref16, "ours_fwd_bwd_ms": ours_fb, "reference_fwd_bwd_ms": ref_fb}
print(json.dumps(res))
