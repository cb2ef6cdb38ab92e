#!/usr/bin/env python3
"""On-the-fly correlation: this library's AlternateCorrBlock against the reference's own
alt_cuda_corr extension (compiled unmodified, oracle/_ref) driven exactly as the reference's
AlternateCorrBlock.__call__ drives it (core/corr.py:74-91: per level two NHWC permute copies, one
extension call, then stack and scale).  One JSON line: ms per lookup at the Sintel and 1080p shapes."""
import json
import math
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import raft_stir_b200 as R      # noqa: E402
from oracle import ref_ext      # noqa: E402

dev = torch.device("cuda:0")
ext = ref_ext.load()


def ref_block(f1, f2, levels=4, radius=4):
    pyr = [(f1, f2)]
    for _ in range(levels):
        f1p, f2p = F.avg_pool2d(pyr[-1][0], 2, stride=2), F.avg_pool2d(pyr[-1][1], 2, stride=2)
        pyr.append((f1p, f2p))

    def call(coords):
        c = coords.permute(0, 2, 3, 1)
        B, H, W, _ = c.shape
        dim = pyr[0][0].shape[1]
        outs = []
        for i in range(levels):
            a = pyr[0][0].permute(0, 2, 3, 1).contiguous()
            b = pyr[i][1].permute(0, 2, 3, 1).contiguous()
            ci = (c / 2 ** i).reshape(B, 1, H, W, 2).contiguous()
            (corr,) = ext.forward(a, b, ci, radius)
            outs.append(corr.squeeze(1))
        corr = torch.stack(outs, dim=1).reshape(B, -1, H, W)
        return corr / math.sqrt(dim)
    return call


def timeit(fn, reps):
    fn(); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


res = {}
with torch.no_grad():
    for name, (H, W) in (("sintel_55x128", (55, 128)), ("video_135x240", (135, 240))):
        g = torch.Generator().manual_seed(0)
        f1 = torch.randn(1, 256, H, W, generator=g).to(dev)
        f2 = torch.randn(1, 256, H, W, generator=g).to(dev)
        coords = R.coords_grid(1, H, W, dev) + 5 * torch.randn(1, 2, H, W, generator=g).to(dev)
        ref = ref_block(f1, f2)
        ours = R.AlternateCorrBlock(f1, f2, 4, 4, math="fp32")
        ours_tc = R.AlternateCorrBlock(f1, f2, 4, 4, math="bf16")
        ms_ref, o_ref = timeit(lambda: ref(coords), 5)
        ms_ours, o_ours = timeit(lambda: ours(coords), 10)
        err = ((o_ours - o_ref).abs().max() / o_ref.abs().max()).item()
        res[name] = {"reference_ext_ms": ms_ref, "ours_ms": ms_ours, "speedup": ms_ref / ms_ours, "max_rel_diff": err}
        # tensor-core variant: a smooth flow (what video gives) and the scattered random one
        ys, xs = torch.meshgrid(torch.arange(H, device=dev), torch.arange(W, device=dev), indexing="ij")
        smooth = R.coords_grid(1, H, W, dev) + torch.stack([3.3 + 0.02 * xs, -2.7 + 0.03 * ys])[None].float()
        for tag, cc in (("smooth_flow", smooth), ("random_flow_sigma5", coords)):
            ms_tc, o_tc = timeit(lambda: ours_tc(cc), 20)
            ms_f, o_f = timeit(lambda: ours(cc), 10)
            res[name][f"ours_tc_ms_{tag}"] = ms_tc
            res[name][f"ours_fp32_ms_{tag}"] = ms_f
            res[name][f"tc_vs_fp32_max_rel_diff_{tag}"] = ((o_tc - o_f).abs().max() / o_f.abs().max()).item()
print(json.dumps(res))
