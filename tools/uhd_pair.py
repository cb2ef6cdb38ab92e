#!/usr/bin/env python3
"""One 2160x3840 (4K UHD) frame pair: 270x480 features, 129 600 queries, a 45 GB bf16 pyramid (22.4 G elements,
far beyond 32-bit offsets) on one 180 GB B200.  Builds it, runs a lookup, checks the result against the on-the-fly
fp32 block (which never materialises the volume), and times both.  One JSON line."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R   # noqa: E402

dev = torch.device("cuda:0")
H, W, D = 270, 480, 256
g = torch.Generator().manual_seed(4)
f1 = torch.randn(1, D, H, W, generator=g).to(dev)
f2 = torch.randn(1, D, H, W, generator=g).to(dev)
coords = R.coords_grid(1, H, W, dev) + 6 * torch.randn(1, 2, H, W, generator=g).to(dev)


def timeit(fn, reps):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


with torch.no_grad():
    build_ms, blk = timeit(lambda: R.CorrBlock(f1, f2, 4, 4), 2)
    look_ms, got = timeit(lambda: blk(coords), 5)
    elems, gb = int(blk.layout.total_elems), int(blk.layout.total_elems) * 2 / 1e9
    alt = R.AlternateCorrBlock(f1, f2, 4, 4, math="fp32")
    alt_ms, ref = timeit(lambda: alt(coords), 2)
    alt_tc = R.AlternateCorrBlock(f1, f2, 4, 4, math="bf16")
    alt_tc_ms, got_tc = timeit(lambda: alt_tc(coords), 5)
    scale = ref.abs().mean().item()
    res = {"shape": f"{H}x{W}x{D}", "pyramid_elements": elems, "pyramid_GB": gb, "build_ms": build_ms,
           "build_TB/s": gb / build_ms, "lookup_ms": look_ms, "on_the_fly_fp32_lookup_ms": alt_ms,
           "on_the_fly_tensor_core_lookup_ms": alt_tc_ms,
           "max_abs_diff_vs_on_the_fly_fp32": (got - ref).abs().max().item(),
           "tensor_core_on_the_fly_max_abs_diff": (got_tc - ref).abs().max().item(), "mean_abs_value": scale,
           "peak_mem_GB": torch.cuda.max_memory_allocated(dev) / 1e9}
print(json.dumps(res))
