#!/usr/bin/env python3
"""Times rc_build_bwd (tensor-core backward of the volume build) alone, both precision settings, at the FlyingChairs
training shape (10 pairs of 46x62) and the Sintel shape (16 pairs of 55x128).  One JSON line per case.
PROFILE=1: one call per case only (for ncu)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from raft_stir_b200 import _lib  # noqa: E402

dev = torch.device("cuda:0")
lib = _lib.lib()
profile = os.environ.get("PROFILE") == "1"
for name, (B, D, H, W, L) in {"chairs_b10": (10, 256, 46, 62, 4), "sintel_b16": (16, 256, 55, 128, 4)}.items():
    f1 = torch.randn(B, D, H, W, device=dev)
    f2 = torch.randn(B, D, H, W, device=dev)
    glay = _lib.pyramid_layout(B, H, W, L, _lib.RC_VOL_F32)
    g = torch.randn(int(glay.total_elems), device=dev)
    ws_bytes = int(lib.rc_build_bwd_workspace_bytes(B, D, H, W, L))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    d1, d2 = torch.empty_like(f1), torch.empty_like(f2)
    st = torch.cuda.current_stream().cuda_stream
    slabs = sum(int(glay.slab[l]) for l in range(L))
    N = H * W
    for math, tag in ((_lib.RC_MATH_TF32, "tf32"), (_lib.RC_MATH_F32, "3xtf32")):
        def call():
            _lib.check(lib.rc_build_bwd(g.data_ptr(), f1.data_ptr(), f2.data_ptr(), B, D, H, W, L, math, d1.data_ptr(),
                                        d2.data_ptr(), ws.data_ptr(), ws_bytes, st), "rc_build_bwd")
        call(); torch.cuda.synchronize()
        if profile:
            continue
        reps = 20
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            call()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        flops = 2 * 2.0 * B * N * D * slabs
        grad_bytes = B * N * slabs * 4
        print(json.dumps({"case": name, "math": tag, "ms": ms, "TFLOP/s": flops / ms / 1e9,
                          "gradient_pyramid_GB": grad_bytes / 1e9,
                          "GB/s_if_gradient_read_once": grad_bytes / ms / 1e6,
                          "GB/s_if_gradient_read_twice": 2 * grad_bytes / ms / 1e6}), flush=True)
