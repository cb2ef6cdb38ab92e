#!/usr/bin/env python3
"""Times the volume build alone (pack + tensor-core build through the C ABI, CUDA events) at the Sintel shape for
16 and 64 pairs in the main storage/math settings, and the NHWC-input build (no pack).  One JSON line per case."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R   # noqa: E402
from raft_stir_b200 import _lib  # noqa: E402

dev = torch.device("cuda:0")
H, W, D, L = 55, 128, 256, 4
N = H * W
E = sum(N * (H >> l) * (W >> l) for l in range(L))
lib = _lib.lib()
st = torch.cuda.current_stream().cuda_stream
for B in (16, 64):
    f1 = torch.randn(B, D, H, W, device=dev)
    f2 = torch.randn(B, D, H, W, device=dev)
    p1, p2 = R.pack_features(f1, f2)
    for storage, math in (("bf16", "bf16"), ("fp32", "bf16"), ("fp32", "tf32"), ("bf16", "nhwc16")):
        vd = _lib.RC_VOL_BF16 if storage == "bf16" else _lib.RC_VOL_F32
        lay = _lib.pyramid_layout(B, H, W, L, vd)
        buf = torch.empty(int(lay.total_elems), dtype=torch.bfloat16 if storage == "bf16" else torch.float32, device=dev)
        if math == "nhwc16":
            def call():
                _lib.check(lib.rc_build_pyramid_nhwc16(p1.data_ptr(), p2.data_ptr(), _lib.RC_VOL_BF16, B, D, H, W, L, vd,
                                                       buf.data_ptr(), st), "nhwc16")
        else:
            m = {"bf16": _lib.RC_MATH_BF16, "tf32": _lib.RC_MATH_TF32}[math]
            wsb = int(lib.rc_build_workspace_bytes(B, D, H, W, m))
            ws = torch.empty(wsb, dtype=torch.uint8, device=dev)

            def call():
                _lib.check(lib.rc_build_pyramid(f1.data_ptr(), f2.data_ptr(), B, D, H, W, L, vd, m, buf.data_ptr(),
                                                ws.data_ptr(), wsb, st), "build")
        for _ in range(3):
            call()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 20
        e0.record()
        for _ in range(reps):
            call()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        es = 2 if storage == "bf16" else 4
        byts = B * ((0 if math == "nhwc16" else 2 * N * D * 4) + (2 * N * D * 2 if math == "nhwc16" else 0) + E * es)
        print(json.dumps({"pairs": B, "storage": storage, "math": math, "ms": round(ms, 4),
                          "GB/s": round(byts / ms / 1e6, 1), "TFLOP/s": round(B * 2.0 * N * N * D / ms / 1e9, 1)}), flush=True)
        del buf
