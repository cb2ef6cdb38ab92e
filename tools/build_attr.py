#!/usr/bin/env python3
"""Times the build from NHWC bf16 features (no pack in the way) at 16 and 64 Sintel pairs, CUDA events, 30 calls back to
back, for whatever RC_BUILD_PAIR / RAFTCORR_LIB + RC_TC_DEBUG / RC_TC_GROUP the environment selects (tools/gpu_attr.sh,
tools/gpu_group.sh).  One JSON line per batch size."""
import json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R
from raft_stir_b200 import _lib
dev = torch.device("cuda:0"); H, W, D, L = 55, 128, 256, 4
lib = _lib.lib(); st = torch.cuda.current_stream().cuda_stream
for B in (16, 64):
    f1 = torch.randn(B, D, H, W, device=dev); f2 = torch.randn(B, D, H, W, device=dev)
    p1, p2 = R.pack_features(f1, f2)
    lay = _lib.pyramid_layout(B, H, W, L, _lib.RC_VOL_BF16)
    buf = torch.empty(int(lay.total_elems), dtype=torch.bfloat16, device=dev)
    call = lambda: _lib.check(lib.rc_build_pyramid_nhwc16(p1.data_ptr(), p2.data_ptr(), _lib.RC_VOL_BF16, B, D, H, W, L, _lib.RC_VOL_BF16, buf.data_ptr(), st), "nhwc16")
    for _ in range(3): call()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 30; e0.record()
    for _ in range(reps): call()
    e1.record(); torch.cuda.synchronize()
    print(json.dumps({"dbg": os.environ.get("RC_TC_DEBUG", "0"), "pair": os.environ.get("RC_BUILD_PAIR", "1"), "pairs": B, "ms": round(e0.elapsed_time(e1) / reps, 4)}), flush=True)
