#!/usr/bin/env python3
"""Clocks and power while ONE kernel family runs back to back for about a second each: the tensor-core build (from
packed features), the lookup, a plain memset and a cuBLAS bf16 GEMM -- does the 1 kW cap pace them, and what does a
launch cost in joules (mean power x time)?  nvidia-smi sampled every 10 ms.

    python tools/power_probe.py [case ...]          # default: all cases
    RAFTCORR_LIB=raft_stir_b200/_build/attrib/libraftcorr.so RC_TC_DEBUG=8 python tools/power_probe.py build_64_pairs_nhwc16
      (the attribution build, python -m raft_stir_b200.build --attrib: 8 = operand loads + MMAs + TMEM reads only,
       1 = everything but the global stores)"""
import json
import os
import subprocess
import sys
import threading
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R   # noqa: E402

dev = torch.device("cuda:0")
B = 64
f1 = torch.randn(B, 256, 55, 128, device=dev)
f2 = torch.randn(B, 256, 55, 128, device=dev)
p1, p2 = R.pack_features(f1, f2)
coords = R.coords_grid(B, 55, 128, dev) + 5 * torch.randn(B, 2, 55, 128, device=dev)


class Sampler:
    def __init__(self):
        self.lines = []
        self.p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap,"
                                   "clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown",
                                   "--format=csv,noheader,nounits", "-lms", "10", "-i", "0"],
                                  stdout=subprocess.PIPE, text=True)
        threading.Thread(target=self._rd, daemon=True).start()

    def _rd(self):
        for ln in self.p.stdout:
            self.lines.append((time.perf_counter(), ln.strip()))

    def window(self, t0, t1):
        sel = [l for t, l in self.lines if t0 <= t <= t1]
        clk = [float(l.split(",")[0]) for l in sel]
        pw = [float(l.split(",")[1]) for l in sel]
        cap = sum("Active" in l.split(",")[2] and "Not" not in l.split(",")[2] for l in sel)
        return {"samples": len(sel), "sm_mhz_median": sorted(clk)[len(clk) // 2] if clk else None,
                "sm_mhz_min": min(clk) if clk else None, "power_w_median": sorted(pw)[len(pw) // 2] if pw else None,
                "power_w_mean": sum(pw) / len(pw) if pw else None,
                "power_w_max": max(pw) if pw else None, "sw_power_cap_samples": cap}


s = Sampler()
time.sleep(0.5)
with torch.no_grad():
    blk = R.CorrBlock(p1, p2, 4, 4)
    cases = {
        "build_64_pairs_nhwc16": lambda: R.CorrBlock(p1, p2, 4, 4),
        "lookup_64_pairs": lambda: blk(coords),
        "zero_2GiB": lambda: big.zero_(),
        "cublas_bf16_8192": lambda: torch.matmul(ga, gb, out=gc),
    }
    big = torch.empty(1 << 31, dtype=torch.uint8, device=dev)
    ga = torch.randn(8192, 8192, device=dev, dtype=torch.bfloat16)
    gb = torch.randn(8192, 8192, device=dev, dtype=torch.bfloat16)
    gc = torch.empty(8192, 8192, device=dev, dtype=torch.bfloat16)
    want = sys.argv[1:] or list(cases)
    for name, fn in cases.items():
        if name not in want:
            continue
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 0
        t0 = time.perf_counter()
        e0.record()
        while time.perf_counter() - t0 < 1.2:
            for _ in range(10):
                fn()
            reps += 10
            torch.cuda.synchronize()
        e1.record(); torch.cuda.synchronize()
        t1 = time.perf_counter()
        out = {"case": name, "lib": os.environ.get("RAFTCORR_LIB", "default"), "RC_TC_DEBUG": os.environ.get("RC_TC_DEBUG", "0"),
               "RC_BUILD_PAIR": os.environ.get("RC_BUILD_PAIR", "1"), "reps": reps, "avg_ms": e0.elapsed_time(e1) / reps}
        out.update(s.window(t0 + 0.3, t1))
        if out["power_w_mean"]:
            out["joule_per_call"] = out["power_w_mean"] * out["avg_ms"] * 1e-3
        print(json.dumps(out), flush=True)
        time.sleep(0.5)
s.p.terminate()
