"""Does cudaLimitMaxL2FetchGranularity change what the lookup costs?  (experiment, GPU box)"""
import ctypes
import glob
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R  # noqa: E402

cands = glob.glob(os.path.join(os.path.dirname(torch.__file__), "lib", "libcudart*.so*")) + \
    glob.glob("/usr/local/cuda/lib64/libcudart.so*")
rt = ctypes.CDLL(cands[0])
LIMIT = 0x05   # cudaLimitMaxL2FetchGranularity
dev = torch.device("cuda:0")
torch.zeros(1, device=dev)
g = torch.Generator().manual_seed(0)
B, D, H, W = 16, 256, 55, 128
f1 = torch.randn(B, D, H, W, generator=g).to(dev)
f2 = torch.randn(B, D, H, W, generator=g).to(dev)
coords = [R.coords_grid(B, H, W, dev) + 5 * torch.randn(B, 2, H, W, device=dev) for _ in range(12)]


def run(tag):
    val = ctypes.c_size_t(0)
    rt.cudaDeviceGetLimit(ctypes.byref(val), LIMIT)
    blk = R.CorrBlock(f1, f2)
    for c in coords:
        blk(c)
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        for c in coords:
            blk(c)
    e1.record(); torch.cuda.synchronize()
    print(f"{tag}: limit={val.value}  lookup {e0.elapsed_time(e1) / 60 * 1e3:.1f} us")


run("default")
for v in (32, 128, 64):
    rc = rt.cudaDeviceSetLimit(LIMIT, ctypes.c_size_t(v))
    run(f"set {v} (rc={rc})")
