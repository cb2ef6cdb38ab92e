"""Diagnostics for the tensor-core build: per-level error maps against the
CUDA-core fp32 build of the same library (run on the GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import raft_stir_b200 as R

dev = torch.device("cuda:0")
shapes = [(1, 64, 8, 32), (1, 64, 16, 64), (1, 256, 55, 128), (2, 128, 46, 62)]
if len(sys.argv) > 1:
    shapes = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]]
for (B, D, H, W) in shapes:
    for storage in ("fp32", "bf16"):
        g = torch.Generator(device="cpu").manual_seed(0)
        f1 = torch.randn(B, D, H, W, generator=g).to(dev)
        f2 = torch.randn(B, D, H, W, generator=g).to(dev)
        f1r, f2r = f1.bfloat16().float(), f2.bfloat16().float()
        L = 4 if min(H, W) >= 8 else 3
        try:
            ref = R.CorrBlock(f1r, f2r, L, 4, storage="fp32", math="fp32")
            tc = R.CorrBlock(f1, f2, L, 4, storage=storage, math="bf16")
            torch.cuda.synchronize()
        except Exception as e:
            print(f"shape {(B,D,H,W)} {storage}: EXCEPTION {type(e).__name__}: {e}")
            raise
        for l in range(L):
            a = tc.corr_pyramid[l][:, 0].float()
            b = ref.corr_pyramid[l][:, 0]
            err = (a - b).abs()
            rms = b.pow(2).mean().sqrt().item()
            print(f"shape {(B,D,H,W)} {storage} level {l}: max_err={err.max().item():.3e} mean_err={err.mean().item():.3e} rms={rms:.3e} "
                  f"nan={torch.isnan(a).sum().item()} zeros={(a==0).float().mean().item():.3f}")
            if err.max().item() > 0.05 * rms and l == 0:
                N = H * W
                e = err.view(B, N, H, W)
                bad = (e > 0.05 * rms)
                print("   bad fraction by query%128 (first 8 buckets of 16):", [round(bad[:, i::128].float().mean().item(), 3) for i in range(0, 128, 16)])
                print("   bad fraction by target row y:", [round(bad[:, :, y].float().mean().item(), 3) for y in range(min(H, 16))])
                print("   bad fraction by target col x (step 4):", [round(bad[:, :, :, x].float().mean().item(), 3) for x in range(0, min(W, 64), 4)])
                print("   sample got/ref:", a.view(B, N, H, W)[0, 0, 0, :8].tolist(), b.view(B, N, H, W)[0, 0, 0, :8].tolist())
print("tc_debug done")
