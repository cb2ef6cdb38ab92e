#!/usr/bin/env python3
"""TEMP diagnosis: build with the current kernel choice and dump levels; parent compares with RC_BUILD_PAIR=0 child."""
import os, subprocess, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_stir_b200 as R
CASES = [(2, 32, 17, 22, 4), (1, 256, 55, 128, 4), (2, 256, 46, 62, 4)]
def run():
    dev = torch.device("cuda:0"); out = {}
    for ci, (B, D, H, W, L) in enumerate(CASES):
        g = torch.Generator().manual_seed(ci)
        f1 = torch.randn(B, D, H, W, generator=g).to(dev); f2 = torch.randn(B, D, H, W, generator=g).to(dev)
        blk = R.CorrBlock(f1, f2, L, 4)
        for l, lv in enumerate(blk.corr_pyramid):
            out[f"c{ci}l{l}"] = lv.float().cpu().numpy()
    return out
if __name__ == "__main__":
    if len(sys.argv) > 1:
        np.savez(sys.argv[1], **run()); sys.exit(0)
    mine = run()
    subprocess.run([sys.executable, __file__, "/tmp/single.npz"], env=dict(os.environ, RC_BUILD_PAIR="0"), check=True)
    ref = np.load("/tmp/single.npz")
    for k in mine:
        a, b = mine[k], ref[k]
        bad = ~((a == b) | (np.isnan(a) & np.isnan(b)))
        print(k, a.shape, "mismatch", int(bad.sum()), "/", bad.size)
        if bad.any() and k.endswith(("l0", "l1")):
            bq = bad.reshape(bad.shape[0], -1).any(1)
            print("   queries bad:", int(bq.sum()), "/", bq.size, " first bad queries:", np.nonzero(bq)[0][:10])
            q = np.nonzero(bq)[0][0]
            print("   bad map of that query (rows x cols):"); m = bad[q, 0]
            for row in m[:17]: print("   ", "".join("X" if v else "." for v in row[:64]))
            print("   got row0:", a[q, 0, 0, :12]); print("   ref row0:", b[q, 0, 0, :12])
