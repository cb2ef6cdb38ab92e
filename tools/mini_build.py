import numpy as np, torch, sys
sys.path.insert(0, '.')
import raft_stir_b200 as R
from oracle import corr_oracle as O
rng = np.random.default_rng(0)
shape = (1, 64, 16, 32)
f1 = rng.standard_normal(shape).astype(np.float32); f2 = rng.standard_normal(shape).astype(np.float32)
blk = R.CorrBlock(torch.from_numpy(f1).cuda(), torch.from_numpy(f2).cuda(), 4, 4, storage="bf16", math="bf16")
torch.cuda.synchronize()
pyr = O.corr_pyramid(f1, f2, 4)
for l, v in enumerate(blk.corr_pyramid):
    err = np.abs(v[:, 0].float().cpu().numpy() - pyr[l]).max()
    print("level", l, "max err", err)
