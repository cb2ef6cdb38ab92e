"""The tensor-core build has two kernels: CTA pairs (tcgen05 cta_group::2, the default) and the single-CTA kernel
(RC_BUILD_PAIR=0 in the environment, read once per process).  Every other GPU test exercises the pairs; this one runs
the single-CTA kernel in a child process and requires the two to agree BIT FOR BIT: both issue the same MMAs over the
same k-blocks in the same order into fp32 accumulators, and both epilogues pool and round the same values."""
import hashlib
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = [  # B, D, H, W, levels, storage, math
    (3, 256, 55, 128, 4, "bf16", "bf16"),
    (2, 256, 46, 62, 4, "bf16", "bf16"),      # odd number of query blocks (23): the last pair's second CTA is idle
    (1, 128, 27, 45, 4, "fp16", "fp16"),
    (2, 64, 17, 22, 3, "fp32", "bf16"),
    (1, 256, 33, 40, 4, "fp32", "tf32"),
    (2, 96, 24, 40, 4, "bf16", "tf32"),
]


def digests():
    """sha256 of every pyramid level of every case, as built by this process's kernel choice"""
    import raft_stir_b200 as R
    dev = torch.device("cuda:0")
    out = []
    for B, D, H, W, L, storage, math in CASES:
        g = torch.Generator().manual_seed(B * 1000 + D + H * W)
        f1 = torch.randn(B, D, H, W, generator=g).to(dev)
        f2 = torch.randn(B, D, H, W, generator=g).to(dev)
        blk = R.CorrBlock(f1, f2, L, 4, storage=storage, math=math)
        row = [hashlib.sha256(lv.contiguous().cpu().view(torch.uint8).numpy().tobytes()).hexdigest() for lv in blk.corr_pyramid]
        p1, p2 = R.pack_features(f1, f2, torch.float16 if math == "fp16" else torch.bfloat16)
        if math != "tf32" and D % 64 == 0:          # the NHWC-input build (4-D TMA boxes of half blocks for a pair)
            nb = R.CorrBlock(p1, p2, L, 4, storage=storage)
            row += [hashlib.sha256(lv.contiguous().cpu().view(torch.uint8).numpy().tobytes()).hexdigest() for lv in nb.corr_pyramid]
        out.append(row)
    return out


def test_single_cta_kernel_equals_cta_pairs():
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but torch.cuda.is_available() is False")
    assert os.environ.get("RC_BUILD_PAIR", "1") != "0", "this test expects the default (CTA pairs) in the parent"
    mine = digests()
    env = dict(os.environ, RC_BUILD_PAIR="0", PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    r = subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    theirs = json.loads(r.stdout.strip().splitlines()[-1])
    for case, a, b in zip(CASES, mine, theirs):
        assert a == b, f"{case}: the single-CTA build differs from the CTA-pair build"


if __name__ == "__main__":
    print(json.dumps(digests()))
