#!/usr/bin/env python3
"""Freeze the reference RAFT's sub-networks for the end-to-end check (test infrastructure).

Runs only where the reference checkout is mounted (this build container):

    python oracle/make_ref_modules.py [--if-missing]

It imports core/raft.py UNMODIFIED, builds RAFT-basic with the reference's own seed
(train.py:241: torch.manual_seed(1234)), and traces the three sub-networks RAFT.forward calls
(core/raft.py:99-100 fnet, :110-114 cnet, :127-128 update_block) to TorchScript under
oracle/_ref/ (git-ignored; it travels to the GPU box with gpurun).  The traced files ARE the
reference's modules and weights; oracle/raft_harness.py re-states only the ~40 lines of control
flow around them so that the correlation block can be swapped.

It also freezes one small end-to-end golden (tests/golden/raft_e2e_small.npz): the reference's
own RAFT.forward on a seeded 1x3x160x224 pair (1/8 map 20x28: every pyramid level at least 2 wide,
the reference divides by zero below that), 6 iterations, CPU fp32 -- the pin for the harness.
"""
import argparse
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("RAFT_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")
FILES = ["raft_basic_fnet.pt", "raft_basic_cnet.pt", "raft_basic_update.pt"]
GOLDEN = os.path.join(ROOT, "tests", "golden", "raft_e2e_small.npz")
# RAFT-small (core/raft.py:29-33, :48-51: D = 128, radius 3, hidden 96 / context 64, no upsampling mask -> upflow8)
SMALL_FILES = ["raft_small_fnet.pt", "raft_small_cnet.pt", "raft_small_update.pt"]
SMALL_GOLDEN = os.path.join(ROOT, "tests", "golden", "raft_small_e2e.npz")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--if-missing", action="store_true")
    a = ap.parse_args()
    have = all(os.path.exists(os.path.join(OUT, f)) for f in FILES + SMALL_FILES)
    if a.if_missing and have and os.path.exists(GOLDEN) and os.path.exists(SMALL_GOLDEN):
        return 0
    if not os.path.isdir(os.path.join(REF, "core")):
        print(f"reference checkout not found at {REF}; nothing to do", file=sys.stderr)
        return 0
    sys.path.insert(0, os.path.join(REF, "core"))
    import warnings
    warnings.filterwarnings("ignore")
    from raft import RAFT  # the reference, unmodified

    args = argparse.Namespace(small=False, mixed_precision=False, alternate_corr=False)
    torch.manual_seed(1234)
    model = RAFT(args).eval()
    os.makedirs(OUT, exist_ok=True)

    g = torch.Generator().manual_seed(0)
    img = torch.rand(2, 3, 96, 128, generator=g) * 2 - 1
    with torch.no_grad():
        fnet = torch.jit.trace(model.fnet, img, check_trace=False)
        cnet = torch.jit.trace(model.cnet, img[:1], check_trace=False)
        net = torch.randn(1, 128, 12, 16, generator=g)
        inp = torch.randn(1, 128, 12, 16, generator=g)
        corr = torch.randn(1, 324, 12, 16, generator=g)
        flow = torch.randn(1, 2, 12, 16, generator=g)
        upd = torch.jit.trace(model.update_block, (net, inp, corr, flow), check_trace=False)
    for mod, f in zip((fnet, cnet, upd), FILES):
        mod.save(os.path.join(OUT, f))

    # the pin: the reference's own forward, CPU fp32
    g = torch.Generator().manual_seed(7)
    im1 = torch.rand(1, 3, 160, 224, generator=g) * 255
    im2 = torch.roll(im1, shifts=(2, -3), dims=(2, 3)) + 4 * torch.rand(1, 3, 160, 224, generator=g)
    with torch.no_grad():
        flow_low, flow_up = model(im1, im2, iters=6, test_mode=True)
    os.makedirs(os.path.dirname(GOLDEN), exist_ok=True)
    np.savez_compressed(GOLDEN, image1=im1.numpy().astype(np.float32), image2=im2.numpy().astype(np.float32),
                        flow_low=flow_low.numpy(), flow_up=flow_up.numpy(), iters=np.int32(6))
    print("wrote", OUT, "and", GOLDEN)

    # ---- RAFT-small, same recipe.  SmallUpdateBlock.forward returns (net, None, delta_flow) (core/update.py:112);
    # a trace cannot return None, so the wrapper drops it -- the harness puts it back.
    class _NoMask(torch.nn.Module):
        def __init__(self, blk):
            super().__init__()
            self.blk = blk

        def forward(self, net, inp, corr, flow):
            net, _, delta = self.blk(net, inp, corr, flow)
            return net, delta

    args = argparse.Namespace(small=True, mixed_precision=False, alternate_corr=False)
    torch.manual_seed(1234)
    small = RAFT(args).eval()
    g = torch.Generator().manual_seed(0)
    img = torch.rand(2, 3, 96, 128, generator=g) * 2 - 1
    with torch.no_grad():
        fnet = torch.jit.trace(small.fnet, img, check_trace=False)
        cnet = torch.jit.trace(small.cnet, img[:1], check_trace=False)
        net = torch.randn(1, 96, 12, 16, generator=g)
        inp = torch.randn(1, 64, 12, 16, generator=g)
        corr = torch.randn(1, 4 * 49, 12, 16, generator=g)
        flow = torch.randn(1, 2, 12, 16, generator=g)
        upd = torch.jit.trace(_NoMask(small.update_block), (net, inp, corr, flow), check_trace=False)
    for mod, f in zip((fnet, cnet, upd), SMALL_FILES):
        mod.save(os.path.join(OUT, f))
    g = torch.Generator().manual_seed(8)
    im1 = torch.rand(1, 3, 160, 224, generator=g) * 255
    im2 = torch.roll(im1, shifts=(-2, 3), dims=(2, 3)) + 4 * torch.rand(1, 3, 160, 224, generator=g)
    with torch.no_grad():
        flow_low, flow_up = small(im1, im2, iters=6, test_mode=True)
    np.savez_compressed(SMALL_GOLDEN, image1=im1.numpy().astype(np.float32), image2=im2.numpy().astype(np.float32),
                        flow_low=flow_low.numpy(), flow_up=flow_up.numpy(), iters=np.int32(6))
    print("wrote the RAFT-small modules and", SMALL_GOLDEN)
    return 0


if __name__ == "__main__":
    sys.exit(main())
