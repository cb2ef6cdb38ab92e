"""Freeze outputs of the reference's flow upsampling into tests/golden/upsample.npz.

Run in the build container (where /root/reference is mounted):

    python oracle/make_golden_upsample.py

Imports core/raft.py and core/utils/utils.py unmodified and calls RAFT.upsample_flow
(core/raft.py:72-83, a method that never touches `self`) and upflow8 (core/utils/utils.py:80-82) on
seeded inputs; stores inputs, outputs and the autograd gradients for a seeded output gradient.
The .npz is committed; this script is what made it.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

REF = os.environ.get("RAFT_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def main():
    core = os.path.join(REF, "core")
    if not os.path.isdir(core):
        raise SystemExit(f"reference not found at {REF}")
    sys.path.insert(0, core)
    from raft import RAFT            # noqa: E402  (the reference's core/raft.py)
    from utils.utils import upflow8  # noqa: E402

    data = {"torch_version": np.array(torch.__version__)}
    for name, (B, H, W, seed) in {"a": (2, 5, 7, 21), "b": (1, 3, 34, 22)}.items():
        g = torch.Generator().manual_seed(seed)
        flow = (4.0 * torch.randn(B, 2, H, W, generator=g)).requires_grad_(True)
        mask = (2.0 * torch.randn(B, 576, H, W, generator=g)).requires_grad_(True)
        gout = torch.randn(B, 2, 8 * H, 8 * W, generator=g)
        up = RAFT.upsample_flow(None, flow, mask)
        dflow, dmask = torch.autograd.grad(up, (flow, mask), gout)
        flow8 = flow.detach().clone().requires_grad_(True)
        up8 = upflow8(flow8)
        (dflow8,) = torch.autograd.grad(up8, (flow8,), gout)
        for k, v in (("flow", flow), ("mask", mask), ("gout", gout), ("up", up), ("dflow", dflow), ("dmask", dmask),
                     ("up8", up8), ("dflow8", dflow8)):
            data[f"{name}_{k}"] = v.detach().numpy().copy()
    path = os.path.join(OUT, "upsample.npz")
    np.savez_compressed(path, **data)
    print(f"wrote {path}: {os.path.getsize(path) / 1e6:.2f} MB")


if __name__ == "__main__":
    main()
