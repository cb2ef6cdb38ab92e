"""GPU parity of rc_build_bwd (the tensor-core backward of the volume build + pyramid, what autograd runs for
core/corr.py:13-27 and :52-60 under train.py:175), called through the C ABI on a dense random gradient pyramid and
compared with fp64 torch autograd of the reference's own calls (oracle/torch_port.py: bmm, / sqrt(D), avg_pool2d).

Tolerances, as |got - ref| <= tol * (|ref| + rms(ref)) for EVERY element, plus a bound on the rms of that ratio:
  math = RC_MATH_TF32 : 2e-3 (rms 5e-4).  One kind::tf32 pass: features rounded to tf32, gradients truncated to tf32
                        by the tensor core; each product carries ~3e-4 of independent relative error, so the worst of
                        10^6..10^7 outputs sits near 5 sigma = 1.5e-3 (measured 1.1e-3..1.5e-3), whatever the rounding.
  math = RC_MATH_F32  : 6e-5 (rms 1.2e-5).  3xTF32 removes the operand rounding; what is left is the tensor core's
                        fp32 accumulator, whose adds truncate: accumulated in one go the error grows linearly with the
                        contraction length (measured 7e-6 at 374 queries, 5e-5 at 2 852, 1.2e-4 at 7 040), so the kernel
                        accumulates 48 k-blocks at a time in TMEM and adds the partial sums in fp32 registers
                        (measured 2.5e-5 max / 6e-6 rms at every size).  The reference's backward is an fp32 SGEMM.
"""
import json
import os

import numpy as np
import pytest
import torch

from oracle import torch_port as P

pytestmark = pytest.mark.gpu

R = pytest.importorskip("raft_stir_b200")
from raft_stir_b200 import _lib  # noqa: E402

SHAPES = [
    (2, 256, 46, 62, 4),     # FlyingChairs features (BASELINE config 5)
    (1, 256, 55, 128, 4),    # Sintel features
    (2, 32, 17, 22, 4),      # the golden fixture's shape: N % 4 != 0
    (1, 320, 17, 33, 2),     # more than 256 channels: two channel tiles
    (1, 24, 16, 24, 3),      # channels not a multiple of 16
    (3, 128, 20, 28, 3),     # RAFT-small channels
]
RESULTS = {}


def dev():
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but torch.cuda.is_available() is False")
    return torch.device("cuda:0")


def reference(f1, f2, grads, L):
    a1 = f1.double().requires_grad_(True)
    a2 = f2.double().requires_grad_(True)
    B, D, H, W = f1.shape
    N = H * W
    vol = torch.bmm(a1.reshape(B, D, N).transpose(1, 2), a2.reshape(B, D, N)) / np.sqrt(D)
    level = vol.reshape(B * N, 1, H, W)
    loss = (level[:, 0] * grads[0].double()).sum()
    for l in range(1, L):
        level = torch.nn.functional.avg_pool2d(level, 2, stride=2)
        loss = loss + (level[:, 0] * grads[l].double()).sum()
    d1, d2 = torch.autograd.grad(loss, [a1, a2])
    return d1, d2


def run_bwd(f1, f2, grads, L, math):
    B, D, H, W = f1.shape
    glay = _lib.pyramid_layout(B, H, W, L, _lib.RC_VOL_F32)
    gbuf = torch.zeros(int(glay.total_elems), device=f1.device)
    for l in range(L):
        _lib.level_store(gbuf, glay, l, grads[l])
    lib = _lib.lib()
    ws_bytes = int(lib.rc_build_bwd_workspace_bytes(B, D, H, W, L))
    assert ws_bytes > 0
    ws = torch.full((ws_bytes,), 0xFF, dtype=torch.uint8, device=f1.device)   # NaN-poisoned
    d1 = torch.full_like(f1, float("nan"))
    d2 = torch.full_like(f2, float("nan"))
    _lib.check(lib.rc_build_bwd(gbuf.data_ptr(), f1.data_ptr(), f2.data_ptr(), B, D, H, W, L, math, d1.data_ptr(),
                                d2.data_ptr(), ws.data_ptr(), ws_bytes, torch.cuda.current_stream().cuda_stream),
               "rc_build_bwd")
    torch.cuda.synchronize()
    return d1, d2


def rel_err(got, ref):
    """(max, rms) over all elements of |got - ref| / (|ref| + rms(ref))"""
    ref = ref.double()
    rms = ref.pow(2).mean().sqrt().item() + 1e-30
    e = (got.double() - ref).abs() / (ref.abs() + rms)
    return e.max().item(), e.pow(2).mean().sqrt().item()


@pytest.mark.parametrize("math", ["tf32", "fp32"])
@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "x".join(map(str, s)))
def test_build_bwd_matches_autograd(shape, math):
    B, D, H, W, L = shape
    g = torch.Generator().manual_seed(17 + B + D + H)
    f1 = torch.randn(B, D, H, W, generator=g).to(dev())
    f2 = torch.randn(B, D, H, W, generator=g).to(dev())
    grads = [torch.randn(B * H * W, H >> l, W >> l, generator=g).to(dev()) for l in range(L)]
    r1, r2 = reference(f1, f2, grads, L)
    d1, d2 = run_bwd(f1, f2, grads, L, _lib.RC_MATH_TF32 if math == "tf32" else _lib.RC_MATH_F32)
    assert torch.isfinite(d1).all() and torch.isfinite(d2).all()
    (e1, s1), (e2, s2) = rel_err(d1, r1), rel_err(d2, r2)
    RESULTS[f"{'x'.join(map(str, shape))}/{math}"] = {"dfmap1_max": e1, "dfmap2_max": e2, "dfmap1_rms": s1, "dfmap2_rms": s2}
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if os.path.isdir(out):
        json.dump(RESULTS, open(os.path.join(out, "build_bwd_errors.json"), "w"), indent=1)
    tol, tol_rms = (2e-3, 5e-4) if math == "tf32" else (6e-5, 1.2e-5)
    assert e1 <= tol and s1 <= tol_rms, f"dfmap1 {shape} {math}: max {e1:.3e} (<= {tol:g}), rms {s1:.3e} (<= {tol_rms:g})"
    assert e2 <= tol and s2 <= tol_rms, f"dfmap2 {shape} {math}: max {e2:.3e} (<= {tol:g}), rms {s2:.3e} (<= {tol_rms:g})"


def test_build_bwd_sparse_gradient_and_padding():
    """A gradient that is zero except for a few entries (what one lookup leaves behind), poisoned workspace: the result
    must not depend on what the workspace held."""
    B, D, H, W, L = 1, 64, 23, 31, 4
    g = torch.Generator().manual_seed(5)
    f1 = torch.randn(B, D, H, W, generator=g).to(dev())
    f2 = torch.randn(B, D, H, W, generator=g).to(dev())
    grads = []
    for l in range(L):
        t = torch.zeros(B * H * W, H >> l, W >> l)
        idx = torch.randint(0, t.numel(), (200,), generator=g)
        t.view(-1)[idx] = torch.randn(200, generator=g)
        grads.append(t.to(dev()))
    r1, r2 = reference(f1, f2, grads, L)
    d1, d2 = run_bwd(f1, f2, grads, L, _lib.RC_MATH_F32)
    assert rel_err(d1, r1)[0] <= 2e-5 and rel_err(d2, r2)[0] <= 2e-5
