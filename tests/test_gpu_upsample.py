"""GPU parity tests of the flow upsampling kernels (through the C ABI, via the ctypes binding) against the
fixtures frozen from the reference's RAFT.upsample_flow / upflow8 (tests/golden/upsample.npz) and against the numpy
oracle on other shapes.  fp32 arithmetic: 1e-5 relative to the value plus 1e-4 absolute on flows of order 100
(exp and summation order differ from torch's kernels)."""
import os

import numpy as np
import pytest
import torch

from oracle import upsample_oracle as U

pytestmark = pytest.mark.gpu

R = pytest.importorskip("raft_stir_b200")

RTOL, ATOL = 1e-5, 1e-4


def dev():
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but torch.cuda.is_available() is False")
    return torch.device("cuda:0")


def t(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev())


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "upsample.npz"))


@pytest.mark.parametrize("case", ["a", "b"])
def test_upsample_flow_matches_reference(golden, case):
    g = golden
    flow, mask = t(g[f"{case}_flow"]).requires_grad_(True), t(g[f"{case}_mask"]).requires_grad_(True)
    up = R.upsample_flow(flow, mask)
    np.testing.assert_allclose(up.detach().cpu().numpy(), g[f"{case}_up"], rtol=RTOL, atol=ATOL)
    dflow, dmask = torch.autograd.grad(up, (flow, mask), t(g[f"{case}_gout"]))
    np.testing.assert_allclose(dflow.cpu().numpy(), g[f"{case}_dflow"], rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(dmask.cpu().numpy(), g[f"{case}_dmask"], rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize("case", ["a", "b"])
def test_upflow8_matches_reference(golden, case):
    g = golden
    flow = t(g[f"{case}_flow"]).requires_grad_(True)
    up = R.upflow8(flow)
    np.testing.assert_allclose(up.detach().cpu().numpy(), g[f"{case}_up8"], rtol=RTOL, atol=ATOL)
    (dflow,) = torch.autograd.grad(up, (flow,), t(g[f"{case}_gout"]))
    np.testing.assert_allclose(dflow.cpu().numpy(), g[f"{case}_dflow8"], rtol=RTOL, atol=5e-4)


@pytest.mark.parametrize("shape", [(1, 1, 1), (2, 3, 31), (1, 7, 32), (3, 6, 33), (1, 55, 128), (2, 46, 62)])
def test_upsample_flow_matches_oracle(shape):
    """Ragged widths around the 32-pixel tile, a single pixel, and the Sintel / Chairs feature-map sizes."""
    B, H, W = shape
    rng = np.random.default_rng(B * 1000 + H * 10 + W)
    flow = rng.normal(0, 5, (B, 2, H, W)).astype(np.float32)
    mask = rng.normal(0, 3, (B, 576, H, W)).astype(np.float32)
    gout = rng.standard_normal((B, 2, 8 * H, 8 * W)).astype(np.float32)
    ft, mt = t(flow).requires_grad_(True), t(mask).requires_grad_(True)
    up = R.upsample_flow(ft, mt)
    f64, m64, g64 = flow.astype(np.float64), mask.astype(np.float64), gout.astype(np.float64)
    np.testing.assert_allclose(up.detach().cpu().numpy(), U.upsample_flow(f64, m64), rtol=RTOL, atol=ATOL)
    dflow, dmask = torch.autograd.grad(up, (ft, mt), t(gout))
    rf, rm = U.upsample_flow_backward(f64, m64, g64)
    np.testing.assert_allclose(dflow.cpu().numpy(), rf, rtol=RTOL, atol=1e-3)     # sums of 576 products of order 10
    np.testing.assert_allclose(dmask.cpu().numpy(), rm, rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize("shape", [(2, 3, 31), (1, 7, 32), (1, 55, 128)])
def test_upsample_flow_fp16_mask(shape):
    """A half-precision mask (what fp16 autocast hands over) is read directly: same result as the fp32 kernel on the
    mask converted by torch, bit for bit (the conversion to fp32 is exact and everything after it is the same code)."""
    B, H, W = shape
    rng = np.random.default_rng(W)
    flow = t(rng.normal(0, 5, (B, 2, H, W)))
    mask16 = t(rng.normal(0, 3, (B, 576, H, W))).half()
    with torch.no_grad():
        got = R.upsample_flow(flow, mask16)
        want = R.upsample_flow(flow, mask16.float())
    assert got.dtype == torch.float32 and torch.equal(got, want)
    # with autograd the mask is taken through fp32 and gradients come back in fp32
    m = mask16.float().requires_grad_(True)
    up = R.upsample_flow(flow, m)
    (dm,) = torch.autograd.grad(up, (m,), torch.ones_like(up))
    assert dm.dtype == torch.float32 and torch.isfinite(dm).all()


def test_upsample_flow_extreme_logits():
    """softmax with logits of +-80 (one neighbour takes all the weight) stays finite and picks that neighbour."""
    B, H, W = 1, 4, 6
    rng = np.random.default_rng(9)
    flow = rng.normal(0, 5, (B, 2, H, W)).astype(np.float32)
    mask = np.full((B, 9, 64, H, W), -80.0, np.float32)
    mask[:, 4] = 80.0                                           # the centre pixel
    up = R.upsample_flow(t(flow), t(mask.reshape(B, 576, H, W))).cpu().numpy()
    assert np.isfinite(up).all()
    np.testing.assert_allclose(up, 8 * np.repeat(np.repeat(flow, 8, axis=2), 8, axis=3), rtol=1e-6)


@pytest.mark.parametrize("shape", [(1, 2, 1, 1), (2, 2, 5, 9), (1, 3, 46, 62)])
def test_upflow8_matches_oracle(shape):
    B, C, H, W = shape
    rng = np.random.default_rng(H * 10 + W)
    flow = rng.normal(0, 5, (B, C, H, W)).astype(np.float32)
    gout = rng.standard_normal((B, C, 8 * H, 8 * W)).astype(np.float32)
    ft = t(flow).requires_grad_(True)
    up = R.upflow8(ft)
    np.testing.assert_allclose(up.detach().cpu().numpy(), U.upflow8(flow.astype(np.float64)), rtol=RTOL, atol=5e-4)
    (dflow,) = torch.autograd.grad(up, (ft,), t(gout))
    np.testing.assert_allclose(dflow.cpu().numpy(), U.upflow8_backward(flow.shape, gout.astype(np.float64)),
                               rtol=RTOL, atol=2e-3)


def test_upsample_rejects_bad_input():
    d = dev()
    with pytest.raises(RuntimeError):
        R.upsample_flow(torch.zeros(1, 2, 4, 4), torch.zeros(1, 576, 4, 4))            # CPU tensors
    with pytest.raises(RuntimeError):
        R.upsample_flow(torch.zeros(1, 2, 4, 4, device=d), torch.zeros(1, 575, 4, 4, device=d))
    with pytest.raises(RuntimeError):
        R.upsample_flow(torch.zeros(1, 3, 4, 4, device=d), torch.zeros(1, 576, 4, 4, device=d))
    with pytest.raises(RuntimeError):
        R.upflow8(torch.zeros(1, 2, 4, 4, device=d), mode="nearest")
