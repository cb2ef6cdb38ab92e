"""Pin the numpy oracle (oracle/corr_oracle.py) and the torch-op port
(oracle/torch_port.py) against tensors frozen from the reference's own
CorrBlock (tests/golden/, made by oracle/make_golden.py)."""
import numpy as np
import pytest
import torch

from oracle import corr_oracle as O
from oracle import torch_port as P

KINDS = ["rand", "int", "far", "edge"]


def test_pyramid_matches_reference(golden):
    pyr = O.corr_pyramid(golden["fmap1"], golden["fmap2"], 4)
    for l, p in enumerate(pyr):
        ref = golden[f"pyr{l}"]
        assert p.shape == ref.shape
        np.testing.assert_allclose(p, ref, rtol=1e-5, atol=2e-5)


def test_pyramid_dims_floor(golden):
    H, W = golden["fmap1"].shape[-2:]
    dims = O.pyramid_dims(H, W, 4)
    for l, (h, w) in enumerate(dims):
        assert golden[f"pyr{l}"].shape[1:] == (h, w)


@pytest.mark.parametrize("kind", KINDS)
def test_lookup_matches_reference(golden, kind):
    # use the reference's own pyramid so this isolates the lookup
    pyr = [golden[f"pyr{l}"] for l in range(4)]
    out = O.window_lookup(pyr, golden[f"coords_{kind}"], int(golden["radius"]))
    ref = golden[f"out_{kind}"]
    assert out.shape == ref.shape and out.dtype == np.float32
    # the reference's normalise/denormalise round trip moves coordinates by ~1 ulp
    np.testing.assert_allclose(out, ref, rtol=1e-4, atol=2e-5)
    if kind == "far":
        assert not out.any()


def test_lookup_channel_order(golden):
    """channel l*n*n + i*n + j samples (x/2^l + i - r, y/2^l + j - r)."""
    r = int(golden["radius"])
    n = 2 * r + 1
    pyr = [golden[f"pyr{l}"] for l in range(4)]
    coords = golden["coords_int"]
    B, _, H, W = coords.shape
    out = golden["out_int"]
    rng = np.random.default_rng(0)
    for _ in range(200):
        b, y, x = rng.integers(B), rng.integers(H), rng.integers(W)
        i, j = rng.integers(n), rng.integers(n)
        X = int(coords[b, 0, y, x]) + i - r
        Y = int(coords[b, 1, y, x]) + j - r
        q = (b * H + y) * W + x
        want = pyr[0][q, Y, X] if (0 <= X < W and 0 <= Y < H) else 0.0
        assert abs(out[b, i * n + j, y, x] - want) < 1e-5


def test_lookup_backward_matches_autograd(golden):
    r = int(golden["radius"])
    pyr = O.corr_pyramid(golden["fmap1"].astype(np.float64), golden["fmap2"].astype(np.float64), 4)
    dpyr, dcoords = O.window_lookup_backward(pyr, golden["coords_rand"], r, golden["grad_out"])
    for l in range(1, 4):
        np.testing.assert_allclose(dpyr[l], golden[f"dpyr{l}"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(dpyr[0].sum(axis=2), golden["dpyr0_rowsum"], rtol=1e-4, atol=1e-4)
    np.testing.assert_allclose(dcoords, golden["dcoords"], rtol=2e-3, atol=2e-3)
    d1, d2 = O.pyramid_backward(dpyr, golden["fmap1"], golden["fmap2"])
    np.testing.assert_allclose(d1, golden["dfmap1"], rtol=1e-3, atol=1e-4)
    np.testing.assert_allclose(d2, golden["dfmap2"], rtol=1e-3, atol=1e-4)


def test_alt_path_equals_default_path(golden):
    """AlternateCorrBlock is mathematically the default path;
    the reference extension is CUDA-only, so it is anchored on CorrBlock."""
    r = int(golden["radius"])
    f1, f2 = golden["fmap1"][:1], golden["fmap2"][:1]
    for kind in ("rand", "edge"):
        c = golden[f"coords_{kind}"][:1]
        out = O.alt_block_lookup(f1, f2, c, 4, r)
        np.testing.assert_allclose(out, golden[f"out_{kind}"][:1], rtol=1e-4, atol=3e-5)


def test_alt_backward_is_transpose_of_forward():
    """<alt_forward(f1,f2), g> differentiated numerically matches alt_backward;
    coords_grad is identically zero as in the reference kernel."""
    rng = np.random.default_rng(3)
    B, H, W, C, r = 1, 5, 6, 4, 2
    f1 = rng.standard_normal((B, H, W, C)).astype(np.float32)
    f2 = rng.standard_normal((B, 4, 5, C)).astype(np.float32)
    coords = (rng.uniform(-1, 5, (B, 1, H, W, 2))).astype(np.float32)
    g = rng.standard_normal((B, 1, (2 * r + 1) ** 2, H, W)).astype(np.float32)
    g1, g2, gc = O.alt_backward(f1, f2, coords, g, r)
    assert not gc.any()
    eps = 1e-2
    for arr, grad, which in ((f1, g1, 0), (f2, g2, 1)):
        for _ in range(5):
            idx = tuple(rng.integers(s) for s in arr.shape)
            a = arr.copy(); a[idx] += eps
            b = arr.copy(); b[idx] -= eps
            fa = O.alt_forward(a if which == 0 else f1, a if which == 1 else f2, coords, r)
            fb = O.alt_forward(b if which == 0 else f1, b if which == 1 else f2, coords, r)
            num = ((fa.astype(np.float64) - fb) * g).sum() / (2 * eps)
            assert abs(num - grad[idx]) < 2e-2 * max(1.0, abs(num))


@pytest.mark.parametrize("kind", KINDS)
def test_torch_port_matches_reference(golden, kind):
    f1 = torch.from_numpy(golden["fmap1"])
    f2 = torch.from_numpy(golden["fmap2"])
    pyr = P.build_pyramid(f1, f2, 4)
    for l, p in enumerate(pyr):
        np.testing.assert_allclose(p[:, 0].numpy(), golden[f"pyr{l}"], rtol=1e-5, atol=2e-5)
    out = P.lookup(pyr, torch.from_numpy(golden[f"coords_{kind}"]), int(golden["radius"]))
    np.testing.assert_allclose(out.numpy(), golden[f"out_{kind}"], rtol=1e-5, atol=1e-5)


def test_coords_grid_convention():
    g = O.coords_grid(2, 3, 5)
    assert g.shape == (2, 2, 3, 5)
    assert g[0, 0, 1, 4] == 4 and g[0, 1, 1, 4] == 1


# ------------------------------------------------------------------ flow upsampling (core/raft.py:72-83, utils.py:80-82)
@pytest.fixture(scope="module")
def up_golden():
    import os
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "upsample.npz"))


@pytest.mark.parametrize("case", ["a", "b"])
def test_upsample_oracle_matches_reference(up_golden, case):
    from oracle import upsample_oracle as U
    g = up_golden
    flow, mask, gout = (g[f"{case}_{k}"].astype(np.float64) for k in ("flow", "mask", "gout"))
    np.testing.assert_allclose(U.upsample_flow(flow, mask), g[f"{case}_up"], rtol=1e-5, atol=1e-4)
    dflow, dmask = U.upsample_flow_backward(flow, mask, gout)
    np.testing.assert_allclose(dflow, g[f"{case}_dflow"], rtol=1e-5, atol=1e-4)
    np.testing.assert_allclose(dmask, g[f"{case}_dmask"], rtol=1e-5, atol=1e-4)
    # torch derives the source position in fp32: 1e-5 relative on values of order 100
    np.testing.assert_allclose(U.upflow8(flow), g[f"{case}_up8"], rtol=1e-5, atol=5e-4)
    np.testing.assert_allclose(U.upflow8_backward(flow.shape, gout), g[f"{case}_dflow8"], rtol=1e-5, atol=5e-4)


def test_upsample_is_a_convex_combination():
    """A constant flow upsamples to 8x that constant in the interior (the weights sum to one); at the border the
    zero padding of F.unfold pulls the result towards zero."""
    from oracle import upsample_oracle as U
    rng = np.random.default_rng(3)
    flow = np.full((1, 2, 4, 5), 1.5)
    up = U.upsample_flow(flow, rng.normal(0, 2, (1, 576, 4, 5)))
    np.testing.assert_allclose(up[:, :, 8:-8, 8:-8], 12.0, rtol=1e-12)
    assert (np.abs(up) <= 12.0 + 1e-9).all()


# ------------------------------------------------------------------ BASELINE shapes (subsampled goldens)
@pytest.mark.parametrize("name", ["sintel", "chairs"])
def test_oracle_matches_reference_at_baseline_shapes(name):
    """oracle/make_golden_shapes.py froze the reference CorrBlock's lookups at 256 random queries of the Sintel and
    FlyingChairs feature shapes; the numpy oracle, evaluated for those queries only, must reproduce them (1e-4)."""
    import os
    from oracle import make_golden_shapes as G
    gold = dict(np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", f"shape_{name}.npz")))
    inp = G.make_inputs(name)
    np.testing.assert_allclose(G.checksums(inp), gold["checksums"], rtol=1e-12, atol=1e-9)
    H, W, _ = G.SHAPES[name]
    q = gold["queries"]
    f1 = inp["f1"].numpy().reshape(G.D, H * W)[:, q].astype(np.float64)      # [D, 256]
    f2 = inp["f2"].numpy().reshape(G.D, H * W).astype(np.float64)
    vol = (f1.T @ f2 / np.sqrt(np.float64(G.D))).reshape(len(q), H, W)
    pyr = [vol]
    for _ in range(G.LEVELS - 1):
        pyr.append(O._pool2(pyr[-1]))
    for l in range(1, G.LEVELS):
        np.testing.assert_allclose(pyr[l][:16], gold[f"pyr{l}"], rtol=1e-4, atol=2e-5, err_msg=f"{name} level {l}")
    for kind in ("rand", "int"):
        c = inp[f"coords_{kind}"].numpy().reshape(1, 2, H * W)[:, :, q].reshape(1, 2, len(q), 1).astype(np.float64)
        out = O.window_lookup(pyr, c, G.RADIUS)                                # [1, 324, 256, 1]
        np.testing.assert_allclose(out[0, :, :, 0], gold[f"out_{kind}"], rtol=1e-4, atol=2e-5, err_msg=f"{name} lookup {kind}")
