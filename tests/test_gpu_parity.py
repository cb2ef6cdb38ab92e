"""GPU parity tests: the CUDA path (through the C ABI, via the ctypes binding)
against the CPU oracle and the fixtures frozen from the reference's CorrBlock.

Tolerances (BASELINE.json north_star): correlation values within 1e-2 relative
for bf16, 1e-4 for fp32; written next to each assert.
"""
import numpy as np
import pytest
import torch

from oracle import corr_oracle as O

pytestmark = pytest.mark.gpu

R = pytest.importorskip("raft_stir_b200")


def dev():
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but torch.cuda.is_available() is False")
    return torch.device("cuda:0")


def t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev())


def bf16_round(a):
    return torch.from_numpy(a).to(torch.bfloat16).float().numpy()


def rel_close(got, ref, rel, what, floor=None):
    """|got-ref| <= rel * (|ref| + floor): relative to the value, with a floor
    for entries near zero.  Default floor: rms(ref)."""
    got = np.asarray(got, np.float64)
    ref = np.asarray(ref, np.float64)
    assert got.shape == ref.shape, (what, got.shape, ref.shape)
    rms = float(np.sqrt(np.mean(ref ** 2))) + 1e-30
    err = np.abs(got - ref)
    bound = rel * (np.abs(ref) + (rms if floor is None else floor))
    bad = err > bound
    assert not bad.any(), f"{what}: {bad.sum()} / {bad.size} outside {rel:g} rel; max err {err.max():.3e}, rms {rms:.3e}"


def bf16_floor(f1, f2):
    """Floor for comparing bf16 arithmetic with the fp32 reference.  A bf16 dot
    product is accurate relative to sum_d |f1_d f2_d| / sqrt(D), not relative to
    the (cancelled) sum itself; for D >= 16 that magnitude sum is above 3x the
    volume's rms, so "1e-2 relative" is taken as 1e-2 * (|ref| + 3 * rms(volume))."""
    B, D = f1.shape[:2]
    a = f1.reshape(B, D, -1).astype(np.float64)
    b = f2.reshape(B, D, -1).astype(np.float64)
    vol_rms = np.sqrt((a ** 2).sum(1).mean() * (b ** 2).sum(1).mean() / D) / np.sqrt(D)
    return 3.0 * float(vol_rms)


KINDS = ["rand", "int", "far", "edge"]


# ------------------------------------------------------------------ build
def test_build_fp32_matches_reference(golden):
    blk = R.CorrBlock(t(golden["fmap1"]), t(golden["fmap2"]), 4, int(golden["radius"]), storage="fp32", math="fp32")
    for l, v in enumerate(blk.corr_pyramid):
        assert v.shape[1] == 1
        rel_close(v[:, 0].cpu().numpy(), golden[f"pyr{l}"], 1e-4, f"fp32 level {l}")   # 1e-4 (fp32)


@pytest.mark.parametrize("storage", ["bf16", "fp32"])
def test_build_tensor_core_matches_reference(golden, storage):
    blk = R.CorrBlock(t(golden["fmap1"]), t(golden["fmap2"]), 4, int(golden["radius"]), storage=storage, math="bf16")
    # (a) against the reference's fp32 values: 1e-2 relative (bf16)
    for l, v in enumerate(blk.corr_pyramid):
        rel_close(v[:, 0].float().cpu().numpy(), golden[f"pyr{l}"], 1e-2, f"bf16/{storage} level {l}",
                  floor=bf16_floor(golden["fmap1"], golden["fmap2"]))
    # (b) against the oracle fed the same bf16-rounded features: only storage rounding remains
    pyr = O.corr_pyramid(bf16_round(golden["fmap1"]).astype(np.float64), bf16_round(golden["fmap2"]).astype(np.float64), 4)
    tol = 2.0 ** -8 if storage == "bf16" else 1e-5
    for l, v in enumerate(blk.corr_pyramid):
        rel_close(v[:, 0].float().cpu().numpy(), pyr[l], tol, f"bf16-exact/{storage} level {l}")


def tf32_round(a):
    """fp32 -> tf32 (10 mantissa bits), round to nearest with ties away from zero: what cvt.rna.tf32.f32 does."""
    u = np.ascontiguousarray(a, np.float32).view(np.uint32).astype(np.uint64)
    u = ((u + 0x1000) & 0xFFFFE000).astype(np.uint32)
    return u.view(np.float32)


@pytest.mark.parametrize("shape", [(1, 256, 55, 128), (2, 128, 46, 62), (1, 64, 17, 33), (3, 40, 8, 9), (2, 32, 24, 30)])
def test_build_tf32_vs_oracle_shapes(shape):
    """math="tf32" (tcgen05 kind::tf32 on the fp32 features rounded to tf32), fp32 storage: equal to the oracle fed
    the same tf32-rounded features up to fp32 summation order, and within 1e-3 (2^-10-mantissa operands) of the
    unrounded fp32 oracle, relative to the magnitude a dot product of that length is accurate to."""
    B, D, H, W = shape
    rng = np.random.default_rng(H * W + D + 1)
    f1 = rng.standard_normal(shape).astype(np.float32)
    f2 = rng.standard_normal(shape).astype(np.float32)
    L = 4 if min(H, W) >= 8 else 3
    blk = R.CorrBlock(t(f1), t(f2), L, 4, storage="fp32", math="tf32")
    # (the pack's power-of-two pre-scale of fmap1 commutes with the rounding)
    pyr = O.corr_pyramid(tf32_round(f1).astype(np.float64), tf32_round(f2).astype(np.float64), L)
    for l, v in enumerate(blk.corr_pyramid):
        rel_close(v[:, 0].float().cpu().numpy(), pyr[l], 2e-5, f"tf32-exact {shape} level {l}")
    ref = O.corr_pyramid(f1, f2, L)
    for l, v in enumerate(blk.corr_pyramid):
        rel_close(v[:, 0].float().cpu().numpy(), ref[l], 1e-3, f"tf32 {shape} level {l} vs fp32", floor=bf16_floor(f1, f2))
    # and a lookup on it agrees with the fp32 block to the same class
    coords = t((O.coords_grid(B, H, W) + rng.normal(0, 3, (B, 2, H, W))).astype(np.float32))
    want = R.CorrBlock(t(f1), t(f2), L, 4, storage="fp32", math="fp32")(coords)
    rel_close(blk(coords).cpu().numpy(), want.cpu().numpy(), 1e-3, f"tf32 lookup {shape}", floor=bf16_floor(f1, f2))


def fp16_round(a):
    return np.asarray(a, np.float32).astype(np.float16).astype(np.float32)


@pytest.mark.parametrize("math", ["bf16", "fp16", "tf32", "fp32"])
@pytest.mark.parametrize("shape", [(1, 256, 55, 128), (2, 128, 19, 45), (3, 32, 8, 9)])
def test_fp16_storage_build_and_lookup(shape, math):
    """storage="fp16" (RC_VOL_F16): every level equals the oracle's value (for the operands that math sees) rounded
    once to IEEE half (2^-11 relative), the padding of edge tiles is zero, and the lookup on it matches the oracle's
    lookup on the fp16-rounded pyramid to fp32 round-off; forward and d/dcoords."""
    B, D, H, W = shape
    rng = np.random.default_rng(H * W + D + len(math))
    f1 = rng.standard_normal(shape).astype(np.float32)
    f2 = rng.standard_normal(shape).astype(np.float32)
    L = 4 if min(H, W) >= 8 else 3
    blk = R.CorrBlock(t(f1), t(f2), L, 4, storage="fp16", math=math)
    assert blk._buf.dtype == torch.float16
    rnd = {"bf16": bf16_round, "fp16": fp16_round, "tf32": tf32_round, "fp32": lambda a: a}[math]
    pyr = O.corr_pyramid(rnd(f1).astype(np.float64), rnd(f2).astype(np.float64), L)
    got = [v[:, 0].float().cpu().numpy() for v in blk.corr_pyramid]
    for l in range(L):
        # tensor-core builds pool the fp32 accumulators (one rounding per level); the CUDA-core build pools the
        # STORED level below: roundings of values several times larger than their mean carry into the mean
        tol = 2.0 ** -11 * 1.01 * (4 if math == "fp32" and l > 0 else 1)
        rel_close(got[l], pyr[l], tol, f"fp16 storage {math} {shape} level {l}")
    coords = (O.coords_grid(B, H, W) + rng.normal(0, 3, (B, 2, H, W))).astype(np.float32)
    out = blk(t(coords)).cpu().numpy()
    rel_close(out, O.window_lookup([g.astype(np.float64) for g in got], coords.astype(np.float64), 4), 2e-5,
              f"lookup on fp16 storage {math} {shape}")
    # d/dcoords through the fp16 volume (reads the forward taps)
    c = t(coords).requires_grad_(True)
    go = torch.randn(B, L * 81, H, W, device=dev(), generator=torch.Generator(device=dev()).manual_seed(1))
    (dc,) = torch.autograd.grad(blk(c), (c,), go)
    ref_dc = O.window_lookup_backward([g.astype(np.float64) for g in got], coords.astype(np.float64), 4,
                                      go.cpu().numpy().astype(np.float64))[1]
    rel_close(dc.cpu().numpy(), ref_dc, 1e-4, f"d/dcoords on fp16 storage {math} {shape}")


@pytest.mark.parametrize("shape", [(1, 256, 55, 128), (2, 128, 46, 62), (1, 64, 17, 33), (3, 32, 8, 9), (2, 32, 24, 30),
                                   (1, 32, 9, 240)])
@pytest.mark.parametrize("storage", ["bf16", "fp32"])
def test_build_tensor_core_vs_oracle_shapes(shape, storage):
    B, D, H, W = shape
    rng = np.random.default_rng(H * W + D)
    f1 = rng.standard_normal(shape).astype(np.float32)
    f2 = rng.standard_normal(shape).astype(np.float32)
    L = 4 if min(H, W) >= 8 else 3
    blk = R.CorrBlock(t(f1), t(f2), L, 4, storage=storage, math="bf16")
    pyr = O.corr_pyramid(bf16_round(f1), bf16_round(f2), L)
    tol = 2.0 ** -8 if storage == "bf16" else 2e-5
    for l, v in enumerate(blk.corr_pyramid):
        rel_close(v[:, 0].float().cpu().numpy(), pyr[l], tol, f"shape {shape} {storage} level {l}")
    # 1e-2 relative against the unrounded fp32 oracle as well
    ref = O.corr_pyramid(f1, f2, L)
    for l, v in enumerate(blk.corr_pyramid):
        rel_close(v[:, 0].float().cpu().numpy(), ref[l], 1e-2, f"shape {shape} {storage} level {l} vs fp32",
                  floor=bf16_floor(f1, f2))


@pytest.mark.parametrize("math", ["bf16", "fp16", "tf32", "fp32"])
@pytest.mark.parametrize("storage", ["bf16", "fp16", "fp32"])
def test_padding_columns_of_edge_tiles_are_zero(storage, math):
    """rc_layout's invariant: inside rows < H_l, tile columns >= W_l hold zeros
    (the lookup reads whole 16-byte tile rows and does not mask columns)."""
    from raft_stir_b200 import _lib
    rng = np.random.default_rng(11)
    for shape in ((2, 32, 24, 30), (1, 16, 46, 62), (1, 16, 17, 33)):
        B, D, H, W = shape
        blk = R.CorrBlock(t(rng.standard_normal(shape).astype(np.float32)),
                          t(rng.standard_normal(shape).astype(np.float32)), 4, 4, storage=storage, math=math)
        for l in range(4):
            v = _lib._tiled(blk._buf, blk.layout, l)
            q, ty, th, tx, tw = v.shape
            full = v.reshape(q, ty * th, tx * tw).float()
            Hl, Wl = blk.layout.H[l], blk.layout.W[l]
            assert torch.isfinite(full[:, :Hl]).all(), f"{shape} level {l}"
            assert not full[:, :Hl, Wl:].any().item(), f"{shape} {storage}/{math} level {l}: nonzero padding columns"


def test_build_fp32_vs_oracle_ragged():
    rng = np.random.default_rng(5)
    f1 = rng.standard_normal((2, 24, 19, 23)).astype(np.float32)
    f2 = rng.standard_normal((2, 24, 19, 23)).astype(np.float32)
    blk = R.CorrBlock(t(f1), t(f2), 4, 3, storage="fp32", math="fp32")
    for l, (v, ref) in enumerate(zip(blk.corr_pyramid, O.corr_pyramid(f1, f2, 4))):
        rel_close(v[:, 0].cpu().numpy(), ref, 1e-4, f"ragged level {l}")


def test_corr_staticmethod_shape_and_values(golden):
    vol = R.CorrBlock.corr(t(golden["fmap1"]), t(golden["fmap2"]))
    B, D, H, W = golden["fmap1"].shape
    assert tuple(vol.shape) == (B, H, W, 1, H, W)
    rel_close(vol.reshape(B * H * W, H, W).cpu().numpy(), golden["pyr0"], 1e-4, "CorrBlock.corr")


# ------------------------------------------------------------------ lookup
@pytest.mark.parametrize("kind", KINDS)
def test_lookup_fp32_matches_reference(golden, kind):
    blk = R.CorrBlock(t(golden["fmap1"]), t(golden["fmap2"]), 4, int(golden["radius"]), storage="fp32", math="fp32")
    out = blk(t(golden[f"coords_{kind}"]))
    ref = golden[f"out_{kind}"]
    assert out.dtype == torch.float32 and out.is_contiguous() and tuple(out.shape) == ref.shape
    rel_close(out.cpu().numpy(), ref, 1e-4, f"lookup fp32 {kind}")   # 1e-4 (fp32)
    if kind == "far":
        assert not out.any().item()


@pytest.mark.parametrize("kind", KINDS)
def test_lookup_bf16_matches_reference(golden, kind):
    blk = R.CorrBlock(t(golden["fmap1"]), t(golden["fmap2"]), 4, int(golden["radius"]))   # defaults: bf16/bf16
    out = blk(t(golden[f"coords_{kind}"]))
    rel_close(out.cpu().numpy(), golden[f"out_{kind}"], 1e-2, f"lookup bf16 {kind}",   # 1e-2 (bf16)
              floor=bf16_floor(golden["fmap1"], golden["fmap2"]))


@pytest.mark.parametrize("radius", [1, 2, 3, 4])
@pytest.mark.parametrize("storage", ["bf16", "fp32"])
def test_lookup_isolated_from_build(radius, storage):
    """Lookup alone: pyramid values written by torch into the library's layout,
    compared with the oracle's lookup on exactly those values."""
    from raft_stir_b200 import _lib
    B, H, W, L = 2, 21, 37, 4
    rng = np.random.default_rng(radius)
    blk = R.CorrBlock(t(np.zeros((B, 8, H, W), np.float32)), t(np.zeros((B, 8, H, W), np.float32)), L, radius,
                      storage=storage, math="fp32")
    pyr = []
    for l, v in enumerate(blk.corr_pyramid):
        vals = torch.from_numpy(rng.standard_normal(tuple(v.shape)).astype(np.float32)).to(v.dtype)
        _lib.level_store(blk._buf, blk.layout, l, vals[:, 0])
        pyr.append(vals.float().numpy()[:, 0])
    coords = (O.coords_grid(B, H, W) + rng.normal(0, 6, (B, 2, H, W))).astype(np.float32)
    out = blk(t(coords)).cpu().numpy()
    rel_close(out, O.window_lookup(pyr, coords, radius), 2e-5, f"isolated lookup r={radius} {storage}")


@pytest.mark.parametrize("radius", [3, 4])
@pytest.mark.parametrize("storage", ["bf16", "fp32"])
def test_lookup_fp16_output_is_the_fp32_output_rounded_once(radius, storage):
    """rc_lookup_fwd_f16 (CorrBlock(out_dtype=torch.float16)): bit-identical to the fp32 result cast to fp16 by torch,
    i.e. exactly what fp16 autocast would make of the reference's output; training calls still return fp32."""
    B, D, H, W = 2, 32, 19, 45
    rng = np.random.default_rng(radius + len(storage))
    f1, f2 = (t(rng.standard_normal((B, D, H, W)).astype(np.float32)) for _ in range(2))
    coords = t((O.coords_grid(B, H, W) + rng.normal(0, 4, (B, 2, H, W))).astype(np.float32))
    math = "fp32" if storage == "fp32" else "bf16"
    ref = R.CorrBlock(f1, f2, 4, radius, storage=storage, math=math)(coords)
    blk = R.CorrBlock(f1, f2, 4, radius, storage=storage, math=math, out_dtype=torch.float16)
    with torch.no_grad():
        got = blk(coords)
    assert got.dtype == torch.float16 and got.is_contiguous() and got.shape == ref.shape
    assert torch.equal(got, ref.half())
    # with autograd in play the block answers in fp32 like the reference
    cg = coords.clone().requires_grad_(True)
    assert blk(cg).dtype == torch.float32
    with pytest.raises(RuntimeError):
        R.CorrBlock(f1, f2, 4, radius, out_dtype=torch.bfloat16)


def test_lookup_far_coords_are_zero_and_nonfinite_coords_are_nan():
    """Finite coordinates far outside the map: zeros (grid_sample's zeros padding).  NaN or infinite coordinates: NaN
    in every sample of that query and nowhere else, as the reference's CorrBlock gives on CPU (probed: NaN, +Inf, -Inf
    -> NaN; 1e30 -> 0) -- a diverging run must not be masked by finite zeros."""
    B, D, H, W = 1, 64, 16, 24
    rng = np.random.default_rng(0)
    f1 = t(rng.standard_normal((B, D, H, W)).astype(np.float32))
    f2 = t(rng.standard_normal((B, D, H, W)).astype(np.float32))
    blocks = [R.CorrBlock(f1, f2, 4, 4), R.CorrBlock(f1, f2, 4, 4, storage="fp32", math="fp32"),
              R.AlternateCorrBlock(f1, f2, 4, 4, math="fp32"), R.AlternateCorrBlock(f1, f2, 4, 4, math="bf16")]
    for blk in blocks:
        for far in (3.0e9, -1.0e4, 1.0e30):
            assert not blk(torch.full((B, 2, H, W), far, device=dev())).any().item()
        c = R.coords_grid(B, H, W, dev())
        bad = [(3, 3, 0, float("nan")), (5, 5, 1, float("inf")), (7, 20, 0, -float("inf"))]
        for y, x, ch, v in bad:
            c[0, ch, y, x] = v
        out = blk(c)
        nan_px = torch.isnan(out).any(dim=1)[0]
        want = torch.zeros(H, W, dtype=torch.bool, device=dev())
        for y, x, _, _ in bad:
            want[y, x] = True
            assert torch.isnan(out[0, :, y, x]).all(), f"{type(blk).__name__}: not every sample of the query is NaN"
        assert torch.equal(nan_px, want), f"{type(blk).__name__}: NaN leaked to other queries"


# ------------------------------------------------------------------ backward
def test_backward_matches_reference_autograd(golden):
    f1 = t(golden["fmap1"]).requires_grad_(True)
    f2 = t(golden["fmap2"]).requires_grad_(True)
    c = t(golden["coords_rand"]).requires_grad_(True)
    blk = R.CorrBlock(f1, f2, 4, int(golden["radius"]), storage="fp32", math="fp32")
    out = blk(c)
    rel_close(out.detach().cpu().numpy(), golden["out_rand"], 1e-4, "fwd under autograd")
    out.backward(t(golden["grad_out"]))
    rel_close(c.grad.cpu().numpy(), golden["dcoords"], 2e-3, "dcoords")
    rel_close(f1.grad.cpu().numpy(), golden["dfmap1"], 1e-3, "dfmap1")
    rel_close(f2.grad.cpu().numpy(), golden["dfmap2"], 1e-3, "dfmap2")


def test_lookup_backward_grad_pyramid(golden):
    """d/dpyramid of the lookup alone (levels as leaves in the reference)."""
    import ctypes as C
    from raft_stir_b200 import _lib
    r = int(golden["radius"])
    blk = R.CorrBlock(t(golden["fmap1"]), t(golden["fmap2"]), 4, r, storage="fp32", math="fp32")
    B, H, W = blk.B, blk.H, blk.W
    glay = _lib.pyramid_layout(B, H, W, 4, _lib.RC_VOL_F32)
    gbuf = torch.zeros(int(glay.total_elems), device=dev())
    coords = t(golden["coords_rand"])
    go = t(golden["grad_out"])
    for rep in range(2):   # accumulates: two calls give twice the gradient
        _lib.check(_lib.lib().rc_lookup_bwd(blk._buf.data_ptr(), blk.vol_dtype, coords.data_ptr(), go.data_ptr(),
                                            B, H, W, 4, r, gbuf.data_ptr(), None,
                                            torch.cuda.current_stream().cuda_stream), "rc_lookup_bwd")
    for l in range(4):
        v = _lib.level_view(gbuf, glay, l)
        got = v.cpu().numpy() / 2.0
        if l == 0:
            rel_close(got.sum(axis=2), golden["dpyr0_rowsum"], 1e-4, "dpyr0 row sums")
        else:
            rel_close(got, golden[f"dpyr{l}"], 1e-4, f"dpyr{l}")


def test_training_accumulates_over_iterations():
    """Several lookups share one gradient pyramid; the build backward runs once."""
    rng = np.random.default_rng(9)
    B, D, H, W, r = 1, 16, 16, 24, 3
    f1n = rng.standard_normal((B, D, H, W)).astype(np.float32)
    f2n = rng.standard_normal((B, D, H, W)).astype(np.float32)
    cs = [(O.coords_grid(B, H, W) + rng.normal(0, 2, (B, 2, H, W))).astype(np.float32) for _ in range(3)]
    gs = [rng.standard_normal((B, 4 * 49, H, W)).astype(np.float32) for _ in range(3)]
    f1 = t(f1n).requires_grad_(True)
    f2 = t(f2n).requires_grad_(True)
    blk = R.CorrBlock(f1, f2, 4, r, storage="fp32", math="fp32")
    loss = sum((blk(t(c)) * t(g)).sum() for c, g in zip(cs, gs))
    loss.backward()
    pyr = O.corr_pyramid(f1n.astype(np.float64), f2n.astype(np.float64), 4)
    acc = None
    for c, g in zip(cs, gs):
        dp, _ = O.window_lookup_backward(pyr, c, r, g)
        acc = dp if acc is None else [a + b for a, b in zip(acc, dp)]
    d1, d2 = O.pyramid_backward(acc, f1n, f2n)
    rel_close(f1.grad.cpu().numpy(), d1, 1e-3, "dfmap1 over 3 iterations")
    rel_close(f2.grad.cpu().numpy(), d2, 1e-3, "dfmap2 over 3 iterations")


# ------------------------------------------------------------------ on-the-fly path
@pytest.mark.parametrize("kind", KINDS)
def test_alternate_block_matches_reference(golden, kind):
    blk = R.AlternateCorrBlock(t(golden["fmap1"]), t(golden["fmap2"]), 4, int(golden["radius"]), math="fp32")
    out = blk(t(golden[f"coords_{kind}"]))
    ref = golden[f"out_{kind}"]
    assert tuple(out.shape) == ref.shape and out.dtype == torch.float32
    rel_close(out.cpu().numpy(), ref, 1e-4, f"alt {kind}")   # fp32 path: 1e-4


def test_alt_cuda_corr_forward_backward_contract():
    rng = np.random.default_rng(4)
    B, H1, W1, H2, W2, Cc, r = 2, 9, 13, 5, 7, 24, 3
    f1 = rng.standard_normal((B, H1, W1, Cc)).astype(np.float32)
    f2 = rng.standard_normal((B, H2, W2, Cc)).astype(np.float32)
    coords = rng.uniform(-3, 9, (B, 1, H1, W1, 2)).astype(np.float32)
    (corr,) = R.alt_cuda_corr.forward(t(f1), t(f2), t(coords), r)
    assert tuple(corr.shape) == (B, 1, 49, H1, W1)
    rel_close(corr.cpu().numpy(), O.alt_forward(f1, f2, coords, r), 1e-4, "alt_cuda_corr.forward")
    g = rng.standard_normal(tuple(corr.shape)).astype(np.float32)
    g1, g2, gc = R.alt_cuda_corr.backward(t(f1), t(f2), t(coords), t(g), r)
    o1, o2, oc = O.alt_backward(f1, f2, coords, g, r)
    rel_close(g1.cpu().numpy(), o1, 1e-4, "fmap1_grad")
    rel_close(g2.cpu().numpy(), o2, 1e-4, "fmap2_grad")
    assert tuple(gc.shape) == coords.shape and not gc.any().item()   # zeros, as in the reference


def test_alt_cuda_corr_many_coordinate_sets():
    """The reference's alt_cuda_corr.forward takes any number N of coordinate sets (coords [B,N,H,W,2],
    correlation.cpp:23-33); round 1 stopped at 6."""
    rng = np.random.default_rng(5)
    B, H1, W1, H2, W2, Cc, r, Nc = 1, 6, 9, 7, 8, 16, 2, 9
    f1 = rng.standard_normal((B, H1, W1, Cc)).astype(np.float32)
    f2 = rng.standard_normal((B, H2, W2, Cc)).astype(np.float32)
    coords = rng.uniform(-2, 9, (B, Nc, H1, W1, 2)).astype(np.float32)
    (corr,) = R.alt_cuda_corr.forward(t(f1), t(f2), t(coords), r)
    assert tuple(corr.shape) == (B, Nc, 25, H1, W1)
    rel_close(corr.cpu().numpy(), O.alt_forward(f1, f2, coords, r), 1e-4, "alt_cuda_corr.forward, 9 coordinate sets")
    g = rng.standard_normal(tuple(corr.shape)).astype(np.float32)
    g1, g2, _ = R.alt_cuda_corr.backward(t(f1), t(f2), t(coords), t(g), r)
    o1, o2, _ = O.alt_backward(f1, f2, coords, g, r)
    rel_close(g1.cpu().numpy(), o1, 1e-4, "fmap1_grad, 9 coordinate sets")
    rel_close(g2.cpu().numpy(), o2, 1e-4, "fmap2_grad, 9 coordinate sets")


def test_alternate_equals_default_at_sintel_shape():
    rng = np.random.default_rng(2)
    B, D, H, W = 1, 256, 55, 128
    f1 = t(rng.standard_normal((B, D, H, W)).astype(np.float32))
    f2 = t(rng.standard_normal((B, D, H, W)).astype(np.float32))
    c = R.coords_grid(B, H, W, dev()) + 5 * torch.randn(B, 2, H, W, device=dev())
    a = R.AlternateCorrBlock(f1, f2, 4, 4, math="fp32")(c)
    d = R.CorrBlock(f1, f2, 4, 4, storage="fp32", math="fp32")(c)
    rel_close(a.cpu().numpy(), d.cpu().numpy(), 1e-4, "alt vs default, fp32")
    h = R.CorrBlock(f1, f2, 4, 4)(c)
    rel_close(h.cpu().numpy(), d.cpu().numpy(), 1e-2, "bf16 tensor-core vs fp32, full size",
              floor=bf16_floor(f1.cpu().numpy(), f2.cpu().numpy()))


# ------------------------------------------------------------------ full-size properties
def test_full_size_properties_sintel_batch():
    """Size-independent checks at the BASELINE shape (55x128x256, batch 2)."""
    B, D, H, W, r = 2, 256, 55, 128, 4
    g = torch.Generator(device="cpu").manual_seed(0)
    f1 = torch.randn(B, D, H, W, generator=g).to(dev())
    f2 = torch.randn(B, D, H, W, generator=g).to(dev())
    blk = R.CorrBlock(f1, f2, 4, r)
    grid = R.coords_grid(B, H, W, dev())
    # integer coordinates: the centre tap of level 0 is the volume entry itself
    out = blk(grid)
    pyr0 = blk.corr_pyramid[0]
    n = 2 * r + 1
    centre = out[:, r * n + r]                                   # [B,H,W]
    q = torch.arange(B * H * W, device=dev())
    yy = (q % (H * W)) // W
    xx = q % W
    direct = pyr0[q, 0, yy, xx].float().view(B, H, W)
    assert torch.equal(centre, direct)
    # linearity in fmap1: doubling features doubles every sample (powers of two are exact in bf16)
    out2 = R.CorrBlock(2 * f1, f2, 4, r)(grid + 0.25)
    out1 = blk(grid + 0.25)
    assert torch.allclose(out2, 2 * out1, rtol=0, atol=0)
    # pooling: each level-1 cell is the mean of its four level-0 cells (up to bf16 rounding)
    p0 = blk.corr_pyramid[0][:, 0].float()
    p1 = blk.corr_pyramid[1][:, 0].float()
    m = 0.25 * (p0[:, 0:2 * (H // 2):2, 0:2 * (W // 2):2] + p0[:, 1:2 * (H // 2):2, 0:2 * (W // 2):2]
                + p0[:, 0:2 * (H // 2):2, 1:2 * (W // 2):2] + p0[:, 1:2 * (H // 2):2, 1:2 * (W // 2):2])
    assert (p1 - m).abs().max().item() < 0.02
    # batch independence: pair 1 alone gives the same bits
    solo = R.CorrBlock(f1[1:], f2[1:], 4, r)(grid[1:] + 0.25)
    assert torch.equal(solo, out1[1:])


# ------------------------------------------------------------------ interface / errors
def test_rejects_cpu_tensors_and_bad_shapes():
    with pytest.raises(RuntimeError):
        R.CorrBlock(torch.zeros(1, 8, 16, 16), torch.zeros(1, 8, 16, 16))
    f = torch.zeros(1, 8, 16, 16, device=dev())
    with pytest.raises(RuntimeError):
        R.CorrBlock(f, torch.zeros(1, 8, 16, 17, device=dev()))
    blk = R.CorrBlock(f, f)
    with pytest.raises(RuntimeError):
        blk(torch.zeros(1, 2, 16, 17, device=dev()))
    with pytest.raises(RuntimeError):
        R.CorrBlock(f, f, radius=7)(torch.zeros(1, 2, 16, 16, device=dev()))
    with pytest.raises(RuntimeError):
        R.alt_cuda_corr.forward(torch.zeros(1, 4, 4, 8), torch.zeros(1, 4, 4, 8), torch.zeros(1, 1, 4, 4, 2), 4)
    with pytest.raises(RuntimeError):   # non-contiguous, like CHECK_CONTIGUOUS
        R.alt_cuda_corr.forward(torch.zeros(1, 4, 8, 4, device=dev()).transpose(2, 3), torch.zeros(1, 4, 4, 8, device=dev()),
                                torch.zeros(1, 1, 4, 4, 2, device=dev()), 4)


def test_non_default_stream_and_noncontiguous_inputs():
    rng = np.random.default_rng(8)
    f1 = t(rng.standard_normal((1, 16, 32, 16)).astype(np.float32)).transpose(2, 3)   # non-contiguous [1,16,16,32]
    f2 = t(rng.standard_normal((1, 16, 16, 32)).astype(np.float32))
    c = R.coords_grid(1, 16, 32, dev()) + 0.5
    ref = R.CorrBlock(f1, f2, storage="fp32", math="fp32")(c)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        got = R.CorrBlock(f1, f2, storage="fp32", math="fp32")(c)
    s.synchronize()
    assert torch.equal(ref, got)


def test_lookup_is_cuda_graph_capturable():
    """The per-iteration call can be captured once and replayed (the natural way to run the 12-iteration
    loop without launch gaps): no synchronisation, no legacy-stream work inside rc_lookup_fwd."""
    rng = np.random.default_rng(3)
    B, D, H, W = 2, 32, 24, 40
    blk = R.CorrBlock(t(rng.standard_normal((B, D, H, W)).astype(np.float32)),
                      t(rng.standard_normal((B, D, H, W)).astype(np.float32)), 4, 4)
    coords = R.coords_grid(B, H, W, dev()) + 2.5
    ref = blk(coords).clone()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        blk(coords)                       # warm-up on the side stream (first-call attribute set-up)
    torch.cuda.current_stream().wait_stream(s)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        out = blk(coords)
    coords.add_(1.0)                      # new coordinates in the captured input buffer
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, blk(coords))
    assert not torch.equal(out, ref)


# ------------------------------------------------------------------ randomized shapes
def _random_cases(n, seed):
    rng = np.random.default_rng(seed)
    cases = []
    for _ in range(n):
        L = int(rng.integers(1, 5))
        lo = max(2 << (L - 1), 3)                      # every level at least 2 wide (the reference's limit)
        H, W = int(rng.integers(lo, 72)), int(rng.integers(lo, 90))
        cases.append((int(rng.integers(1, 4)), int(rng.integers(1, 9)) * 8, H, W, L, int(rng.integers(1, 5)),
                      ["bf16", "fp32"][int(rng.integers(0, 2))], int(rng.integers(0, 1 << 30))))
    return cases


@pytest.mark.parametrize("case", _random_cases(24, 2026), ids=lambda c: "B%d_D%d_%dx%d_L%d_r%d_%s" % c[:7])
def test_random_shapes_build_and_lookup_vs_oracle(case):
    """Seeded random (B, D, H, W, levels, radius, storage): tensor-core build + lookup against the oracle
    fed the same bf16-rounded features (so only storage rounding and fp32 summation order remain)."""
    B, D, H, W, L, r, storage, seed = case
    rng = np.random.default_rng(seed)
    f1 = rng.standard_normal((B, D, H, W)).astype(np.float32)
    f2 = rng.standard_normal((B, D, H, W)).astype(np.float32)
    coords = (O.coords_grid(B, H, W) + rng.normal(0, max(H, W) / 6.0, (B, 2, H, W))).astype(np.float32)
    coords[:, :, ::2, ::3] = np.round(coords[:, :, ::2, ::3])          # exact integers too
    blk = R.CorrBlock(t(f1), t(f2), L, r, storage=storage, math="bf16")
    pyr = O.corr_pyramid(bf16_round(f1).astype(np.float64), bf16_round(f2).astype(np.float64), L)
    tol = 2.0 ** -8 if storage == "bf16" else 2e-5
    for l, v in enumerate(blk.corr_pyramid):
        rel_close(v[:, 0].float().cpu().numpy(), pyr[l], tol, f"{case[:7]} level {l}")
    got_pyr = [v[:, 0].double().cpu().numpy() for v in blk.corr_pyramid]
    out = blk(t(coords)).cpu().numpy()
    # lookup on the stored values, oracle in fp64 (in fp32 the oracle's own x/2^l + offset rounds the
    # fractional part by up to 4e-6 at |x| ~ 50, which shows up as 1e-5 on steep taps)
    rel_close(out, O.window_lookup(got_pyr, coords.astype(np.float64), r), 2e-5, f"{case[:7]} lookup")


def test_video_shape_beyond_32bit_offsets():
    """BASELINE config 4's shape (1080x1920 -> 135x240 features), two pairs: the bf16 pyramid has 2.9 G
    elements (> 2^31), so every slab offset must be 64-bit.  Checker: this library's on-the-fly block
    (fp32, pinned to the reference extension in test_gpu_ref_extension.py), which never forms the volume."""
    B, D, H, W = 2, 64, 135, 240
    g = torch.Generator(device="cpu").manual_seed(5)
    f1 = torch.randn(B, D, H, W, generator=g).to(dev())
    f2 = torch.randn(B, D, H, W, generator=g).to(dev())
    c = R.coords_grid(B, H, W, dev()) + 8 * torch.randn(B, 2, H, W, generator=g).to(dev())
    blk = R.CorrBlock(f1, f2, 4, 4)
    assert blk.layout.total_elems > 2 ** 31
    got = blk(c)
    ref = R.AlternateCorrBlock(f1, f2, 4, 4, math="fp32")(c)
    rel_close(got.cpu().numpy(), ref.cpu().numpy(), 1e-2, "1080p, bf16 volume vs on-the-fly fp32",
              floor=bf16_floor(f1.cpu().numpy(), f2.cpu().numpy()))
    # the last query of the last pair sits at the far end of every level
    assert torch.isfinite(got[-1, :, -1, -1]).all() and got[-1, :, -1, -1].abs().sum() > 0


def test_alternate_tensor_core_fp16_operands():
    """AlternateCorrBlock(math="fp16"): the same kernels with IEEE-half operands.  Against the fp32 block fed
    fp16-rounded features level 0 differs by fp32 summation order only; against the unrounded fp32 block the whole
    output is within 1e-3 (bf16 operands: 1e-2)."""
    for shape in ((2, 128, 19, 45, 3), (1, 256, 55, 128, 4)):
        B, D, H, W, r = shape
        rng = np.random.default_rng(D + H)
        f1 = rng.standard_normal((B, D, H, W)).astype(np.float32)
        f2 = rng.standard_normal((B, D, H, W)).astype(np.float32)
        coords = (O.coords_grid(B, H, W) + rng.normal(0, 4, (B, 2, H, W))).astype(np.float32)
        got = R.AlternateCorrBlock(t(f1), t(f2), 4, r, math="fp16")(t(coords)).cpu().numpy()
        n = (2 * r + 1) ** 2
        ref16 = R.AlternateCorrBlock(t(fp16_round(f1)), t(fp16_round(f2)), 4, r, math="fp32")(t(coords)).cpu().numpy()
        rel_close(got[:, :n], ref16[:, :n], 2e-5, f"alt fp16 operands {shape} level 0")
        ref = R.AlternateCorrBlock(t(f1), t(f2), 4, r, math="fp32")(t(coords)).cpu().numpy()
        rel_close(got, ref, 1e-3, f"alt fp16 operands {shape} vs fp32", floor=bf16_floor(f1, f2))
    with pytest.raises(RuntimeError):
        R.AlternateCorrBlock(t(f1), t(f2), 4, 4, math="fp8")


# ------------------------------------------------------------------ on-the-fly path on the tensor cores
@pytest.mark.parametrize("kind", ["smooth", "rand", "int", "far", "edge"])
@pytest.mark.parametrize("shape", [(1, 64, 24, 40, 4), (2, 128, 19, 45, 3), (1, 256, 55, 128, 4),
                                   # persistent kernel: several units per CTA, one level per unit / all four levels
                                   (8, 64, 40, 64, 4), (37, 64, 16, 64, 4),
                                   # D > 256: the per-(block, level) kernel without resident query features
                                   (1, 320, 17, 33, 2)])
def test_alternate_tensor_core_matches_fp32_path(shape, kind):
    """rc_alt_lookup_fwd_tc against rc_alt_lookup_fwd (fp32, itself pinned to the reference extension) fed
    bf16-rounded features: what remains is fp32 summation order.  Coordinates: a smooth flow (small
    boxes), random (boxes of many chunks), exact integers, far outside (all zeros), straddling the border."""
    B, D, H, W, r = shape
    rng = np.random.default_rng(H * W + D + len(kind))
    f1 = bf16_round(rng.standard_normal((B, D, H, W)).astype(np.float32))
    f2 = bf16_round(rng.standard_normal((B, D, H, W)).astype(np.float32))
    grid = O.coords_grid(B, H, W)
    if kind == "smooth":
        yy, xx = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
        flow = np.stack([3.3 + 0.05 * xx + 0.02 * yy, -2.7 + 0.03 * yy - 0.01 * xx])[None]
        coords = grid + flow
    elif kind == "rand":
        coords = grid + rng.normal(0, 6, (B, 2, H, W))
    elif kind == "int":
        coords = grid + np.round(rng.normal(0, 3, (B, 2, 1, 1)))
    elif kind == "far":
        coords = grid + 1.0e4
    else:
        coords = grid + np.array([W - 6.5, -(H - 5.25)]).reshape(1, 2, 1, 1) * (grid[:, :1] > W / 2)
    coords = coords.astype(np.float32)
    ref = R.AlternateCorrBlock(t(f1), t(f2), 4, r, math="fp32")(t(coords))
    got = R.AlternateCorrBlock(t(f1), t(f2), 4, r, math="bf16")(t(coords))
    assert got.shape == ref.shape
    n = (2 * r + 1) ** 2
    # level 0 sees the same bf16 features in both paths: only fp32 summation order differs
    rel_close(got[:, :n].cpu().numpy(), ref[:, :n].cpu().numpy(), 2e-5, f"alt tensor-core {shape} {kind} level 0")
    # levels 1..3: the pooled features are rounded to bf16 only on the tensor-core path: 1e-2 (bf16)
    rel_close(got.cpu().numpy(), ref.cpu().numpy(), 1e-2, f"alt tensor-core {shape} {kind}", floor=bf16_floor(f1, f2))
    if kind == "far":
        assert not got.any().item()
