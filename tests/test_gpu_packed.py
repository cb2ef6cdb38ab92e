"""16-bit channels-last features consumed in place (rc_pack_features_nhwc16 /
rc_build_pyramid_nhwc16).  The NHWC build feeds the SAME 16-bit operands to the same MMAs as the default path (which
repacks fp32 NCHW itself), so with bf16 operands every level must be bit-identical: scaling by the exact power-of-two
part of 1/sqrt(D) before (default path) or after (this path) the fp32 accumulation commutes with every rounding.
With fp16 operands the default path's pre-scaled features can land in half's subnormal range (|f|/16 < 6e-5) and lose
bits there, so single stored values may differ by one rounding step of the storage type."""
import pytest
import torch

pytestmark = pytest.mark.gpu

R = pytest.importorskip("raft_stir_b200")

SHAPES = [(2, 256, 55, 128), (3, 256, 46, 62), (1, 128, 27, 45), (2, 64, 17, 22), (1, 192, 9, 70)]


def dev():
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but torch.cuda.is_available() is False")
    return torch.device("cuda:0")


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "x".join(map(str, s)))
def test_pack_and_nhwc_build_equal_the_default_path(shape, dtype):
    B, D, H, W = shape
    g = torch.Generator().manual_seed(B * D + H)
    f1 = torch.randn(shape, generator=g).to(dev())
    f2 = torch.randn(shape, generator=g).to(dev())
    p1, p2 = R.pack_features(f1, f2, dtype)
    assert p1.shape == f1.shape and p1.is_contiguous(memory_format=torch.channels_last)
    # the pack is a transposing round-to-nearest cast (fp16: saturating, no value here is near 65 504)
    assert torch.equal(p1, f1.to(dtype).contiguous(memory_format=torch.channels_last))
    assert torch.equal(p2, f2.to(dtype).contiguous(memory_format=torch.channels_last))
    L = 4 if min(H, W) >= 16 else 3
    name = "bf16" if dtype is torch.bfloat16 else "fp16"
    for storage in ("bf16", "fp32"):
        ref = R.CorrBlock(f1, f2, L, 4, storage=storage, math=name)
        got = R.CorrBlock(p1, p2, L, 4, storage=storage)
        assert got._nhwc16 and got.math == ref.math
        for l, (a, b) in enumerate(zip(got.corr_pyramid, ref.corr_pyramid)):     # levels, not raw buffers: the
            if dtype is torch.bfloat16:                                          # gaps between levels are never written
                assert torch.equal(a, b), f"{shape} {name} {storage} level {l}: NHWC build differs from the default path"
            else:
                step = 2.0 ** -7 if storage == "bf16" else 1e-5      # one unit in the last place of the storage type
                d = (a.float() - b.float()).abs()
                worst = float((d / (b.float().abs() + 1.0)).max())
                same = float((d == 0).float().mean())
                assert worst <= step and same > (0.99 if storage == "bf16" else 0.5), \
                    f"{shape} fp16 {storage} level {l}: worst {worst:.3e} (<= {step:g}), identical {same:.4f}"
    c = R.coords_grid(B, H, W, dev()) + 3 * torch.randn(B, 2, H, W, device=dev())
    if dtype is torch.bfloat16:
        assert torch.equal(got(c), ref(c))


def test_encoder_style_input_is_consumed_in_place():
    """What RAFT.forward would hand over with a channels_last bf16 encoder: one [2B, D, H, W] tensor split along the
    batch (core/extractor.py:189-190); both halves are views and must be used without a copy."""
    B, D, H, W = 2, 256, 46, 62
    g = torch.Generator().manual_seed(3)
    both = torch.randn(2 * B, D, H, W, generator=g).to(dev()).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    f1, f2 = torch.split(both, [B, B], dim=0)
    blk = R.CorrBlock(f1, f2, radius=4)
    assert blk._nhwc16
    ref = R.CorrBlock(f1.float(), f2.float(), radius=4)        # bf16 values are exact in fp32: same operands after repack
    for a, b in zip(blk.corr_pyramid, ref.corr_pyramid):
        assert torch.equal(a, b)
    # a tensor that wants gradients takes the ordinary (fp32, autograd) route
    fg = f1.float().requires_grad_(True)
    assert not R.CorrBlock(fg, f2.float(), radius=4)._nhwc16
