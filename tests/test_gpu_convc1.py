"""the lookup fused with the motion encoder's first layer, cor = relu(convc1(corr))
(core/update.py:70-71, :88-90), through rc_lookup_convc1_fwd.

Checker: torch's fp32 conv2d + relu applied to this library's own fp32 lookup (itself pinned to the reference's
CorrBlock by tests/test_gpu_parity.py and tests/test_gpu_golden_shapes.py), and, where oracle/_ref/*.pt exists, the
reference's traced convc1 on the golden lookup.  Tolerance: the fused path rounds samples and weights to IEEE half
(2^-11 relative each, fp32 accumulate): |got - ref| <= 2e-3 * (|ref| + rms(pre-activation)); the north star's bf16
bound would be 1e-2."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

R = pytest.importorskip("raft_stir_b200")

CASES = [
    # B, D, H, W, levels, radius, O
    (2, 256, 55, 128, 4, 4, 256),     # RAFT-basic, Sintel features
    (1, 256, 46, 62, 4, 4, 256),      # FlyingChairs features: N % 8 != 0 (padded row pitch of the intermediate)
    (2, 128, 20, 28, 4, 3, 96),       # RAFT-small: C = 196, O = 96 (core/update.py:62)
    (1, 64, 17, 22, 4, 4, 256),       # the golden fixture's odd shape
    (1, 64, 24, 40, 2, 2, 32),        # C = 50: no full 64-channel block, tail steps only
]


def dev():
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but torch.cuda.is_available() is False")
    return torch.device("cuda:0")


def check(got, pre, what, tol=2e-3):
    ref = torch.relu(pre).double()
    rms = pre.double().pow(2).mean().sqrt().item()
    err = (got.double() - ref).abs()
    bad = err > tol * (ref.abs() + rms)
    assert not bool(bad.any()), f"{what}: {int(bad.sum())} / {bad.numel()} outside {tol:g}; max err {err.max().item():.3e}, rms {rms:.3e}"


@pytest.mark.parametrize("storage", ["bf16", "fp32"])
@pytest.mark.parametrize("case", CASES, ids=lambda c: "x".join(map(str, c)))
def test_lookup_convc1_matches_conv_of_lookup(case, storage):
    B, D, H, W, L, r, O = case
    g = torch.Generator().manual_seed(D + H * W + O)
    f1 = torch.randn(B, D, H, W, generator=g).to(dev())
    f2 = torch.randn(B, D, H, W, generator=g).to(dev())
    C = L * (2 * r + 1) ** 2
    w = (torch.randn(O, C, 1, 1, generator=g) / np.sqrt(C)).to(dev())
    b = (0.1 * torch.randn(O, generator=g)).to(dev())
    coords = R.coords_grid(B, H, W, dev()) + 3 * torch.randn(B, 2, H, W, generator=g).to(dev())
    blk = R.CorrBlock(f1, f2, L, r, storage=storage, math="bf16" if storage == "bf16" else "fp32")
    pre = F.conv2d(blk(coords), w, b)                      # the same stored volume: isolates the fused kernels
    got = blk.lookup_convc1(coords, w, b)
    assert got.shape == (B, O, H, W) and got.dtype == torch.float32
    check(got, pre, f"{case} {storage}")
    got16 = blk.lookup_convc1(coords, w, b, out_dtype=torch.float16)
    assert got16.dtype == torch.float16
    check(got16.float(), pre, f"{case} {storage} fp16 out", tol=3e-3)


def test_lazy_corr_through_a_motion_encoder():
    """The drop-in route: a block built with fuse_convc1=True answers with a LazyCorr, and an encoder whose convc1 was
    wrapped by fuse_motion_encoder resolves it; tensors still go through the plain convolution."""
    class Encoder(torch.nn.Module):                        # the first two lines of BasicMotionEncoder (core/update.py:79-90)
        def __init__(self):
            super().__init__()
            self.convc1 = torch.nn.Conv2d(324, 256, 1, padding=0)

        def forward(self, corr):
            return F.relu(self.convc1(corr))

    torch.manual_seed(5)
    enc = Encoder().to(dev()).eval()
    assert R.fuse_motion_encoder(enc) == 1
    B, D, H, W = 1, 256, 46, 62
    f1, f2 = torch.randn(B, D, H, W, device=dev()), torch.randn(B, D, H, W, device=dev())
    coords = R.coords_grid(B, H, W, dev()) + 2 * torch.randn(B, 2, H, W, device=dev())
    with torch.no_grad():
        plain = R.CorrBlock(f1, f2, radius=4)
        lazy = R.CorrBlock(f1, f2, radius=4, fuse_convc1=True)(coords)
        assert isinstance(lazy, R.LazyCorr) and tuple(lazy.shape) == (B, 324, H, W)
        ref = enc(plain(coords))                           # tensor in -> the wrapped fp32 convolution
        got = enc(lazy)
        assert torch.equal(lazy.materialize(), plain(coords))
    check(got, enc.convc1.conv(plain(coords)), "lazy route")
    assert (got - ref).abs().max().item() < 0.05
    # with gradients wanted the block behaves as always
    c = coords.clone().requires_grad_(True)
    assert isinstance(R.CorrBlock(f1, f2, radius=4, fuse_convc1=True)(c), torch.Tensor)


def test_flow_with_fused_lookup_convc1():
    """End to end (BASELINE config 1): the reference's networks with lookup + convc1 fused; <= 0.01 px of the reference."""
    from oracle import raft_harness as Hn
    if not Hn.available():
        pytest.skip("oracle/_ref/*.pt not built (needs the reference checkout)")
    nets = Hn.load(dev())
    g = torch.Generator(device="cpu").manual_seed(0)
    im1 = (torch.rand(1, 3, 440, 1024, generator=g) * 255.0).to(dev())
    im2 = (torch.rand(1, 3, 440, 1024, generator=g) * 255.0).to(dev())
    _, up = Hn.raft_forward(nets, im1, im2, iters=12, corr_cls=Hn.TorchCorrBlock)
    # the restated wiring alone (reference block, convc1 applied outside) must reproduce the traced block
    class Outside(Hn.TorchCorrBlock):
        def lookup_convc1(self, coords, w, b):
            return F.relu(F.conv2d(self(coords), w, b))
    _, same = Hn.raft_forward(nets, im1, im2, iters=12, corr_cls=Outside, fuse_convc1=True)
    # (not bit-identical: an eager cuDNN convolution outside the traced block; the recurrence amplifies its last bits --
    # tests/test_gpu_raft_e2e.py measures the same yardstick with 1e-7 relative noise on the block's output)
    assert torch.sqrt(((same - up) ** 2).sum(1)).mean().item() <= 2e-3
    _, got = Hn.raft_forward(nets, im1, im2, iters=12, corr_cls=R.CorrBlock, fuse_convc1=True)
    d = torch.sqrt(((got - up) ** 2).sum(1)).mean().item()
    assert d <= 0.01, f"mean EPE delta {d:.5f} px > 0.01 px"
