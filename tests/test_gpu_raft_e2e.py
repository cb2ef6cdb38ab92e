"""End-to-end check of BASELINE config 1: RAFT-basic, random-init weights (seed 1234), one
440x1024 pair, 12 iterations -- the reference's sub-networks (frozen by
oracle/make_ref_modules.py) around (a) the reference CorrBlock's torch arithmetic and (b) this
library's correlation blocks.  north_star tolerance: mean end-point-error delta of the final
flow <= 0.01 px against the fp32 reference.
"""
from functools import partial

import pytest
import torch

from oracle import raft_harness as Hn

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not Hn.available(), reason="oracle/_ref/*.pt not built (needs the reference checkout)")]

R = pytest.importorskip("raft_stir_b200")


def _inputs(B, H, W, dev):
    g = torch.Generator(device="cpu").manual_seed(0)
    im1 = torch.rand(B, 3, H, W, generator=g) * 255.0
    im2 = torch.rand(B, 3, H, W, generator=g) * 255.0
    return im1.to(dev), im2.to(dev)


def _epe(a, b):
    return torch.sqrt(((a - b) ** 2).sum(dim=1)).mean().item()


@pytest.fixture(scope="module")
def reference_run():
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but no CUDA device is visible")
    dev = torch.device("cuda:0")
    nets = Hn.load(dev)
    im1, im2 = _inputs(1, 440, 1024, dev)
    low, up = Hn.raft_forward(nets, im1, im2, iters=12, corr_cls=Hn.TorchCorrBlock)
    assert torch.isfinite(up).all()
    return nets, im1, im2, low, up


def test_flow_with_bf16_tensor_core_volume(reference_run):
    nets, im1, im2, low, up = reference_run
    _, got = Hn.raft_forward(nets, im1, im2, iters=12, corr_cls=R.CorrBlock)      # defaults: bf16 storage, tcgen05 build
    d = _epe(got, up)
    assert d <= 0.01, f"mean EPE delta {d:.5f} px > 0.01 px (mean |flow| {up.norm(dim=1).mean().item():.2f} px)"


def test_flow_with_library_upsampling(reference_run):
    """The fused convex upsampling (rc_upsample_flow) in place of the reference's softmax/unfold/sum: with the
    reference's own correlation it must reproduce the flow to fp32 rounding, with this library's to the same 0.01 px."""
    nets, im1, im2, low, up = reference_run
    diffs = []

    def both(flow, mask):          # the two upsamplings on the very same flow and mask of a real run
        ref = Hn.upsample_flow(flow, mask)
        diffs.append(_epe(R.upsample_flow(flow, mask), ref))
        return ref

    Hn.raft_forward(nets, im1, im2, iters=12, corr_cls=Hn.TorchCorrBlock, last_only=False, upsample_fn=both)
    assert len(diffs) == 12 and max(diffs) <= 1e-4, f"upsampling differs from the reference's by {max(diffs):.2e} px"
    _, got = Hn.raft_forward(nets, im1, im2, iters=12, corr_cls=R.CorrBlock, upsample_fn=R.upsample_flow)
    d = _epe(got, up)
    assert d <= 0.01, f"mean EPE delta {d:.5f} px > 0.01 px"


def test_flow_with_tf32_build_fp32_volume(reference_run):
    nets, im1, im2, low, up = reference_run
    _, got = Hn.raft_forward(nets, im1, im2, iters=12, corr_cls=partial(R.CorrBlock, storage="fp32", math="tf32"))
    d = _epe(got, up)
    assert d <= 0.002, f"mean EPE delta {d:.5f} px with the tf32 build (bf16 build: ~0.0023, fp32: ~0.0003)"


def test_flow_with_fp32_volume(reference_run):
    nets, im1, im2, low, up = reference_run
    _, got = Hn.raft_forward(nets, im1, im2, iters=12, corr_cls=partial(R.CorrBlock, storage="fp32", math="fp32"))
    assert _epe(got, up) <= 1e-3


def test_flow_with_on_the_fly_correlation(reference_run):
    nets, im1, im2, low, up = reference_run
    _, got = Hn.raft_forward(nets, im1, im2, iters=12, corr_cls=partial(R.AlternateCorrBlock, math="fp32"))
    assert _epe(got, up) <= 1e-3


def test_flow_with_on_the_fly_correlation_default_settings(reference_run):
    """AlternateCorrBlock as a user swapping the two names gets it: D = 256 -> the tensor-core path (bf16 operands)."""
    nets, im1, im2, low, up = reference_run
    blocks = []

    def make(f1, f2, **kw):
        blocks.append(R.AlternateCorrBlock(f1, f2, **kw))
        return blocks[-1]
    _, got = Hn.raft_forward(nets, im1, im2, iters=12, corr_cls=make)
    assert blocks and blocks[0].math == "bf16"
    d = _epe(got, up)
    assert d <= 0.01, f"mean EPE delta {d:.5f} px > 0.01 px"


def test_flow_with_on_the_fly_correlation_on_tensor_cores(reference_run):
    nets, im1, im2, low, up = reference_run
    _, got = Hn.raft_forward(nets, im1, im2, iters=12, corr_cls=partial(R.AlternateCorrBlock, math="bf16"))
    d = _epe(got, up)
    assert d <= 0.01, f"mean EPE delta {d:.5f} px > 0.01 px"


class _JitteredReference(Hn.TorchCorrBlock):
    """The reference block with 1e-7 relative noise on its output: the yardstick for how much the rest
    of the network amplifies a last-bit difference in the correlation features."""

    def __call__(self, coords):
        out = super().__call__(coords)
        return out * (1 + 1e-7 * torch.randn_like(out))


def test_training_gradients_through_the_block():
    """BASELINE config 5 in miniature: loss.backward() through the whole network with this library's
    autograd-enabled CorrBlock (fp32).  The block itself matches torch autograd of the reference ops to
    1e-6 (test_gpu_parity.py); the random-init network around it turns a 1e-7 difference in the
    correlation features into a ~1e-2 difference in the parameter gradients (measured here with the
    jittered reference), so that is the scale the comparison is made on."""
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but no CUDA device is visible")
    dev = torch.device("cuda:0")
    im1, im2 = _inputs(2, 160, 224, dev)
    gt = torch.zeros(2, 2, 160, 224, device=dev)
    valid = torch.ones(2, 160, 224, device=dev)
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(0)
    try:
        grads, losses = [], []
        for cls in (Hn.TorchCorrBlock, _JitteredReference, partial(R.CorrBlock, storage="fp32", math="fp32")):
            nets = Hn.load(dev)
            loss = Hn.sequence_loss(Hn.raft_forward_train(nets, im1, im2, 4, cls), gt, valid)
            loss.backward()
            losses.append(loss.item())
            grads.append(torch.cat([p.grad.flatten() for p in nets[0].parameters() if p.grad is not None]))
    finally:
        torch.backends.cudnn.allow_tf32 = old
    a, aj, b = grads
    assert a.numel() > 1000 and torch.isfinite(b).all() and b.abs().max() > 0
    assert abs(losses[2] - losses[0]) <= 1e-4 * abs(losses[0])
    floor = ((a - aj).norm() / a.norm()).item()
    rel = ((a - b).norm() / a.norm()).item()
    cos = torch.nn.functional.cosine_similarity(a, b, dim=0).item()
    assert cos > 0.999 and rel < 3 * floor + 2e-3, f"gradient difference {rel:.2e} (cos {cos:.5f}); 1e-7 jitter floor {floor:.2e}"


def test_raft_small_with_library_block_and_upflow8():
    """RAFT-small end to end (core/raft.py:29-33, :48-51): 128-channel features, radius-3 lookup, no mask so the flow
    leaves through upflow8.  Traced small sub-networks around (a) the reference block's torch ops + F.interpolate and
    (b) this library's CorrBlock (bf16 tensor-core build, post-scaled since 128 is not a power of 4) + rc_upflow8."""
    if not Hn.available(small=True):
        pytest.skip("oracle/_ref/raft_small_*.pt not built")
    dev = torch.device("cuda:0")
    nets = Hn.load(dev, small=True)
    im1, im2 = _inputs(1, 440, 1024, dev)
    _, ref = Hn.raft_small_forward(nets, im1, im2, iters=12)
    assert torch.isfinite(ref).all()
    _, got = Hn.raft_small_forward(nets, im1, im2, iters=12, corr_cls=R.CorrBlock, upflow_fn=R.upflow8)
    d = _epe(got, ref)
    assert d <= 0.01, f"RAFT-small: mean EPE delta {d:.5f} px > 0.01 px (mean |flow| {ref.norm(dim=1).mean().item():.2f} px)"
    _, got = Hn.raft_small_forward(nets, im1, im2, iters=12, corr_cls=partial(R.CorrBlock, storage="fp32", math="tf32"),
                                   upflow_fn=R.upflow8)
    d32 = _epe(got, ref)
    assert d32 <= 0.002, f"RAFT-small, tf32 build: mean EPE delta {d32:.5f} px"


def test_mixed_precision_with_fp16_lookup_and_fp16_mask(reference_run):
    """--mixed_precision (core/raft.py:98-100, :110-114, :127-128: the three networks under fp16 autocast).  The block
    then returns fp16 (out_dtype); against the reference block
    under the same autocast the flow moves by less than autocast itself moves it from the fp32 run, and stays within
    0.01 px of the fp32 reference."""
    nets, im1, im2, low, up = reference_run
    _, ref16 = Hn.raft_forward(nets, im1, im2, iters=12, corr_cls=Hn.TorchCorrBlock, mixed_precision=True)
    autocast_noise = _epe(ref16, up)
    assert 1e-5 < autocast_noise < 0.01, f"autocast run differs from fp32 by {autocast_noise:.5f} px (is autocast active?)"
    # (the traced update block hands the mask over in whatever type its last traced op produces; the half-precision
    # mask path of the upsampling has its own test in test_gpu_upsample.py)
    _, got = Hn.raft_forward(nets, im1, im2, iters=12, corr_cls=partial(R.CorrBlock, out_dtype=torch.float16),
                             upsample_fn=R.upsample_flow, mixed_precision=True)
    assert _epe(got, ref16) <= 0.01 and _epe(got, up) <= 0.01, (_epe(got, ref16), _epe(got, up))



def test_whole_forward_as_one_cuda_graph(reference_run):
    """the reference's 12-iteration loop (core/raft.py:122-139) with this library's block and
    upsampling, captured once by raft_stir_b200.GraphedCallable and replayed on NEW images: same flow as the eager run
    on those images (bit-identical kernels, same order) and within the 0.01 px budget of the reference."""
    nets, im1, im2, low, up = reference_run

    def fwd(a, b):
        return Hn.raft_forward(nets, a, b, iters=12, corr_cls=R.CorrBlock, upsample_fn=R.upsample_flow)

    g = torch.Generator(device="cpu").manual_seed(77)
    other1 = (torch.rand(im1.shape, generator=g) * 255).to(im1.device)
    other2 = (torch.rand(im2.shape, generator=g) * 255).to(im1.device)
    graphed = R.GraphedCallable(fwd, (other1, other2))          # captured on other images ...
    _, got = graphed(im1, im2)                                  # ... replayed on the fixture's
    _, eager = fwd(im1, im2)
    assert (got - eager).abs().max().item() <= 1e-3      # same kernels in the same order (cuDNN may pick another algorithm)
    assert _epe(got, up) <= 0.01
    _, again = graphed(other1, other2)
    _, eager2 = fwd(other1, other2)
    assert (again - eager2).abs().max().item() <= 1e-3
    with pytest.raises(RuntimeError):
        graphed(im1[:, :, :-8], im2[:, :, :-8])
