"""This library's rc_alt_fwd / rc_alt_bwd against the REFERENCE'S OWN alt_cuda_corr extension,
compiled unmodified from /root/reference/alt_cuda_corr by oracle/build_ref_ext.sh and shipped in
oracle/_ref/.  Same tensors in (alt_cuda_corr/correlation.cpp:23-48), same lists out.  fp32: 1e-4.
The reference kernel reads channels in blocks of 32 without a bound (correlation_kernel.cu:43-55), so
only C % 32 == 0 is defined behaviour for it.  Skipped where the extension was not built."""
import numpy as np
import pytest
import torch

from oracle import ref_ext

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_ext.available(), reason="oracle/_ref/alt_cuda_corr_ref*.so not built")]

R = pytest.importorskip("raft_stir_b200")


def _close(got, ref, rel, what):
    got, ref = got.double().cpu().numpy(), ref.double().cpu().numpy()
    rms = float(np.sqrt((ref ** 2).mean())) + 1e-30
    err = np.abs(got - ref)
    assert (err <= rel * (np.abs(ref) + rms)).all(), f"{what}: max err {err.max():.3e}, rms {rms:.3e}"


@pytest.mark.parametrize("shape", [(2, 9, 13, 5, 7, 32, 3), (1, 20, 28, 20, 28, 64, 4), (2, 16, 24, 8, 12, 256, 4)])
def test_forward_and_backward_match_the_reference_extension(shape):
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but no CUDA device is visible")
    dev = torch.device("cuda:0")
    ext = ref_ext.load()
    B, H1, W1, H2, W2, C, r = shape
    g = torch.Generator().manual_seed(H1 * W1 + C)
    f1 = torch.randn(B, H1, W1, C, generator=g).to(dev)
    f2 = torch.randn(B, H2, W2, C, generator=g).to(dev)
    # window centres spread over and beyond the map, plus exact integers (the first GRU iteration)
    coords = torch.empty(B, 1, H1, W1, 2)
    coords[..., 0] = torch.rand(B, 1, H1, W1, generator=g) * (W2 + 6) - 3
    coords[..., 1] = torch.rand(B, 1, H1, W1, generator=g) * (H2 + 6) - 3
    coords[:, :, ::3] = coords[:, :, ::3].round()
    coords = coords.to(dev)
    with torch.cuda.device(dev):
        torch.cuda.synchronize()
        (ref,) = ext.forward(f1, f2, coords, r)          # launches on the legacy default stream
        torch.cuda.synchronize()
    (got,) = R.alt_cuda_corr.forward(f1, f2, coords, r)
    assert got.shape == ref.shape
    _close(got, ref, 1e-4, f"forward {shape}")
    grad = torch.randn(ref.shape, generator=g).to(dev)
    torch.cuda.synchronize()
    r1, r2, rc = ext.backward(f1, f2, coords, grad, r)
    torch.cuda.synchronize()
    g1, g2, gc = R.alt_cuda_corr.backward(f1, f2, coords, grad, r)
    _close(g1, r1, 1e-4, f"fmap1_grad {shape}")
    _close(g2, r2, 1e-4, f"fmap2_grad {shape}")
    assert gc.shape == rc.shape and not gc.any().item() and not rc.any().item()   # the reference never writes it
