"""rc_build_pyramid_static: the volume build for hosts that must not allocate.

A tracker that froze its memory plan at start-up, or a plugin living inside
another framework's arena, can hand this library its inputs and an output
pyramid but has no spare device buffer for the repacked bf16 features, and may
not call cudaMalloc to get one.  The entry point under test keeps that scratch
in a fixed 16 MiB array inside the library instead.  It has to produce exactly
what rc_build_pyramid produces with a caller-owned workspace -- on the first
call, and again on a second call with other inputs, because the array is shared
and still holds the previous call's features.

Collected last on purpose (file name): a failure here cannot hide any other test
from a run that stops at the first failure.  Runs alone or after the others.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

_lib = pytest.importorskip("raft_stir_b200._lib")


def _build_both(f1, f2, L, vol_dtype):
    """(pyramid via caller workspace, pyramid via the built-in scratch), same buffers layout."""
    lib = _lib.lib()
    B, D, H, W = f1.shape
    lay = _lib.pyramid_layout(B, H, W, L, vol_dtype)
    dt = torch.bfloat16 if vol_dtype == _lib.RC_VOL_BF16 else torch.float32
    stream = torch.cuda.current_stream().cuda_stream
    # fill with a sentinel so that "nothing was written" cannot pass
    ref = torch.full((int(lay.total_elems),), 7.0, dtype=dt, device=f1.device)
    got = torch.full((int(lay.total_elems),), -7.0, dtype=dt, device=f1.device)
    ws_bytes = int(lib.rc_build_workspace_bytes(B, D, H, W, _lib.RC_MATH_BF16))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=f1.device)
    _lib.check(lib.rc_build_pyramid(f1.data_ptr(), f2.data_ptr(), B, D, H, W, L, vol_dtype, _lib.RC_MATH_BF16,
                                    ref.data_ptr(), ws.data_ptr(), ws_bytes, stream), "rc_build_pyramid")
    _lib.check(lib.rc_build_pyramid_static(f1.data_ptr(), f2.data_ptr(), B, D, H, W, L, vol_dtype,
                                           got.data_ptr(), stream), "rc_build_pyramid_static")
    torch.cuda.synchronize()
    return [[_lib.level_view(buf, lay, l).float() for l in range(L)] for buf in (ref, got)]


def test_static_scratch_build_matches_workspace_build_twice():
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but no CUDA device is visible")
    dev = torch.device("cuda:0")
    g = torch.Generator(device="cpu").manual_seed(123)
    # two calls in a row, different sizes and contents, sharing the library's array;
    # the second is smaller, so stale features of the first sit right behind its data
    cases = [((1, 256, 55, 128), _lib.RC_VOL_BF16), ((2, 64, 19, 45), _lib.RC_VOL_F32)]
    for shape, vol_dtype in cases:
        f1 = torch.randn(*shape, generator=g).to(dev)
        f2 = torch.randn(*shape, generator=g).to(dev)
        ref, got = _build_both(f1, f2, 4, vol_dtype)
        for l, (a, b) in enumerate(zip(ref, got)):
            assert torch.isfinite(b).all(), f"level {l}: static-scratch build left non-finite values"
            # same kernels, same operands: the tests' fp32 tolerance (1e-4) is generous here
            scale = a.abs().max().item()
            err = (a - b).abs().max().item()
            assert err <= 1e-4 * scale, f"shape {shape} level {l}: max |diff| {err:.3e} vs scale {scale:.3e}"


def test_static_scratch_rejects_oversize_request():
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but no CUDA device is visible")
    lib = _lib.lib()
    B, D, H, W = 4, 256, 55, 128          # 4 Sintel pairs need ~29 MB of scratch, the array holds 16 MiB
    assert lib.rc_build_workspace_bytes(B, D, H, W, _lib.RC_MATH_BF16) > _lib.RC_STATIC_SCRATCH_BYTES
    dev = torch.device("cuda:0")
    f = torch.zeros(B, D, H, W, device=dev)
    lay = _lib.pyramid_layout(B, H, W, 4, _lib.RC_VOL_BF16)
    out = torch.zeros(int(lay.total_elems), dtype=torch.bfloat16, device=dev)
    rc = lib.rc_build_pyramid_static(f.data_ptr(), f.data_ptr(), B, D, H, W, 4, _lib.RC_VOL_BF16,
                                     out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == -3                                      # RC_ESIZE
    msg = lib.rc_last_error().decode()
    assert "scratch" in msg and "rc_build_pyramid" in msg
    torch.cuda.synchronize()
    assert not out.any().item()                          # nothing was launched
