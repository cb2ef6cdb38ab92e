"""Drop-in `upsample_flow` / `upflow8` (reference: core/raft.py:72-83, core/utils/utils.py:80-82).

Same arguments and results as the reference's two functions, autograd included (training backpropagates
through every iteration's upsampled flow, train.py:47-72), computed by one kernel each in libraftcorr.so:

    flow_up = raft_stir_b200.upsample_flow(coords1 - coords0, up_mask)     # [B,2,H,W], [B,576,H,W] -> [B,2,8H,8W]
    flow_up = raft_stir_b200.upflow8(coords1 - coords0)                    # RAFT-small: no mask

`install_into(raft_module)` in corr.py rebinds the corr blocks; these two are rebound with

    raft.RAFT.upsample_flow = lambda self, flow, mask: raft_stir_b200.upsample_flow(flow, mask)
    raft.upflow8 = raft_stir_b200.upflow8

CPU tensors are rejected: there is no fallback.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import RaftCorrError, check
from .corr import _guard, _require_cuda_f32, _stream


def _upsample_fwd(flow, mask):
    B, _, H, W = flow.shape
    out = torch.empty(B, 2, 8 * H, 8 * W, device=flow.device, dtype=torch.float32)
    fn = _lib.lib().rc_upsample_flow_f16mask if mask.dtype is torch.float16 else _lib.lib().rc_upsample_flow
    with _guard(flow.device):
        check(fn(flow.data_ptr(), mask.data_ptr(), B, H, W, out.data_ptr(), _stream(flow)), "rc_upsample_flow")
    return out


class _UpsampleFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, flow, mask):
        ctx.save_for_backward(flow, mask)
        return _upsample_fwd(flow, mask)

    @staticmethod
    def backward(ctx, gout):
        flow, mask = ctx.saved_tensors
        B, _, H, W = flow.shape
        gout = gout.contiguous().float()
        dflow = torch.empty_like(flow)
        dmask = torch.empty_like(mask)
        with _guard(flow.device):
            check(_lib.lib().rc_upsample_flow_bwd(flow.data_ptr(), mask.data_ptr(), gout.data_ptr(), B, H, W,
                                                  dflow.data_ptr(), dmask.data_ptr(), _stream(flow)),
                  "rc_upsample_flow_bwd")
        return dflow, dmask


class _Upflow8Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, flow):
        B, Cn, H, W = flow.shape
        out = torch.empty(B, Cn, 8 * H, 8 * W, device=flow.device, dtype=torch.float32)
        with _guard(flow.device):
            check(_lib.lib().rc_upflow8(flow.data_ptr(), B, Cn, H, W, out.data_ptr(), _stream(flow)), "rc_upflow8")
        ctx.shape = (B, Cn, H, W)
        return out

    @staticmethod
    def backward(ctx, gout):
        B, Cn, H, W = ctx.shape
        gout = gout.contiguous().float()
        dflow = torch.empty(B, Cn, H, W, device=gout.device, dtype=torch.float32)
        with _guard(gout.device):
            check(_lib.lib().rc_upflow8_bwd(gout.data_ptr(), B, Cn, H, W, dflow.data_ptr(), _stream(gout)),
                  "rc_upflow8_bwd")
        return dflow


def upsample_flow(flow: torch.Tensor, mask: torch.Tensor) -> torch.Tensor:
    """Convex-combination upsampling [B,2,H,W] -> [B,2,8H,8W] (core/raft.py:72-83).  `mask` is the update block's
    [B, 576, H, W] output (any float dtype: autocast hands over fp16; it is read as fp32 like the reference's
    softmax does under autocast)."""
    flow = _require_cuda_f32(flow, "flow")
    grad = torch.is_grad_enabled() and (flow.requires_grad or (isinstance(mask, torch.Tensor) and mask.requires_grad))
    if isinstance(mask, torch.Tensor) and mask.is_cuda and mask.dtype is torch.float16 and not grad:
        mask = mask.contiguous()             # inference under fp16 autocast: read as it is, no fp32 copy
    else:
        mask = _require_cuda_f32(mask, "mask")
    if flow.dim() != 4 or flow.shape[1] != 2:
        raise RaftCorrError(f"flow must be [B,2,H,W], got {tuple(flow.shape)}")
    if mask.dim() != 4 or mask.shape[1] != 576 or mask.shape[0] != flow.shape[0] or mask.shape[2:] != flow.shape[2:]:
        raise RaftCorrError(f"mask must be [B,576,H,W] matching flow {tuple(flow.shape)}, got {tuple(mask.shape)}")
    if grad:
        return _UpsampleFn.apply(flow, mask)
    return _upsample_fwd(flow, mask)          # inference: no autograd node to build


def upflow8(flow: torch.Tensor, mode: str = "bilinear") -> torch.Tensor:
    """8 * bilinear x8 interpolation, align_corners=True (core/utils/utils.py:80-82)."""
    if mode != "bilinear":
        raise RaftCorrError(f"upflow8: only mode='bilinear' (the reference's default and only use), got {mode!r}")
    flow = _require_cuda_f32(flow, "flow")
    if flow.dim() != 4:
        raise RaftCorrError(f"flow must be [B,C,H,W], got {tuple(flow.shape)}")
    return _Upflow8Fn.apply(flow)
