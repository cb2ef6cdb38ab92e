"""Drop-in `CorrBlock` / `AlternateCorrBlock` (reference: core/corr.py).

Same constructor `(fmap1, fmap2, num_levels=4, radius=4)`, same call
`corr_fn(coords) -> [B, num_levels*(2r+1)^2, H, W] fp32`, same attributes
(`num_levels`, `radius`, `corr_pyramid` / `pyramid`), so `RAFT.forward`
(core/raft.py:104-107, :124) runs unchanged once the two names are swapped:

    import raft, raft_stir_b200
    raft_stir_b200.install_into(raft)

Everything numeric happens in libraftcorr.so through the C ABI
(include/raft_corr.h); torch is used for device memory, the current stream and
autograd bookkeeping only.  CPU tensors are rejected: there is no fallback.
"""
from __future__ import annotations

import ctypes as C
import os
import weakref
from typing import List, Optional

import torch

from . import _lib
from ._lib import RaftCorrError, check

_STORAGE = {"bf16": _lib.RC_VOL_BF16, "fp32": _lib.RC_VOL_F32, "fp16": _lib.RC_VOL_F16}
_MATH = {"bf16": _lib.RC_MATH_BF16, "fp32": _lib.RC_MATH_F32, "tf32": _lib.RC_MATH_TF32, "fp16": _lib.RC_MATH_F16}
_TORCH_DT = {_lib.RC_VOL_BF16: torch.bfloat16, _lib.RC_VOL_F32: torch.float32, _lib.RC_VOL_F16: torch.float16}


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def _stream(t: torch.Tensor) -> int:
    """The current stream of the tensor's device as a raw cudaStream_t (the fast private accessor when this torch
    has it: these wrappers run once per refinement iteration and their host time is what a single pair waits for)."""
    if _raw_stream is not None:
        return _raw_stream(t.device.index)
    return torch.cuda.current_stream(t.device).cuda_stream


class _guard:
    """`with torch.cuda.device(dev)` without its cost when dev already is the current device (the C ABI launches on
    the current device; nn.DataParallel replicas call from threads whose current device is their own)."""
    __slots__ = ("idx", "prev")

    def __init__(self, device: torch.device):
        self.idx = device.index

    def __enter__(self):
        self.prev = torch.cuda.current_device()
        if self.idx is not None and self.prev != self.idx:
            torch.cuda.set_device(self.idx)

    def __exit__(self, *exc):
        if self.idx is not None and self.prev != self.idx:
            torch.cuda.set_device(self.prev)
        return False


def _require_cuda_f32(t: torch.Tensor, name: str) -> torch.Tensor:
    if isinstance(t, torch.Tensor) and t.is_cuda and t.dtype is torch.float32 and t.is_contiguous():
        return t
    if not isinstance(t, torch.Tensor):
        raise RaftCorrError(f"{name} must be a torch.Tensor")
    if not t.is_cuda:
        raise RaftCorrError(f"{name} must be a CUDA tensor (libraftcorr has no CPU path)")
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def coords_grid(batch: int, ht: int, wd: int, device) -> torch.Tensor:
    """[B,2,H,W], channel 0 = x, channel 1 = y (core/utils/utils.py:74-77)."""
    ys, xs = torch.meshgrid(torch.arange(ht, device=device), torch.arange(wd, device=device), indexing="ij")
    return torch.stack([xs, ys], dim=0).float()[None].repeat(batch, 1, 1, 1)


_OP16 = {torch.bfloat16: _lib.RC_VOL_BF16, torch.float16: _lib.RC_VOL_F16}


def _is_nhwc16(t) -> bool:
    return (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype in _OP16 and t.dim() == 4
            and t.shape[1] % 64 == 0)


def pack_features(fmap1: torch.Tensor, fmap2: torch.Tensor, dtype: torch.dtype = torch.bfloat16):
    """fp32 NCHW feature maps -> 16-bit channels-last tensors (logical shape [B, D, H, W], memory NHWC), the form
    `CorrBlock` consumes without any repack (rc_build_pyramid_nhwc16) and the form a channels_last bf16 encoder would
    emit directly (core/extractor.py:144,184 -> core/raft.py:99-103).  One transposing cast per map."""
    if dtype not in _OP16:
        raise RaftCorrError(f"dtype={dtype!r}: expected torch.bfloat16 or torch.float16")
    f1 = _require_cuda_f32(fmap1.detach(), "fmap1")
    f2 = _require_cuda_f32(fmap2.detach(), "fmap2")
    if f1.shape != f2.shape or f1.dim() != 4:
        raise RaftCorrError(f"fmap shapes {tuple(f1.shape)} vs {tuple(f2.shape)}: expected equal [B,D,H,W]")
    B, D, H, W = (int(v) for v in f1.shape)
    o1 = torch.empty((B, D, H, W), dtype=dtype, device=f1.device, memory_format=torch.channels_last)
    o2 = torch.empty((B, D, H, W), dtype=dtype, device=f1.device, memory_format=torch.channels_last)
    with _guard(f1.device):
        check(_lib.lib().rc_pack_features_nhwc16(f1.data_ptr(), f2.data_ptr(), B, D, H, W, _OP16[dtype], o1.data_ptr(),
                                                 o2.data_ptr(), _stream(f1)), "rc_pack_features_nhwc16")
    return o1, o2


def pack_convc1_weights(weight: torch.Tensor) -> torch.Tensor:
    """[O, C(,1,1)] fp32 convolution weight -> the fp16 [O, C rounded up to 8] operand rc_lookup_convc1_fwd reads."""
    w = _require_cuda_f32(weight.detach().reshape(weight.shape[0], -1), "weight")
    O, C_ = int(w.shape[0]), int(w.shape[1])
    L = _lib.lib()
    nbytes = int(L.rc_convc1_weight_bytes(O, C_))
    if nbytes <= 0:
        raise RaftCorrError(f"convc1 weight [{O}, {C_}]: the output channels must be a multiple of 16, at most 256")
    out = torch.empty(nbytes // 2, dtype=torch.float16, device=w.device)
    with _guard(w.device):
        check(L.rc_convc1_pack_weights(w.data_ptr(), O, C_, out.data_ptr(), _stream(w)), "rc_convc1_pack_weights")
    return out


class LazyCorr:
    """What a CorrBlock(fuse_convc1=True) returns from `corr_fn(coords)`: the block and the coordinates, not yet
    looked up.  RAFT.forward only passes it on to the update block (core/raft.py:124-128), whose motion encoder
    applies convc1 first (core/update.py:89): a FusedConvc1 in that place computes lookup + convolution + ReLU in
    one fused pass.  `materialize()` gives the plain [B, L*(2r+1)^2, H, W] tensor for any other consumer."""
    __slots__ = ("block", "coords")

    def __init__(self, block, coords):
        self.block, self.coords = block, coords

    def materialize(self) -> torch.Tensor:
        return self.block._lookup(self.coords, self.block.out_dtype)

    @property
    def shape(self):
        n = 2 * self.block.radius + 1
        return torch.Size((self.block.B, self.block.num_levels * n * n, self.block.H, self.block.W))


class FusedConvc1(torch.nn.Module):
    """Drop-in for the motion encoder's `convc1` (core/update.py:62, :82).  Given a LazyCorr it returns
    relu(convc1(lookup)) from the fused kernels (the encoder's own F.relu on top is then the identity); given a
    tensor it is the wrapped convolution, so training and every other path are unchanged."""

    def __init__(self, conv: torch.nn.Conv2d):
        super().__init__()
        if tuple(conv.kernel_size) != (1, 1) or conv.bias is None:
            raise RaftCorrError("FusedConvc1 wraps a 1x1 convolution with bias")
        self.conv = conv
        self._packed = None
        self._key = None

    @property
    def weight(self):
        return self.conv.weight

    @property
    def bias(self):
        return self.conv.bias

    def forward(self, corr):
        if not isinstance(corr, LazyCorr):
            return self.conv(corr)
        w = self.conv.weight
        key = (w.data_ptr(), w._version, w.device)
        if self._key != key:                      # converted once per weight version (inference: once)
            self._packed, self._key = pack_convc1_weights(w), key
        out_dtype = torch.float16 if torch.is_autocast_enabled() and torch.get_autocast_gpu_dtype() is torch.float16 \
            else torch.float32
        return corr.block.lookup_convc1(corr.coords, w, self.conv.bias, out_dtype, self._packed)


def fuse_motion_encoder(model) -> int:
    """Wrap `convc1` of every motion encoder inside `model` (a reference RAFT, possibly inside nn.DataParallel) with
    FusedConvc1; returns how many were wrapped.  Use together with install_into(raft_module, fuse_convc1=True)."""
    n = 0
    for mod in model.modules():
        conv = getattr(mod, "convc1", None)
        if isinstance(conv, torch.nn.Conv2d) and tuple(conv.kernel_size) == (1, 1):
            mod.convc1 = FusedConvc1(conv)
            n += 1
    return n


class _BuildFn(torch.autograd.Function):
    """Autograd node of the volume build.  Its output is a one-element token
    that every lookup takes as an input, so autograd runs this backward once,
    after all lookup backwards have accumulated into the gradient pyramid."""

    @staticmethod
    def forward(ctx, fmap1, fmap2, blk):
        blk._build(fmap1, fmap2)
        # weak: the block owns the token whose grad_fn is this node; a strong reference back would make a cycle that
        # only the cyclic GC frees, keeping a step's pyramid and gradient pyramid (GBs at training batch sizes) alive.
        # Every lookup node holds the block strongly, so it lives as long as any lookup output is part of the graph.
        ctx.blk_ref = weakref.ref(blk)
        ctx.save_for_backward(fmap1, fmap2)
        return torch.zeros(1, device=fmap1.device, dtype=torch.float32)

    @staticmethod
    def backward(ctx, _token_grad):
        fmap1, fmap2 = ctx.saved_tensors
        blk = ctx.blk_ref()
        if blk is None:                    # no lookup output survived: nothing flowed into the volume
            return torch.zeros_like(fmap1), torch.zeros_like(fmap2), None
        d1, d2 = blk._build_backward(fmap1, fmap2)
        return d1, d2, None


class _LookupFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, coords, token, blk):
        ctx.blk = blk
        ctx.save_for_backward(coords)
        ctx.has_token = token is not None
        return blk._lookup(coords)

    @staticmethod
    def backward(ctx, grad_out):
        (coords,) = ctx.saved_tensors
        blk = ctx.blk
        want_vol = ctx.has_token and ctx.needs_input_grad[1]
        dcoords = blk._lookup_backward(coords, grad_out.contiguous(), want_vol, ctx.needs_input_grad[0])
        token_grad = torch.zeros(1, device=coords.device) if want_vol else None
        return dcoords, token_grad, None


class CorrBlock:
    """All-pairs correlation volume, 4-level pooled pyramid and windowed lookup.

    storage: "bf16" (default), "fp16" or "fp32" -- element type of the stored pyramid.  fp16 costs the same bytes
             as bf16 and keeps three more mantissa bits; its range ends at 65 504, which correlations of
             normalised features (the encoder ends in instance norm) stay far below.  With math="tf32" it is the
             near-fp32 setting at the default's speed.
    math:    "bf16" (default; tcgen05 tensor cores, fp32 accumulate), "fp16" (the same at the same speed on
             IEEE-half operands: 10 mantissa bits instead of 7, features beyond +-65 504 saturate), "tf32" (tensor
             cores on the fp32 features rounded to tf32; half the MMA rate, twice the operand bytes) or "fp32"
             (CUDA-core fp32 FMA, the reference's SGEMM arithmetic).
    out_dtype: torch.float32 (default, the reference's) or torch.float16 -- element type of what the call returns.
             fp16 is for inference under fp16 autocast (--mixed_precision): the update block's first convolution
             would cast the fp32 result itself; the lookup then writes half the bytes and the cast pass disappears.
    Defaults can be set with RAFT_CORR_STORAGE / RAFT_CORR_MATH / RAFT_CORR_OUT (fp32|fp16).

    fuse_convc1: in inference the call returns a LazyCorr (block + coordinates) instead of the 324-channel tensor; the
             motion encoder's first layer, wrapped by `fuse_motion_encoder(model)`, resolves it with ONE fused pass
             (lookup -> fp16 samples -> tensor-core 1x1 convolution + bias + ReLU, rc_lookup_convc1_fwd).  Anything else
             that touches the LazyCorr as a tensor gets the ordinary lookup through `.materialize()`.

    Features that are already 16-bit (torch.bfloat16 / torch.float16, D % 64 == 0, no gradient wanted) -- what an
    encoder run in channels_last bf16 hands over, or `pack_features` -- are consumed in place: no fp32 round trip, no
    repack, no workspace (rc_build_pyramid_nhwc16); `math` is then the tensors' own type.
    """

    def __init__(self, fmap1, fmap2, num_levels: int = 4, radius: int = 4, *,
                 storage: Optional[str] = None, math: Optional[str] = None, out_dtype: Optional[torch.dtype] = None,
                 fuse_convc1: bool = False):
        self.num_levels = num_levels
        self.radius = radius
        self.fuse_convc1 = bool(fuse_convc1)
        if out_dtype is None:
            out_dtype = {"fp32": torch.float32, "fp16": torch.float16}.get(os.environ.get("RAFT_CORR_OUT", "fp32"))
        if out_dtype not in (torch.float32, torch.float16):
            raise RaftCorrError(f"out_dtype={out_dtype!r}: expected torch.float32 or torch.float16")
        self.out_dtype = out_dtype
        storage = storage or os.environ.get("RAFT_CORR_STORAGE", "bf16")
        math = math or os.environ.get("RAFT_CORR_MATH", "bf16")
        if storage not in _STORAGE or math not in _MATH:
            raise RaftCorrError(f"storage={storage!r} math={math!r}: expected storage bf16|fp16|fp32, math bf16|fp16|tf32|fp32")
        self.vol_dtype = _STORAGE[storage]
        self.math = _MATH[math]
        _lib.lib()   # fail loudly before touching tensors if the extension is missing

        self._nhwc16 = (_is_nhwc16(fmap1) and _is_nhwc16(fmap2) and fmap1.dtype == fmap2.dtype
                        and not (torch.is_grad_enabled() and (fmap1.requires_grad or fmap2.requires_grad)))
        if self._nhwc16:
            f1 = fmap1.detach().contiguous(memory_format=torch.channels_last)     # no copy when the encoder ran channels_last
            f2 = fmap2.detach().contiguous(memory_format=torch.channels_last)
            self.math = _lib.RC_MATH_BF16 if f1.dtype is torch.bfloat16 else _lib.RC_MATH_F16
        else:
            f1 = _require_cuda_f32(fmap1, "fmap1")
            f2 = _require_cuda_f32(fmap2, "fmap2")
        if f1.shape != f2.shape or f1.dim() != 4:
            raise RaftCorrError(f"fmap shapes {tuple(f1.shape)} vs {tuple(f2.shape)}: expected equal [B,D,H,W]")
        if f1.device != f2.device:
            raise RaftCorrError("fmap1 and fmap2 must be on the same device")
        self.B, self.D, self.H, self.W = (int(v) for v in f1.shape)
        self.device = f1.device
        self.layout = _lib.pyramid_layout(self.B, self.H, self.W, num_levels, self.vol_dtype)
        self._buf = None
        self._grad_buf = None
        self._token = None

        if torch.is_grad_enabled() and (f1.requires_grad or f2.requires_grad):
            self._token = _BuildFn.apply(f1, f2, self)
        else:
            self._build(f1, f2)

    # ------------------------------------------------------------------ build
    def _build(self, f1: torch.Tensor, f2: torch.Tensor) -> None:
        L = _lib.lib()
        lay = self.layout
        self._buf = torch.empty(int(lay.total_elems), dtype=_TORCH_DT[self.vol_dtype], device=self.device)
        if self._nhwc16:
            with _guard(self.device):
                check(L.rc_build_pyramid_nhwc16(f1.data_ptr(), f2.data_ptr(), _OP16[f1.dtype], self.B, self.D, self.H,
                                                self.W, self.num_levels, self.vol_dtype, self._buf.data_ptr(),
                                                _stream(f1)), "rc_build_pyramid_nhwc16")
            return
        ws_bytes = int(L.rc_build_workspace_bytes(self.B, self.D, self.H, self.W, self.math))
        if ws_bytes < 0:
            raise RaftCorrError("rc_build_workspace_bytes rejected the sizes")
        ws = torch.empty(max(ws_bytes, 1), dtype=torch.uint8, device=self.device)
        with _guard(self.device):
            check(L.rc_build_pyramid(f1.data_ptr(), f2.data_ptr(), self.B, self.D, self.H, self.W,
                                     self.num_levels, self.vol_dtype, self.math, self._buf.data_ptr(),
                                     ws.data_ptr(), ws_bytes, _stream(f1)), "rc_build_pyramid")
        # (ws was allocated on the current stream and the kernels run on it: the caching allocator's own stream order
        # covers its reuse, no record_stream needed -- that call costs more host time than the launch)

    @property
    def corr_pyramid(self) -> List[torch.Tensor]:
        """Levels as [B*N, 1, H_l, W_l] tensors, the reference's shape
        (core/corr.py:21-27).  The buffer itself is tiled (rc_layout), so these
        are gathered copies for inspection and tests; the lookup reads the buffer."""
        return [_lib.level_view(self._buf, self.layout, l).unsqueeze(1) for l in range(self.num_levels)]

    @staticmethod
    def corr(fmap1, fmap2):
        """[B,H,W,1,H,W] fp32 volume, 1/sqrt(D)-scaled (core/corr.py:52-60)."""
        blk = CorrBlock(fmap1.detach(), fmap2.detach(), num_levels=1, storage="fp32", math="fp32")
        B, H, W = blk.B, blk.H, blk.W
        return blk.corr_pyramid[0].reshape(B, H, W, 1, H, W)

    # ----------------------------------------------------------------- lookup
    def _check_coords(self, coords: torch.Tensor) -> torch.Tensor:
        c = _require_cuda_f32(coords, "coords")
        if tuple(c.shape) != (self.B, 2, self.H, self.W):
            raise RaftCorrError(f"coords shape {tuple(c.shape)}, expected {(self.B, 2, self.H, self.W)}")
        if c.device != self.device:
            raise RaftCorrError("coords must be on the pyramid's device")
        return c

    def _lookup(self, coords: torch.Tensor, out_dtype: torch.dtype = torch.float32) -> torch.Tensor:
        n = 2 * self.radius + 1
        out = torch.empty(self.B, self.num_levels * n * n, self.H, self.W, dtype=out_dtype, device=self.device)
        fn = _lib.lib().rc_lookup_fwd if out_dtype is torch.float32 else _lib.lib().rc_lookup_fwd_f16
        with _guard(self.device):
            check(fn(self._buf.data_ptr(), self.vol_dtype, coords.data_ptr(), self.B, self.H, self.W,
                     self.num_levels, self.radius, out.data_ptr(), _stream(coords)), "rc_lookup_fwd")
        return out

    def __call__(self, coords: torch.Tensor):
        c = self._check_coords(coords)
        if torch.is_grad_enabled() and (self._token is not None or c.requires_grad):
            return _LookupFn.apply(c, self._token, self)      # training: fp32, as the reference
        if self.fuse_convc1:
            return LazyCorr(self, c)                          # resolved by the motion encoder's first layer (FusedConvc1)
        return self._lookup(c, self.out_dtype)

    def lookup_convc1(self, coords: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor,
                      out_dtype: torch.dtype = torch.float32, packed_weight: Optional[torch.Tensor] = None) -> torch.Tensor:
        """relu(conv1x1(lookup(coords), weight, bias)): the lookup fused with the motion encoder's first layer
        (core/update.py:70-71, :88-90), [B, O, H, W].  weight: [O, C] or [O, C, 1, 1] fp32, C = num_levels*(2r+1)^2;
        packed_weight: what pack_convc1_weights(weight) returned (converted here when absent).  Inference only."""
        c = self._check_coords(coords.detach())
        L = _lib.lib()
        n = 2 * self.radius + 1
        C_ = self.num_levels * n * n
        O = int(weight.shape[0])
        if packed_weight is None:
            packed_weight = pack_convc1_weights(weight)
        if tuple(weight.shape[:2]) != (O, C_) or bias.numel() != O:
            raise RaftCorrError(f"weight {tuple(weight.shape)} / bias {tuple(bias.shape)}: expected [{O}, {C_}(,1,1)] and [{O}]")
        if out_dtype not in (torch.float32, torch.float16):
            raise RaftCorrError(f"out_dtype={out_dtype!r}: expected torch.float32 or torch.float16")
        b32 = _require_cuda_f32(bias.detach(), "bias")
        ws_bytes = int(L.rc_lookup_convc1_workspace_bytes(self.B, self.H, self.W, self.num_levels, self.radius))
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=self.device)
        out = torch.empty(self.B, O, self.H, self.W, dtype=out_dtype, device=self.device)
        with _guard(self.device):
            check(L.rc_lookup_convc1_fwd(self._buf.data_ptr(), self.vol_dtype, c.data_ptr(), self.B, self.H, self.W,
                                         self.num_levels, self.radius, packed_weight.data_ptr(), b32.data_ptr(), O,
                                         out.data_ptr(), _lib.RC_VOL_F32 if out_dtype is torch.float32 else _lib.RC_VOL_F16,
                                         ws.data_ptr(), ws_bytes, _stream(c)), "rc_lookup_convc1_fwd")
        return out

    # --------------------------------------------------------------- backward
    def _lookup_backward(self, coords, grad_out, want_vol: bool, want_coords: bool):
        L = _lib.lib()
        if want_vol and self._grad_buf is None:
            glay = _lib.pyramid_layout(self.B, self.H, self.W, self.num_levels, _lib.RC_VOL_F32)
            self._grad_buf = torch.zeros(int(glay.total_elems), dtype=torch.float32, device=self.device)
        dcoords = torch.empty_like(coords) if want_coords else None
        if not (want_vol or want_coords):
            return None
        with _guard(self.device):
            check(L.rc_lookup_bwd(self._buf.data_ptr(), self.vol_dtype, coords.data_ptr(), grad_out.data_ptr(),
                                  self.B, self.H, self.W, self.num_levels, self.radius,
                                  self._grad_buf.data_ptr() if want_vol else None,
                                  dcoords.data_ptr() if want_coords else None, _stream(coords)),
                  "rc_lookup_bwd")
        return dcoords

    def _build_backward(self, f1, f2):
        d1 = torch.empty_like(f1)
        d2 = torch.empty_like(f2)
        if self._grad_buf is None:   # no lookup contributed
            return d1.zero_(), d2.zero_()
        L = _lib.lib()
        ws_bytes = int(L.rc_build_bwd_workspace_bytes(self.B, self.D, self.H, self.W, self.num_levels))
        ws = torch.empty(max(ws_bytes, 1), dtype=torch.uint8, device=self.device)
        with _guard(self.device):
            # fp32-math blocks keep fp32-grade gradients (3xTF32); the 16-bit / tf32 builds take the one-pass tf32 backward
            bwd_math = _lib.RC_MATH_F32 if self.math == _lib.RC_MATH_F32 else _lib.RC_MATH_TF32
            check(L.rc_build_bwd(self._grad_buf.data_ptr(), f1.data_ptr(), f2.data_ptr(), self.B, self.D,
                                 self.H, self.W, self.num_levels, bwd_math, d1.data_ptr(), d2.data_ptr(),
                                 ws.data_ptr(), ws_bytes, _stream(f1)), "rc_build_bwd")
        ws.record_stream(torch.cuda.current_stream(self.device))
        self._grad_buf.record_stream(torch.cuda.current_stream(self.device))
        self._grad_buf = None
        return d1, d2


class AlternateCorrBlock:
    """Memory-lean variant: feature dot products fused with the bilinear taps,
    the volume is never built (reference: core/corr.py:63-91 calling
    alt_cuda_corr.forward once per level; here one launch covers all levels).
    Like the reference, the output carries no autograd history."""

    def __init__(self, fmap1, fmap2, num_levels: int = 4, radius: int = 4, *, math: Optional[str] = None):
        """math: "bf16" (the default when D % 64 == 0, which RAFT-basic's 256 and RAFT-small's 128 channels are: the
        block's local volume on the tensor cores, 16-bit operands, fp32 accumulate -- the precision the north star
        names, and 20-30x faster than the CUDA-core path), "fp16" (the same with three more mantissa bits of the
        features; saturates beyond +-65 504) or "fp32" (CUDA-core FMA, the reference kernel's arithmetic; the default
        for other channel counts and the opt-in for 1e-4 parity).  RAFT_ALT_MATH sets the default."""
        self.num_levels = num_levels
        self.radius = radius
        if math is None:
            math = os.environ.get("RAFT_ALT_MATH")
        if math is None:
            D_ = int(fmap1.shape[1]) if isinstance(fmap1, torch.Tensor) and fmap1.dim() == 4 else 0
            math = "bf16" if D_ >= 64 and D_ % 64 == 0 else "fp32"
        if math not in ("fp32", "bf16", "fp16"):
            raise RaftCorrError(f"math={math!r}: expected fp32|bf16|fp16")
        self.math = math
        L = _lib.lib()
        f1 = _require_cuda_f32(fmap1.detach(), "fmap1")
        f2 = _require_cuda_f32(fmap2.detach(), "fmap2")
        if f1.shape != f2.shape or f1.dim() != 4:
            raise RaftCorrError(f"fmap shapes {tuple(f1.shape)} vs {tuple(f2.shape)}: expected equal [B,D,H,W]")
        self.B, self.D, self.H, self.W = (int(v) for v in f1.shape)
        self.device = f1.device
        offs = (C.c_int64 * _lib.RC_MAX_LEVELS)()
        total = int(L.rc_alt_feature_pyramid_elems(self.B, self.D, self.H, self.W, num_levels, offs))
        if total <= 0:
            raise RaftCorrError("rc_alt_feature_pyramid_elems rejected the sizes")
        self._offs = [int(offs[l]) for l in range(num_levels)]
        self._f1 = torch.empty(self.B, self.H, self.W, self.D, dtype=torch.float32, device=self.device)
        self._f2 = torch.empty(total, dtype=torch.float32, device=self.device)
        with _guard(self.device):
            check(L.rc_alt_pool_features(f1.data_ptr(), f2.data_ptr(), self.B, self.D, self.H, self.W,
                                         num_levels, self._f1.data_ptr(), self._f2.data_ptr(), _stream(f1)),
                  "rc_alt_pool_features")
        if self.math != "fp32":
            if self.D % 64:
                raise RaftCorrError(f"math={self.math!r} needs D % 64 == 0 (got D={self.D})")
            dt = torch.bfloat16 if self.math == "bf16" else torch.float16
            self._op_dtype = _lib.RC_VOL_BF16 if self.math == "bf16" else _lib.RC_VOL_F16
            self._f1h = torch.empty(self.B, self.H, self.W, self.D, dtype=dt, device=self.device)
            self._f2h = torch.empty(total, dtype=dt, device=self.device)
            with _guard(self.device):
                check(L.rc_alt_features_16(self._f1.data_ptr(), self._f2.data_ptr(), self.B, self.D, self.H,
                                           self.W, num_levels, self._op_dtype, self._f1h.data_ptr(),
                                           self._f2h.data_ptr(), _stream(f1)), "rc_alt_features_16")
        # same attribute as the reference: [(fmap1, fmap2), pooled levels ...] as NCHW views
        self.pyramid = [(fmap1, fmap2)]
        for l in range(1, num_levels):
            h, w = self.H >> l, self.W >> l
            v = self._f2[self._offs[l]: self._offs[l] + self.B * h * w * self.D].view(self.B, h, w, self.D)
            self.pyramid.append((None, v.permute(0, 3, 1, 2)))

    def __call__(self, coords: torch.Tensor) -> torch.Tensor:
        c = _require_cuda_f32(coords.detach(), "coords")
        if tuple(c.shape) != (self.B, 2, self.H, self.W):
            raise RaftCorrError(f"coords shape {tuple(c.shape)}, expected {(self.B, 2, self.H, self.W)}")
        n = 2 * self.radius + 1
        out = torch.empty(self.B, self.num_levels * n * n, self.H, self.W, dtype=torch.float32, device=self.device)
        with _guard(self.device):
            if self.math != "fp32":
                check(_lib.lib().rc_alt_lookup_fwd_tc(self._f1h.data_ptr(), self._f2h.data_ptr(), self._op_dtype,
                                                      c.data_ptr(), self.B, self.D, self.H, self.W, self.num_levels,
                                                      self.radius, out.data_ptr(), _stream(c)), "rc_alt_lookup_fwd_tc")
            else:
                check(_lib.lib().rc_alt_lookup_fwd(self._f1.data_ptr(), self._f2.data_ptr(), c.data_ptr(), self.B,
                                                   self.D, self.H, self.W, self.num_levels, self.radius,
                                                   out.data_ptr(), _stream(c)), "rc_alt_lookup_fwd")
        return out


def install_into(raft_module, upsampling: bool = True, fuse_convc1: bool = False) -> None:
    """Swap the two names `core/raft.py:8` imports, e.g. install_into(sys.modules['raft']); with `upsampling`
    also rebind the flow upsampling that follows every iteration (core/raft.py:72-83, :133-135).  fuse_convc1: the
    blocks RAFT.forward creates hand the update block a LazyCorr (see CorrBlock); call fuse_motion_encoder(model) on
    the model as well, or the encoder's convolution will not know what to do with it."""
    if fuse_convc1:
        import functools
        raft_module.CorrBlock = functools.partial(CorrBlock, fuse_convc1=True)
    else:
        raft_module.CorrBlock = CorrBlock
    raft_module.AlternateCorrBlock = AlternateCorrBlock
    if upsampling:
        from . import upsample as _up
        if hasattr(raft_module, "RAFT"):
            raft_module.RAFT.upsample_flow = lambda self, flow, mask: _up.upsample_flow(flow, mask)
        if hasattr(raft_module, "upflow8"):
            raft_module.upflow8 = _up.upflow8
