"""RAFT end-to-end harness around the frozen reference sub-networks (test infrastructure).

The three networks (feature encoder, context encoder, update block) are the reference's own
modules, traced by oracle/make_ref_modules.py; this file re-states only the control flow of
RAFT.forward (core/raft.py:86-144) so that the correlation block is a parameter:

    corr_cls(fmap1, fmap2, radius=4) -> corr_fn;  corr_fn(coords1) -> [B, 324, H/8, W/8]

which is exactly how the reference uses CorrBlock / AlternateCorrBlock (core/raft.py:104-107,
:124).  Used by the GPU end-to-end test, bench.py's extras and nothing else.
"""
from __future__ import annotations

import os

import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
FILES = ("raft_basic_fnet.pt", "raft_basic_cnet.pt", "raft_basic_update.pt")
SMALL_FILES = ("raft_small_fnet.pt", "raft_small_cnet.pt", "raft_small_update.pt")
HDIM = CDIM = 128   # core/raft.py:35-36
SMALL_HDIM, SMALL_CDIM, SMALL_RADIUS = 96, 64, 3   # core/raft.py:29-33


def available(small: bool = False) -> bool:
    return all(os.path.exists(os.path.join(REF_DIR, f)) for f in (SMALL_FILES if small else FILES))


def load(device, small: bool = False):
    """(fnet, cnet, update_block) TorchScript modules on `device` (RAFT-basic, or RAFT-small)."""
    return tuple(torch.jit.load(os.path.join(REF_DIR, f), map_location=device).eval()
                 for f in (SMALL_FILES if small else FILES))


def upflow8(flow):
    """core/utils/utils.py:80-82"""
    return 8 * F.interpolate(flow, size=(8 * flow.shape[2], 8 * flow.shape[3]), mode="bilinear", align_corners=True)


@torch.no_grad()
def raft_small_forward(nets, image1, image2, iters=12, corr_cls=None, upflow_fn=None):
    """RAFT.forward(test_mode=True) of the SMALL model (core/raft.py:86-144 with args.small): D = 128 features,
    correlation radius 3, no upsampling mask -> upflow8 (core/raft.py:131-133)."""
    fnet, cnet, update = nets
    corr_cls = corr_cls or TorchCorrBlock
    image1 = (2 * (image1 / 255.0) - 1.0).contiguous()
    image2 = (2 * (image2 / 255.0) - 1.0).contiguous()
    B = image1.shape[0]
    fmaps = fnet(torch.cat([image1, image2], dim=0)).float()
    corr_fn = corr_cls(fmaps[:B], fmaps[B:], radius=SMALL_RADIUS)
    net, inp = torch.split(cnet(image1), [SMALL_HDIM, SMALL_CDIM], dim=1)
    net, inp = torch.tanh(net), torch.relu(inp)
    H8, W8 = image1.shape[2] // 8, image1.shape[3] // 8
    coords0 = coords_grid(B, H8, W8, image1.device)
    coords1 = coords0.clone()
    for _ in range(iters):
        corr = corr_fn(coords1)
        net, delta = update(net, inp, corr, coords1 - coords0)     # the traced wrapper drops the None mask
        coords1 = coords1 + delta
    return coords1 - coords0, (upflow_fn or upflow8)(coords1 - coords0)


def coords_grid(batch, ht, wd, device):
    ys, xs = torch.meshgrid(torch.arange(ht, device=device), torch.arange(wd, device=device), indexing="ij")
    return torch.stack([xs, ys], dim=0).float()[None].repeat(batch, 1, 1, 1)   # core/utils/utils.py:74-77


def upsample_flow(flow, mask):
    """convex 8x upsampling, core/raft.py:72-83"""
    N, _, H, W = flow.shape
    mask = torch.softmax(mask.view(N, 1, 9, 8, 8, H, W), dim=2)
    up = F.unfold(8 * flow, [3, 3], padding=1).view(N, 2, 9, 1, 1, H, W)
    up = torch.sum(mask * up, dim=2).permute(0, 1, 4, 2, 5, 3)
    return up.reshape(N, 2, 8 * H, 8 * W)


class TorchCorrBlock:
    """The reference CorrBlock's arithmetic (torch library calls, oracle/torch_port.py) on whatever
    device the features live on: the 'reference through stock torch' arm of the comparison."""

    def __init__(self, fmap1, fmap2, num_levels=4, radius=4):
        from . import torch_port as P
        self._P = P
        self.radius = radius
        self.pyr = P.build_pyramid(fmap1, fmap2, num_levels)

    def __call__(self, coords):
        return self._P.lookup(self.pyr, coords, self.radius)


def update_after_convc1(update, net, inp, cor, flow):
    """BasicUpdateBlock.forward (core/update.py:131-139) with BasicMotionEncoder.forward (:88-95) inlined and its first
    line, `cor = F.relu(self.convc1(corr))`, already done by the caller: `cor` is that activation.  Every layer is the
    traced reference submodule; only the wiring between them is restated here (for the fused lookup + convc1 check)."""
    enc = update.encoder
    cor = F.relu(enc.convc2(cor))
    flo = F.relu(enc.convf1(flow))
    flo = F.relu(enc.convf2(flo))
    out = F.relu(enc.conv(torch.cat([cor, flo], dim=1)))
    motion = torch.cat([out, flow], dim=1)
    net = update.gru(net, torch.cat([inp, motion], dim=1))
    return net, 0.25 * update.mask(net), update.flow_head(net)


@torch.no_grad()
def raft_forward(nets, image1, image2, iters=12, corr_cls=TorchCorrBlock, last_only=True, upsample_fn=None,
                 mixed_precision=False, fuse_convc1=False):
    """RAFT.forward(test_mode=True) (core/raft.py:86-144) -> (flow_low, flow_up).  mixed_precision wraps the three
    network calls in fp16 autocast exactly where the reference does (core/raft.py:98-100, :110-114, :127-128).
    fuse_convc1: corr_cls's block offers lookup_convc1(coords, weight, bias) (lookup + the motion encoder's first
    layer in one pass); the rest of the update block runs through update_after_convc1."""
    fnet, cnet, update = nets
    amp = lambda: torch.autocast("cuda", dtype=torch.float16, enabled=mixed_precision)   # noqa: E731
    image1 = (2 * (image1 / 255.0) - 1.0).contiguous()
    image2 = (2 * (image2 / 255.0) - 1.0).contiguous()
    B = image1.shape[0]
    # the reference feeds [image1, image2] as a list, which its encoder concatenates on the batch
    # axis and splits again (core/extractor.py:171-174, :189-190)
    with amp():
        fmaps = fnet(torch.cat([image1, image2], dim=0))
    fmaps = fmaps.float()
    fmap1, fmap2 = fmaps[:B], fmaps[B:]
    corr_fn = corr_cls(fmap1, fmap2, radius=4)
    with amp():
        net, inp = torch.split(cnet(image1), [HDIM, CDIM], dim=1)
        net, inp = torch.tanh(net), torch.relu(inp)
    H8, W8 = image1.shape[2] // 8, image1.shape[3] // 8
    coords0 = coords_grid(B, H8, W8, image1.device)
    coords1 = coords0.clone()
    flow_up = None
    if fuse_convc1:
        w1, b1 = update.encoder.convc1.weight, update.encoder.convc1.bias
    for it in range(iters):
        if fuse_convc1:
            cor = corr_fn.lookup_convc1(coords1, w1, b1)
            with amp():
                net, up_mask, delta = update_after_convc1(update, net, inp, cor, coords1 - coords0)
        else:
            corr = corr_fn(coords1)
            with amp():
                net, up_mask, delta = update(net, inp, corr, coords1 - coords0)
        coords1 = coords1 + delta
        if not last_only or it == iters - 1:
            flow_up = (upsample_fn or upsample_flow)(coords1 - coords0, up_mask)   # the reference: every iteration
    return coords1 - coords0, flow_up


def raft_forward_train(nets, image1, image2, iters=12, corr_cls=TorchCorrBlock, upsample_fn=None):
    """RAFT.forward(test_mode=False) (core/raft.py:86-144): the list of upsampled flow predictions, with
    autograd history.  coords1 is detached every iteration exactly as the reference does (:123)."""
    fnet, cnet, update = nets
    image1 = (2 * (image1 / 255.0) - 1.0).contiguous()
    image2 = (2 * (image2 / 255.0) - 1.0).contiguous()
    B = image1.shape[0]
    fmaps = fnet(torch.cat([image1, image2], dim=0)).float()
    fmap1, fmap2 = fmaps[:B], fmaps[B:]
    corr_fn = corr_cls(fmap1, fmap2, radius=4)
    net, inp = torch.split(cnet(image1), [HDIM, CDIM], dim=1)
    net, inp = torch.tanh(net), torch.relu(inp)
    H8, W8 = image1.shape[2] // 8, image1.shape[3] // 8
    coords0 = coords_grid(B, H8, W8, image1.device)
    coords1 = coords0.clone()
    preds = []
    for _ in range(iters):
        coords1 = coords1.detach()
        corr = corr_fn(coords1)
        net, up_mask, delta = update(net, inp, corr, coords1 - coords0)
        coords1 = coords1 + delta
        preds.append((upsample_fn or upsample_flow)(coords1 - coords0, up_mask))
    return preds


def sequence_loss(preds, flow_gt, valid, gamma=0.8, max_flow=400.0):
    """train.py:47-72 (the loss only; metrics left out)."""
    n = len(preds)
    mag = torch.sum(flow_gt ** 2, dim=1).sqrt()
    valid = (valid >= 0.5) & (mag < max_flow)
    loss = 0.0
    for i, p in enumerate(preds):
        w = gamma ** (n - i - 1)
        loss = loss + w * (valid[:, None] * (p - flow_gt).abs()).mean()
    return loss
