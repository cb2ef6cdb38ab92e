#!/usr/bin/env python3
"""BASELINE config 5: one RAFT training step at the FlyingChairs shape (368x496, batch 10 per GPU, 12
iterations, sequence loss gamma 0.8, AdamW, clip 1.0), reference sub-networks (oracle/_ref) around
(a) the reference CorrBlock's torch ops and (b) this library's autograd-enabled CorrBlock.  One process
per GPU; with WORLD_SIZE > 1 the networks are wrapped in DistributedDataParallel (NCCL all-reduce of the
5.26 M parameter gradients).  Prints one JSON line on rank 0: ms per step (max over ranks) and pairs/s.

    python tests/perf/train_step_bench.py                       # 1 GPU
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tests/perf/train_step_bench.py
"""
import json
import os
import sys

import torch
import torch.nn as nn

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import raft_stir_b200 as R              # noqa: E402
from oracle import raft_harness as Hn   # noqa: E402


class Net(nn.Module):
    def __init__(self, nets, corr_cls, upsample_fn=None):
        super().__init__()
        self.fnet, self.cnet, self.update = nets
        self.corr_cls = corr_cls
        self.upsample_fn = upsample_fn

    def forward(self, im1, im2):
        return Hn.raft_forward_train((self.fnet, self.cnet, self.update), im1, im2, 12, self.corr_cls,
                                     upsample_fn=self.upsample_fn)


def main():
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    B, H, W = 10, 368, 496
    g = torch.Generator(device="cpu").manual_seed(100 + rank)
    im1 = (torch.rand(B, 3, H, W, generator=g) * 255).to(dev)
    im2 = (torch.rand(B, 3, H, W, generator=g) * 255).to(dev)
    gt = (5 * torch.randn(B, 2, H, W, generator=g)).to(dev)
    valid = torch.ones(B, H, W, device=dev)
    res = {"config": f"RAFT-basic train step, {B} pairs/GPU 368x496, 12 iters, fp32, {world} GPU(s)"}
    for name, cls, ups in (("torch_corrblock", Hn.TorchCorrBlock, None), ("ours_bf16", R.CorrBlock, None),
                           ("ours_bf16_and_upsampling", R.CorrBlock, R.upsample_flow)):
        model = Net(Hn.load(dev), cls, ups).to(dev)
        if world > 1:
            model = nn.parallel.DistributedDataParallel(model, device_ids=[local])
        opt = torch.optim.AdamW(model.parameters(), lr=1e-5, weight_decay=1e-4)

        def step():
            opt.zero_grad(set_to_none=True)
            loss = Hn.sequence_loss(model(im1, im2), gt, valid)
            loss.backward()
            torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
            opt.step()
            return loss

        for _ in range(3):
            step()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            loss = step()
        e1.record(); torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / 3], device=dev)
        if dist is not None:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        res[name] = {"ms_per_step": float(ms.item()), "pairs_per_s": 1e3 * B * world / float(ms.item()),
                     "loss": float(loss.item()), "peak_mem_GB": torch.cuda.max_memory_allocated(dev) / 1e9}
        del model, opt
        torch.cuda.empty_cache(); torch.cuda.reset_peak_memory_stats(dev)
    if rank == 0:
        print(json.dumps(res))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
