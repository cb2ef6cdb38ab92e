"""The refinement loop as ONE CUDA graph (reference: core/raft.py:122-139).

At one frame pair RAFT's forward is launch-bound: 12 iterations x (lookup + the update block's ~25 cuDNN / elementwise
kernels + upsampling) are some 400 launches of a few microseconds each, issued from Python.  Every entry point of
libraftcorr is capturable (no allocation and no synchronisation inside the C ABI, launches go to the capturing stream,
TMA descriptors are encoded on the host at capture time), so the whole forward -- encoders, volume build, the twelve
lookups, update blocks, upsampling -- can be recorded once and replayed.

    fwd = raft_stir_b200.GraphedCallable(lambda im1, im2: model(im1, im2, iters=12, test_mode=True), (im1, im2))
    flow_low, flow_up = fwd(next_im1, next_im2)      # copies the inputs in, replays, returns the captured outputs

The callable must be shape-static and inference-only (torch.no_grad is applied); tensors it allocates live in the
graph's private memory pool and are reused by every replay, so outputs are valid until the next call (clone to keep).
"""
from __future__ import annotations

from typing import Callable, Sequence

import torch


def _flatten(out):
    if isinstance(out, torch.Tensor):
        return [out]
    if isinstance(out, (list, tuple)):
        flat = []
        for o in out:
            flat.extend(_flatten(o))
        return flat
    return []


class GraphedCallable:
    """fn(*inputs) captured into a CUDA graph on the inputs' device.

    warmup: eager runs on a side stream before the capture (cuDNN picks its algorithms, the allocator settles, this
            library resolves its per-device state); at least 1.
    """

    def __init__(self, fn: Callable, example_inputs: Sequence[torch.Tensor], warmup: int = 3):
        if not example_inputs or not all(isinstance(t, torch.Tensor) and t.is_cuda for t in example_inputs):
            raise RuntimeError("GraphedCallable needs CUDA tensors as example inputs (there is no CPU path)")
        self.device = example_inputs[0].device
        self._static_in = [t.detach().clone() for t in example_inputs]
        self._graph = torch.cuda.CUDAGraph()
        with torch.no_grad(), torch.cuda.device(self.device):
            side = torch.cuda.Stream(self.device)
            side.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(side):
                for _ in range(max(1, warmup)):
                    fn(*self._static_in)
            torch.cuda.current_stream(self.device).wait_stream(side)
            torch.cuda.synchronize(self.device)
            with torch.cuda.graph(self._graph):
                self._out = fn(*self._static_in)
        self._flat_out = _flatten(self._out)

    def __call__(self, *inputs: torch.Tensor):
        if len(inputs) != len(self._static_in):
            raise RuntimeError(f"expected {len(self._static_in)} inputs, got {len(inputs)}")
        for dst, src in zip(self._static_in, inputs):
            if src.shape != dst.shape or src.dtype != dst.dtype:
                raise RuntimeError(f"input {tuple(src.shape)} {src.dtype} differs from the captured "
                                   f"{tuple(dst.shape)} {dst.dtype}: a CUDA graph is shape-static")
            if src.data_ptr() != dst.data_ptr():
                dst.copy_(src, non_blocking=True)
        self._graph.replay()
        return self._out

    def replay(self):
        """Replay on whatever the static input buffers hold (no copies): `static_inputs` may be written in place."""
        self._graph.replay()
        return self._out

    @property
    def static_inputs(self):
        return self._static_in
