"""HBM probes on the GPU box: pure write (fill / memset) and copy, CUDA events, best of 10.
Why it exists: a write-only kernel cannot reach the copy bandwidth MEASURED_PEAKS.json quotes
(copy = read + write sharing the bus); the build kernel's ceiling is the fill number."""
import json
import torch

dev = torch.device("cuda:0")
nbytes = 2 * 1024 ** 3
a = torch.empty(nbytes // 2, dtype=torch.bfloat16, device=dev)
b = torch.empty(nbytes // 2, dtype=torch.bfloat16, device=dev)


def best_ms(fn, reps=10):
    fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


out = {
    "fill_gbs": nbytes / best_ms(lambda: a.fill_(1.0)) / 1e6,
    "memset_gbs": nbytes / best_ms(lambda: a.zero_()) / 1e6,
    "copy_gbs": 2 * nbytes / best_ms(lambda: b.copy_(a)) / 1e6,
    "bytes": nbytes,
}
print(json.dumps(out))
