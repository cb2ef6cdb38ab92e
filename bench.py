#!/usr/bin/env python
"""Benchmark of RAFT's correlation hot path on B200 (contract: see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One STEP = one pass of the hot path over one batch of synthetic frame pairs at
the Sintel shape (1/8-resolution features 55 x 128 x 256: BASELINE.json
configs[1], at the batch of configs[2], 64 pairs per GPU): volume build + 4-level pyramid, then 12 windowed lookups (r = 4)
at 12 different coordinate fields.  `value` = frame pairs per second over all
ranks with inputs resident in HBM; `e2e` = the same work through the public
CorrBlock API from pinned HOST buffers (features and coordinates copied in,
all 12 lookup outputs copied out) inside the timed region.

For N > 1 launch with torch.distributed.run; ranks shard the frame pairs
(weak scaling, no collective on the data path).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H, W, D, LEVELS, RADIUS, ITERS = 55, 128, 256, 4, 4, 12
N = H * W
PYR_ELEMS = sum(N * (H >> l) * (W >> l) for l in range(LEVELS))          # 65 331 200
BUILD_FLOPS = 2.0 * N * N * D                                           # 25.376 GFLOP / pair
WORKLOAD = "corrblock_microbench_55x128x256_L4_r4_build+12lookups"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


def ncu_traffic(kernel, batch, storage):
    """DRAM bytes per launch of `kernel` from the committed ncu capture (profiles/traffic.json),
    or None when the capture was taken on another configuration."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return None
    ent = json.load(open(p)).get(kernel)
    if not ent or ent.get("pairs") != batch or ent.get("storage") != storage:
        return None
    return ent["dram_bytes_per_launch"]


def write_only_gbs(device):
    """HBM bandwidth of a pure write stream, measured live with cudaMemsetAsync over 2 GiB (called through
    libcudart; torch's zero_ / fill_ kernels stop at 3.9 TB/s).  Round 1 quoted torch's fill kernel here (3.9 TB/s) as "the" write-only ceiling; tools/store_probe.cu
    (profiles/r2b_store_probe.jsonl, hand-written stores) shows that was the fill kernel's limit, not the memory's:
    memset 7.3 TB/s, bulk shared->global copies 6.4 TB/s, the build's own access pattern (256 contiguous bytes per
    query row per visit) 5.9 TB/s with plain 16-byte stores."""
    import ctypes
    buf = torch.empty(1 << 31, dtype=torch.uint8, device=device)
    try:
        rt = ctypes.CDLL("libcudart.so.12")
        rt.cudaMemsetAsync.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p]
        st = torch.cuda.current_stream().cuda_stream

        def fill():
            if rt.cudaMemsetAsync(buf.data_ptr(), 0, 1 << 31, st) != 0:
                raise RuntimeError("cudaMemsetAsync failed")
    except OSError:          # no libcudart on the loader path: torch's fill kernel (slower than the memory)
        def fill():
            buf.zero_()
    best = 1e9
    for i in range(5):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); fill(); e1.record(); torch.cuda.synchronize()
        if i:
            best = min(best, e0.elapsed_time(e1))
    return (1 << 31) / best / 1e6


# hand-written store probe (tools/store_probe.cu on this pool's B200, profiles/r2b_store_probe.jsonl): plain 16-byte
# stores in the build's access pattern, 256 contiguous bytes per query row per visit
PROBE_CHUNK256_GBS = 5914.5


def make_inputs(batch, device, seed):
    """Seeded synthetic features and 12 coordinate fields (grid + N(0, 5^2) flow)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    f1 = torch.randn(batch, D, H, W, generator=g)
    f2 = torch.randn(batch, D, H, W, generator=g)
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    grid = torch.stack([xs, ys], 0).float()[None]
    coords = [grid + 5.0 * torch.randn(batch, 2, H, W, generator=g) for _ in range(ITERS)]
    if device is not None:
        f1, f2 = f1.to(device), f2.to(device)
        coords = [c.to(device) for c in coords]
    return f1, f2, coords


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20", "-i", str(self.index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
            if len(parts) > 6:
                try:
                    pw.append(float(parts[6]))
                except ValueError:
                    pass
        # power_w: the board's draw over the same window (both hot kernels run at its power cap)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons),
                "power_w": {"median": statistics.median(pw), "max": max(pw)} if pw else None}


def cpu_hot_path():
    """The reference's CPU implementation of the path as a callable (f1, f2, coords_list) -> outputs.
    kind "reference": the reference's own core/corr.py + core/utils/utils.py, unmodified, staged under
    oracle/_ref/refcore/ by oracle/stage_ref_python.py (the checkout itself is absent on the GPU box).
    kind "port": oracle/torch_port.py (the same torch calls) when those files were not staged."""
    ref_dir = os.path.join(ROOT, "oracle", "_ref", "refcore")
    if os.path.exists(os.path.join(ref_dir, "corr.py")):
        if ref_dir not in sys.path:
            sys.path.insert(0, ref_dir)
        import warnings
        warnings.filterwarnings("ignore")          # torch.meshgrid indexing warning of the 2020 code
        from corr import CorrBlock as RefCorrBlock   # the reference, unmodified

        def run(f1, f2, coords):
            blk = RefCorrBlock(f1, f2, num_levels=LEVELS, radius=RADIUS)
            return [blk(c) for c in coords]
        return run, "reference", "reference core/corr.py CorrBlock (unmodified, oracle/_ref/refcore) on the host CPU"
    from oracle import torch_port as P

    def run(f1, f2, coords):
        return P.hot_path(f1, f2, coords, RADIUS, LEVELS)
    return run, "port", "the reference's torch CPU calls via oracle/torch_port.py"


def cpu_reference_pairs_per_s(budget_s, max_pairs):
    """The reference's CPU path timed on the host cores over a bounded sample of the workload."""
    run, kind, what = cpu_hot_path()
    torch.set_num_threads(os.cpu_count() or 1)
    f1, f2, coords = make_inputs(1, None, seed=0)
    with torch.no_grad():
        run(f1, f2, coords[:1])            # warm-up (thread pools, allocator)
        times = []
        t_start = time.perf_counter()
        while len(times) < max_pairs and (len(times) < 2 or time.perf_counter() - t_start < budget_s):
            t0 = time.perf_counter()
            run(f1, f2, coords)
            times.append(time.perf_counter() - t0)
    return len(times) / sum(times), len(times), torch.get_num_threads(), kind, what


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path on all host cores.  One step = ONE
    frame pair (build + 12 lookups): a bounded sample of the GPU arm's 64-pair step; both arms report pairs/s."""
    if rank != 0:
        return
    run, kind, what = cpu_hot_path()
    torch.set_num_threads(os.cpu_count() or 1)
    f1, f2, coords = make_inputs(1, None, seed=0)
    with torch.no_grad():
        for _ in range(max(args.warmup, 1)):
            run(f1, f2, coords)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            run(f1, f2, coords)
        dt = time.perf_counter() - t0
    v = args.steps / dt
    cores = torch.get_num_threads()
    line = {
        "impl": "reference", "metric": "frame_pairs_per_s_corr_hot_path_55x128x256_12iters", "value": v,
        "unit": "pairs/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": max(args.warmup, 1),
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "pairs_per_step": 1, "device": "host CPU",
                   "note": "per-pair throughput of the same workload; the GPU arm runs 64 pairs per step"},
        "cpu_baseline": {"value": v, "unit": "pairs/s", "cores": cores, "kind": kind,
                         "sample": f"1 pair per step (build + 12 lookups), {what}"},
        "e2e": {"value": v, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=64,
                    help="frame pairs per GPU per step (BASELINE configs[2]: batch 64 Sintel-shape pairs)")
    ap.add_argument("--total-batch", type=int, default=0,
                    help="strong scaling: this many pairs in total, split evenly over the ranks (overrides --batch; "
                         "SURVEY 8(d) config 3 asks for 64 total next to 64 per GPU)")
    ap.add_argument("--e2e-batch", type=int, default=0, help="frame pairs per end-to-end step (0 = the same as --batch)")
    ap.add_argument("--storage", default="bf16", choices=["bf16", "fp16", "fp32"])
    ap.add_argument("--math", default="bf16", choices=["bf16", "fp16", "tf32"], help="operand type of the tensor-core build")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-budget", type=float, default=12.0, help="seconds of CPU baseline work")
    ap.add_argument("--no-extras", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the correlation path has no CPU fallback")
    import raft_stir_b200 as R
    from raft_stir_b200 import _lib
    _lib.lib()

    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)

    if args.total_batch:
        if args.total_batch % world:
            raise SystemExit(f"--total-batch {args.total_batch} does not split over {world} ranks")
        args.batch = args.total_batch // world
    B = args.batch
    f1, f2, coords = make_inputs(B, device, seed=1000 + rank)
    stream = torch.cuda.current_stream()
    es = 4 if args.storage == "fp32" else 2

    def step(rec=None):
        if rec is not None:
            rec.append(torch.cuda.Event(enable_timing=True)); rec[-1].record(stream)
        blk = R.CorrBlock(f1, f2, LEVELS, RADIUS, storage=args.storage, math=args.math)
        if rec is not None:
            rec.append(torch.cuda.Event(enable_timing=True)); rec[-1].record(stream)
        out = None
        for c in coords:
            out = blk(c)
            if rec is not None:
                rec.append(torch.cuda.Event(enable_timing=True)); rec[-1].record(stream)
        return out

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.no_grad():
        sampler = ClockSampler(local_rank)
        sampler.start()
        for _ in range(args.warmup):
            step()
        # the timed region is tens of milliseconds: keep the same workload running until nvidia-smi delivers
        # its first sample, so that the clock record brackets the timed steps under load
        t_wait = time.perf_counter()
        while not sampler.lines and time.perf_counter() - t_wait < 3.0:
            step()
            torch.cuda.synchronize()
        n_before = len(sampler.lines)
        recs = []
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        for _ in range(args.steps):
            r = []
            step(r)
            recs.append(r)
        e1.record(stream)
        barrier()
        t_wait = time.perf_counter()
        while len(sampler.lines) < n_before + 3 and time.perf_counter() - t_wait < 1.0:
            step()                      # untimed: same load until a few samples past the timed region exist
            torch.cuda.synchronize()
        clocks = sampler.stop()
        clocks["window"] = "same workload from before to after the timed steps (20 ms sampling)"
    ms = e0.elapsed_time(e1)
    if dist is not None:
        tms = torch.tensor([ms], device=device)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        ms = float(tms.item())

    # per-kernel durations from the events recorded inside the timed region
    build_ms = [r[0].elapsed_time(r[1]) for r in recs]
    look_ms = [r[i].elapsed_time(r[i + 1]) for r in recs for i in range(1, 1 + ITERS)]
    build_avg, look_avg = statistics.mean(build_ms), statistics.mean(look_ms)

    pk = peaks()
    look_bytes = B * N * (LEVELS * (2 * RADIUS + 2) ** 2 * es + 8 + LEVELS * (2 * RADIUS + 1) ** 2 * 4)
    build_bytes = B * (2 * N * D * 4 + PYR_ELEMS * es)       # fp32 features in, pyramid out (repack traffic not counted)
    look_gbs = look_bytes / (look_avg * 1e-3) / 1e9
    build_gbs = build_bytes / (build_avg * 1e-3) / 1e9
    build_tf = B * BUILD_FLOPS / (build_avg * 1e-3) / 1e12
    # `roofline` describes the per-iteration kernel (12 of the 14 launches of a step, half of its time);
    # `roofline_build` the other half (pack + tcgen05 build, timed together as the API issues them)
    wr_gbs = write_only_gbs(device)
    roof = {"kernel": "lookup_fwd_kernel", "bound": "hbm", "achieved": look_gbs, "peak": pk["hbm_gbs"], "unit": "GB/s",
            "frac": look_gbs / pk["hbm_gbs"], "traffic": ncu_traffic("lookup_fwd_kernel", B, args.storage),
            "peak_source": pk["source"],
            "algorithmic_bytes_per_launch": look_bytes, "avg_launch_ms": look_avg}
    roof_build = {"kernel": "pack_rows16_kernel + build_pyramid_tc_kernel (CTA pairs)", "bound": "hbm", "achieved": build_gbs,
                  "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": build_gbs / pk["hbm_gbs"],
                  "traffic": ncu_traffic("build_pyramid_tc_kernel", B, args.storage), "peak_source": pk["source"],
                  "algorithmic_bytes_per_launch": build_bytes, "avg_launch_ms": build_avg,
                  "tensor": {"achieved": build_tf, "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s",
                             "frac": build_tf / pk["bf16_tflops_sustained"]},
                  "write_stream": {"memset_GB/s_live": wr_gbs, "frac_of_memset": build_gbs / wr_gbs,
                                   "probe_256B_pieces_GB/s": PROBE_CHUNK256_GBS,
                                   "frac_of_probe": build_gbs / PROBE_CHUNK256_GBS}}
    kernels = {
        "lookup_fwd": {"avg_ms": look_avg, "GB/s": look_gbs, "frac_hbm": look_gbs / pk["hbm_gbs"], "share_of_step": ITERS * look_avg / (ITERS * look_avg + build_avg)},
        "build(pack+tcgen05)": {"avg_ms": build_avg, "GB/s": build_gbs, "frac_hbm": build_gbs / pk["hbm_gbs"], "TFLOP/s": build_tf,
                                "frac_tensor": build_tf / pk["bf16_tflops_sustained"], "share_of_step": build_avg / (ITERS * look_avg + build_avg)},
    }

    # ---- the lookup with fp16 output (what --mixed_precision callers get), same pyramid, as a third roofline object
    with torch.no_grad():
        blk16 = R.CorrBlock(f1, f2, LEVELS, RADIUS, storage=args.storage, math=args.math, out_dtype=torch.float16)
        for c in coords[:3]:
            blk16(c)
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(ITERS + 1)]
        ev[0].record(stream)
        for i, c in enumerate(coords):
            blk16(c)
            ev[i + 1].record(stream)
        torch.cuda.synchronize()
        l16_ms = statistics.mean(ev[i].elapsed_time(ev[i + 1]) for i in range(ITERS))
        l16_bytes = B * N * (LEVELS * (2 * RADIUS + 2) ** 2 * es + 8 + LEVELS * (2 * RADIUS + 1) ** 2 * 2)
        roof_l16 = {"kernel": "lookup_fwd_kernel<fp16 out>", "bound": "hbm", "achieved": l16_bytes / l16_ms / 1e6,
                    "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": l16_bytes / l16_ms / 1e6 / pk["hbm_gbs"],
                    "traffic": ncu_traffic("lookup_fwd_kernel_f16out", B, args.storage), "peak_source": pk["source"],
                    "algorithmic_bytes_per_launch": l16_bytes, "avg_launch_ms": l16_ms}
        del blk16

    # ---- parity of the timed configuration itself: 1 000 random outputs of the last lookup of a 64-pair step against
    # the fp32 on-the-fly block (another kernel, fp32 FMA, no stored volume) on the same inputs, outside the timed region
    with torch.no_grad():
        out = step()
        ref = R.AlternateCorrBlock(f1, f2, LEVELS, RADIUS, math="fp32")(coords[-1])
        gsel = torch.Generator(device="cpu").manual_seed(5)
        idx = torch.randint(0, out.numel(), (1000,), generator=gsel).to(device)
        a, r_ = out.reshape(-1)[idx].double(), ref.reshape(-1)[idx].double()
        err = (a - r_).abs()
        floor = 3.0       # 3 rms(volume): N(0,1) features give a unit-variance volume (tests/test_gpu_parity.py: bf16_floor)
        parity = {"checked": 1000, "against": "AlternateCorrBlock(math='fp32') on the same inputs",
                  "max_rel_err_with_floor": float((err / (r_.abs() + floor)).max()),
                  "share_within_plain_1e-2_relative": float((err <= 1e-2 * r_.abs()).double().mean()),
                  "max_abs_err": float(err.max()), "tolerance": "1e-2 * (|ref| + 3 rms(volume)) for a bf16 volume"}
        del out, ref

    # ---- end to end through the public API from pinned host buffers: the SAME step as `value` (B pairs per GPU)
    e2e = None
    e2e16 = None
    with torch.no_grad():
        Be = B if args.e2e_batch <= 0 else min(B, args.e2e_batch)
        hf1, hf2 = f1[:Be].cpu().pin_memory(), f2[:Be].cpu().pin_memory()
        hcoords = [c[:Be].cpu().pin_memory() for c in coords]
        nch = LEVELS * (2 * RADIUS + 1) ** 2
        copy_stream = torch.cuda.Stream(device)

        def run_e2e(out_dtype):
            # Every lookup result goes to pinned host memory.  Two result buffers alternate, as in an application
            # that consumes result i while result i+1 is on its way (12 distinct buffers of 584 MB per rank would
            # pin 56 GB of host memory on an 8-GPU box); a buffer is reused only after its previous copy finished.
            houts = [torch.empty(Be, nch, H, W, dtype=out_dtype).pin_memory() for _ in range(2)]
            done = [None, None]

            def e2e_step():
                d1 = hf1.to(device, non_blocking=True)
                d2 = hf2.to(device, non_blocking=True)
                blk = R.CorrBlock(d1, d2, LEVELS, RADIUS, storage=args.storage, math=args.math, out_dtype=out_dtype)
                for i in range(ITERS):
                    c = hcoords[i].to(device, non_blocking=True)
                    o = blk(c)
                    ev = torch.cuda.Event(); ev.record(stream)
                    copy_stream.wait_event(ev)
                    with torch.cuda.stream(copy_stream):
                        houts[i & 1].copy_(o, non_blocking=True)     # stream order keeps the two buffers' copies apart
                    o.record_stream(copy_stream)

            # Steps are pipelined as an application streaming batches would: the next step's uploads and build run
            # while the previous step's last results are still on their way out; the clock stops only after the
            # last result has landed.
            e2e_step()
            stream.wait_stream(copy_stream)
            barrier()
            t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
            t0.record(stream)
            for _ in range(args.e2e_steps):
                e2e_step()
            stream.wait_stream(copy_stream)
            t1.record(stream)
            barrier()
            my_ms = t0.elapsed_time(t1)
            e2e_ms = my_ms
            if dist is not None:
                tms = torch.tensor([e2e_ms], device=device)
                dist.all_reduce(tms, op=dist.ReduceOp.MAX)
                e2e_ms = float(tms.item())
            d2h = int(ITERS * Be * nch * N * houts[0].element_size())
            del houts, done
            return {"value": world * Be * args.e2e_steps / (e2e_ms * 1e-3), "unit": "pairs/s",
                    "h2d_bytes_per_step": int(2 * Be * D * N * 4 + ITERS * Be * 2 * N * 4),
                    "d2h_bytes_per_step": d2h, "steps": args.e2e_steps, "pairs_per_step": Be,
                    "d2h_gbs_per_rank": d2h * args.e2e_steps / (my_ms * 1e-3) / 1e9,
                    "result_dtype": str(out_dtype).replace("torch.", ""),
                    "api": "raft_stir_b200.CorrBlock(fmap1, fmap2)(coords) x12, pinned host in / pinned host out"}

        e2e = run_e2e(torch.float32)
        # the same with the block answering in fp16 (out_dtype=torch.float16, what --mixed_precision consumes): half
        # the bytes on the PCIe link, which is what bounds this number (d2h_gbs_per_rank)
        e2e16 = run_e2e(torch.float16)
        e2e["fp16_results"] = {k: e2e16[k] for k in ("value", "d2h_bytes_per_step", "d2h_gbs_per_rank", "result_dtype")}
        e2e["note"] = ("bounded by the device->host link: every pair returns 12 x 324 x 7040 outputs (109 MB fp32); "
                       "d2h_gbs_per_rank is this rank's achieved result bandwidth")

    # ---- CPU baseline (rank 0, N = 1 only) and extras
    cpu = None
    extras = {}
    if rank == 0 and world == 1:
        v, npairs, cores, kind, what = cpu_reference_pairs_per_s(args.cpu_budget, max_pairs=40)
        cpu = {"value": v, "unit": "pairs/s", "cores": cores, "kind": kind,
               "sample": f"{npairs} pairs of {WORKLOAD}: {what}"}
        if not args.no_extras:
            extras = run_extras(R, device, args, pk)

    if rank == 0:
        line = {
            "metric": "frame_pairs_per_s_corr_hot_path_55x128x256_12iters",
            "value": world * B * args.steps / (ms * 1e-3), "unit": "pairs/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "strong" if args.total_batch else "weak", "vs_baseline": None,
            "dtype": {"bf16": "bf16", "fp16": "f16", "tf32": "tf32"}[args.math],
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "pairs_per_gpu_per_step": B, "volume_storage": args.storage,
                       "build_math": f"{args.math} tcgen05, fp32 accumulate", "l2": "working set per step (pyramid %.0f MB) exceeds the 126 MB L2" % (B * PYR_ELEMS * es / 1e6),
                       "parallelism": f"pairs sharded over {world} rank(s), no collective"},
            "roofline": roof, "roofline_build": roof_build, "roofline_lookup_fp16_out": roof_l16, "kernels": kernels,
            "parity_spot_check": parity, "cpu_baseline": cpu, "e2e": e2e,
            "gpu_launches": world * args.steps * (2 + ITERS), "clocks": clocks, "extras": extras,
        }
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def run_extras(R, device, args, pk):
    """Secondary measurements (not the headline): single-pair latency, training
    direction, on-the-fly path.  CUDA events, warm, reported as-is."""
    out = {}
    stream = torch.cuda.current_stream()

    def timeit(fn, reps):
        fn(); torch.cuda.synchronize()
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(reps):
            fn()
        b.record(stream); torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    try:
        with torch.no_grad():
            f1, f2, coords = make_inputs(1, device, seed=7)
            out["b1_build_ms"] = timeit(lambda: R.CorrBlock(f1, f2, LEVELS, RADIUS, storage=args.storage, math=args.math), 20)
            blk = R.CorrBlock(f1, f2, LEVELS, RADIUS, storage=args.storage, math=args.math)
            out["b1_lookup_ms"] = timeit(lambda: blk(coords[0]), 50)
            if args.storage != "fp32" and args.math in ("bf16", "fp16"):
                # the same build from 16-bit channels-last features (what a channels_last bf16 encoder emits):
                # no pack kernel, the 8x32 blocks are cut out of NHWC by the TMA box
                dt = torch.float16 if args.math == "fp16" else torch.bfloat16
                p1, p2 = R.pack_features(f1, f2, dt)
                out["b1_build_from_nhwc16_ms"] = timeit(lambda: R.CorrBlock(p1, p2, LEVELS, RADIUS, storage=args.storage), 20)
                g64 = make_inputs(args.batch, device, seed=8)
                q1, q2 = R.pack_features(g64[0], g64[1], dt)
                out["b%d_build_from_nhwc16_ms" % args.batch] = timeit(lambda: R.CorrBlock(q1, q2, LEVELS, RADIUS, storage=args.storage), 5)
                out["b%d_build_with_pack_ms" % args.batch] = timeit(
                    lambda: R.CorrBlock(g64[0], g64[1], LEVELS, RADIUS, storage=args.storage, math=args.math), 5)
                del g64, q1, q2
            out["b1_build_fp32_simt_ms"] = timeit(lambda: R.CorrBlock(f1, f2, LEVELS, RADIUS, storage="fp32", math="fp32"), 3)
            out["b1_build_tf32_fp32_storage_ms"] = timeit(lambda: R.CorrBlock(f1, f2, LEVELS, RADIUS, storage="fp32", math="tf32"), 10)
            out["b1_build_bf16_fp32_storage_ms"] = timeit(lambda: R.CorrBlock(f1, f2, LEVELS, RADIUS, storage="fp32", math="bf16"), 10)
            alt = R.AlternateCorrBlock(f1, f2, LEVELS, RADIUS, math="fp32")
            out["b1_alt_lookup_ms"] = timeit(lambda: alt(coords[0]), 10)
            alt_tc = R.AlternateCorrBlock(f1, f2, LEVELS, RADIUS, math="bf16")
            out["b1_alt_tc_lookup_ms"] = timeit(lambda: alt_tc(coords[0]), 30)
        # training direction at the FlyingChairs shape (config 5): 46x62, batch 10
        g = torch.Generator(device="cpu").manual_seed(3)
        Bc, Hc, Wc = 10, 46, 62
        a1 = torch.randn(Bc, D, Hc, Wc, generator=g).to(device).requires_grad_(True)
        a2 = torch.randn(Bc, D, Hc, Wc, generator=g).to(device).requires_grad_(True)
        cc = [R.coords_grid(Bc, Hc, Wc, device) + 3 * torch.randn(Bc, 2, Hc, Wc, device=device) for _ in range(ITERS)]

        def train_step():
            a1.grad = None; a2.grad = None
            blk = R.CorrBlock(a1, a2, LEVELS, RADIUS, storage=args.storage, math=args.math)
            loss = sum(blk(c).sum() for c in cc)
            loss.backward()
        out["chairs_b10_fwd_bwd_ms"] = timeit(train_step, 5)
        # the two backward kernels alone, same shape: 12 lookup backwards into one gradient pyramid, then the build backward
        with torch.no_grad():
            blk = R.CorrBlock(a1.detach(), a2.detach(), LEVELS, RADIUS, storage=args.storage, math=args.math)
            go = torch.randn(Bc, LEVELS * (2 * RADIUS + 1) ** 2, Hc, Wc, device=device)
            blk._lookup_backward(cc[0], go, True, False)
            ms = timeit(lambda: blk._lookup_backward(cc[0], go, True, False), 20)
            nq = Bc * Hc * Wc
            bwd_bytes = nq * (LEVELS * (2 * RADIUS + 1) ** 2 * 4 + 8 + LEVELS * (2 * RADIUS + 2) ** 2 * 4 * 2)
            out["chairs_b10_lookup_bwd_ms"] = ms
            out["chairs_b10_lookup_bwd_GB/s"] = bwd_bytes / (ms * 1e-3) / 1e9
            out["chairs_b10_lookup_bwd_frac_hbm"] = out["chairs_b10_lookup_bwd_GB/s"] / pk["hbm_gbs"]
        # STIR-scale video (config 4): one 1080x1920 pair, fmap 135x240
        with torch.no_grad():
            g = torch.Generator(device="cpu").manual_seed(11)
            Hv, Wv = 135, 240
            v1 = torch.randn(1, D, Hv, Wv, generator=g).to(device)
            v2 = torch.randn(1, D, Hv, Wv, generator=g).to(device)
            cv = R.coords_grid(1, Hv, Wv, device) + 5 * torch.randn(1, 2, Hv, Wv, device=device)
            out["v1080_build_ms"] = timeit(lambda: R.CorrBlock(v1, v2, LEVELS, RADIUS, storage=args.storage, math=args.math), 3)
            blk = R.CorrBlock(v1, v2, LEVELS, RADIUS, storage=args.storage, math=args.math)
            out["v1080_lookup_ms"] = timeit(lambda: blk(cv), 20)
            del blk
            altv = R.AlternateCorrBlock(v1, v2, LEVELS, RADIUS, math="fp32")
            out["v1080_alt_lookup_ms"] = timeit(lambda: altv(cv), 5)
            altv_tc = R.AlternateCorrBlock(v1, v2, LEVELS, RADIUS, math="bf16")
            out["v1080_alt_tc_lookup_ms"] = timeit(lambda: altv_tc(cv), 20)
            out["v1080_note"] = "12 iterations: default = build + 12 lookups; on-the-fly = 12 alt lookups (no volume in memory)"
        # the headline's lookup with fp16 output (what an update block under fp16 autocast consumes), same batch
        with torch.no_grad():
            f1b = torch.randn(args.batch, D, H, W, device=device)
            f2b = torch.randn(args.batch, D, H, W, device=device)
            cb = R.coords_grid(args.batch, H, W, device) + 5 * torch.randn(args.batch, 2, H, W, device=device)
            blk16 = R.CorrBlock(f1b, f2b, LEVELS, RADIUS, storage=args.storage, math=args.math, out_dtype=torch.float16)
            del f1b, f2b
            ms = timeit(lambda: blk16(cb), 20)
            nq = args.batch * H * W
            b16 = nq * (LEVELS * (2 * RADIUS + 2) ** 2 * (4 if args.storage == "fp32" else 2) + 8
                        + LEVELS * (2 * RADIUS + 1) ** 2 * 2)
            out["lookup_fp16_out_ms"] = ms
            out["lookup_fp16_out_GB/s"] = b16 / (ms * 1e-3) / 1e9
            out["lookup_fp16_out_frac_hbm"] = out["lookup_fp16_out_GB/s"] / pk["hbm_gbs"]
            del blk16, cb
        # the step after the lookup + update block: convex 8x upsampling of the batch's flow
        with torch.no_grad():
            g = torch.Generator(device="cpu").manual_seed(13)
            Bu = args.batch
            fl = (5 * torch.randn(Bu, 2, H, W, generator=g)).to(device)
            mk = (2 * torch.randn(Bu, 576, H, W, generator=g)).to(device)
            ms = timeit(lambda: R.upsample_flow(fl, mk), 20)
            up_bytes = Bu * H * W * (576 * 4 + 2 * 4 + 128 * 4)
            out["upsample_flow_ms"] = ms
            out["upsample_flow_GB/s"] = up_bytes / (ms * 1e-3) / 1e9
            out["upsample_flow_frac_hbm"] = out["upsample_flow_GB/s"] / pk["hbm_gbs"]
            out["upsample_flow_note"] = f"{Bu} maps of {H}x{W}: 2 304 B mask + 8 B flow in, 512 B out per coarse pixel"
            del fl, mk
        # lookup -> convc1 -> ReLU (core/update.py:88-90), fused against composed, same batch
        with torch.no_grad():
            import torch.nn.functional as F
            f1b = torch.randn(args.batch, D, H, W, device=device)
            f2b = torch.randn(args.batch, D, H, W, device=device)
            cb = R.coords_grid(args.batch, H, W, device) + 5 * torch.randn(args.batch, 2, H, W, device=device)
            blk = R.CorrBlock(f1b, f2b, LEVELS, RADIUS, storage=args.storage, math=args.math)
            del f1b, f2b
            nch = LEVELS * (2 * RADIUS + 1) ** 2
            wc = torch.randn(256, nch, 1, 1, device=device) / nch ** 0.5
            bc = 0.1 * torch.randn(256, device=device)
            wp = R.pack_convc1_weights(wc)
            out["lookup_convc1_fused_ms"] = timeit(lambda: blk.lookup_convc1(cb, wc, bc, packed_weight=wp), 20)
            out["lookup_then_cudnn_convc1_relu_ms"] = timeit(lambda: F.relu(F.conv2d(blk(cb), wc, bc)), 10)
            prev = torch.backends.cudnn.allow_tf32
            torch.backends.cudnn.allow_tf32 = False
            out["lookup_then_cudnn_convc1_relu_fp32_ms"] = timeit(lambda: F.relu(F.conv2d(blk(cb), wc, bc)), 5)
            torch.backends.cudnn.allow_tf32 = prev
            out["lookup_convc1_note"] = (f"{args.batch} pairs; fused = rc_lookup_convc1_fwd (fp16 samples, tcgen05 1x1 conv + bias + "
                                         "ReLU, fp32 out); composed = this library's fp32-out lookup + torch conv2d + relu "
                                         "(cuDNN with its default TF32 setting, and with TF32 off)")
            del blk, cb
    except Exception as e:  # extras must never sink the headline
        out["error"] = f"{type(e).__name__}: {e}"
    try:
        out.update(run_network_extras(R, device))
    except Exception as e:
        out["network_error"] = f"{type(e).__name__}: {str(e)[:300]}"
    return out


def run_network_extras(R, device):
    """BASELINE configs 1 and 5 under the driver (part of the untimed CPU-baseline leg; rank 0, N = 1): the whole
    RAFT-basic network with the reference's own sub-networks (traced to oracle/_ref/*.pt by oracle/make_ref_modules.py,
    random-init weights, fp32) around (a) the reference CorrBlock's torch ops on this GPU and (b) this library.
    Skipped when the traced networks are not on the machine."""
    from oracle import raft_harness as Hn      # checker-side harness: control flow of RAFT.forward around the nets
    if not Hn.available():
        return {"network": "oracle/_ref/*.pt not present: configs 1 and 5 not run"}
    out = {}
    nets = Hn.load(device)

    def timeit(fn, warm, reps):
        for _ in range(warm):
            r = fn()
        torch.cuda.synchronize()
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            r = fn()
        b.record(); torch.cuda.synchronize()
        return a.elapsed_time(b) / reps, r

    # config 1: one 440x1024 pair, 12 iterations, test mode
    g = torch.Generator(device="cpu").manual_seed(0)
    im1 = (torch.rand(1, 3, 440, 1024, generator=g) * 255).to(device)
    im2 = (torch.rand(1, 3, 440, 1024, generator=g) * 255).to(device)
    with torch.no_grad():
        timeit(lambda: Hn.raft_forward(nets, im1, im2, 12, R.CorrBlock), 2, 1)        # process warm-up
        ms_ref, (_, up_ref) = timeit(lambda: Hn.raft_forward(nets, im1, im2, 12, Hn.TorchCorrBlock), 3, 5)
        ms_our, (_, up) = timeit(lambda: Hn.raft_forward(nets, im1, im2, 12, R.CorrBlock), 3, 5)
        ms_alt, (_, up_alt) = timeit(lambda: Hn.raft_forward(nets, im1, im2, 12, R.AlternateCorrBlock), 3, 5)
        try:      # the whole forward as one CUDA graph (raft_stir_b200.GraphedCallable)
            graphed = R.GraphedCallable(lambda a, b: Hn.raft_forward(nets, a, b, 12, R.CorrBlock), (im1, im2))
            ms_graph, (_, up_g) = timeit(graphed.replay, 2, 10)
            graph_epe = torch.sqrt(((up_g - up_ref) ** 2).sum(1)).mean().item()
            del graphed
        except Exception as e:
            ms_graph, graph_epe = None, f"{type(e).__name__}: {str(e)[:200]}"
        out["config1_raft_440x1024_12it"] = {
            "torch_corrblock_ms": ms_ref, "this_library_bf16_ms": ms_our, "this_library_on_the_fly_ms": ms_alt,
            "this_library_bf16_one_cuda_graph_ms": ms_graph, "mean_epe_delta_px_cuda_graph": graph_epe,
            "pairs_per_s": {"torch_corrblock": 1e3 / ms_ref, "this_library_bf16": 1e3 / ms_our},
            "mean_epe_delta_px": torch.sqrt(((up - up_ref) ** 2).sum(1)).mean().item(),
            "mean_epe_delta_px_on_the_fly": torch.sqrt(((up_alt - up_ref) ** 2).sum(1)).mean().item(),
            "mean_flow_px": up_ref.norm(dim=1).mean().item(),
            "note": "reference sub-networks (cuDNN, fp32) around the block; budget 0.01 px"}
    del im1, im2

    # config 5 on one GPU: training step, 10 pairs of 368x496, 12 iterations, sequence loss, AdamW, clip 1.0
    import torch.nn as nn

    class Net(nn.Module):
        def __init__(self, nets, corr_cls, upsample_fn=None):
            super().__init__()
            self.fnet, self.cnet, self.update = nets
            self.corr_cls, self.upsample_fn = corr_cls, upsample_fn

        def forward(self, a, b):
            return Hn.raft_forward_train((self.fnet, self.cnet, self.update), a, b, 12, self.corr_cls,
                                         upsample_fn=self.upsample_fn)

    Bt, Ht, Wt = 10, 368, 496
    g = torch.Generator(device="cpu").manual_seed(100)
    im1 = (torch.rand(Bt, 3, Ht, Wt, generator=g) * 255).to(device)
    im2 = (torch.rand(Bt, 3, Ht, Wt, generator=g) * 255).to(device)
    gt = (5 * torch.randn(Bt, 2, Ht, Wt, generator=g)).to(device)
    valid = torch.ones(Bt, Ht, Wt, device=device)
    res = {}
    for name, cls, ups in (("torch_corrblock", Hn.TorchCorrBlock, None),
                           ("this_library_block_and_upsampling", R.CorrBlock, R.upsample_flow)):
        model = Net(Hn.load(device), cls, ups).to(device)
        opt = torch.optim.AdamW(model.parameters(), lr=1e-5, weight_decay=1e-4)

        def step():
            opt.zero_grad(set_to_none=True)
            loss = Hn.sequence_loss(model(im1, im2), gt, valid)
            loss.backward()
            torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
            opt.step()
            return loss
        torch.cuda.empty_cache()
        ms, loss = timeit(step, 3, 3)
        res[name] = {"ms_per_step": ms, "pairs_per_s": 1e3 * Bt / ms, "loss": float(loss.item())}
        del model, opt
        torch.cuda.empty_cache()
    res["note"] = "one GPU, no DDP (tests/perf/train_step_bench.py runs the 8-GPU DistributedDataParallel form)"
    out["config5_train_step_chairs_b10"] = res
    return out


if __name__ == "__main__":
    main()
