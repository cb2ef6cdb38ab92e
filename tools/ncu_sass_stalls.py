#!/usr/bin/env python3
"""Warp-stall samples of one kernel from an .ncu-rep, summed per stall reason and per SASS opcode, plus the
instructions with the most samples (with the instruction BEFORE them: a scoreboard stall is charged to the consumer).

    python tools/ncu_sass_stalls.py gpurun_out/x.ncu-rep <kernel regex> [top N]
"""
import csv
import subprocess
import sys
from collections import Counter


def main():
    rep, pat = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass",
                          "--kernel-name", f"regex:{pat}"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[1]
    i_src, i_n, i_ex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    stall = [(i, h[6:]) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    by_reason, by_op, execd = Counter(), Counter(), Counter()
    ins = []
    for r in rows[2:]:
        if len(r) < len(hdr) or not r[i_n].isdigit():
            continue
        src = " ".join(r[i_src].split())
        toks = src.split()
        op = toks[1] if toks and toks[0].startswith("@") and len(toks) > 1 else (toks[0] if toks else "?")
        op = op.split(".")[0]
        n = int(r[i_n])
        st = {name: int(r[i]) for i, name in stall if r[i].isdigit() and int(r[i])}
        for k, v in st.items():
            by_reason[k] += v
        by_op[op] += n
        execd[op] += int(r[i_ex]) if r[i_ex].isdigit() else 0
        ins.append((n, src, st))
    tot = sum(by_reason.values())
    print(f"samples {tot}")
    print("by reason: " + ", ".join(f"{k}={v} ({100.0 * v / tot:.1f}%)" for k, v in by_reason.most_common()))
    print("by opcode (samples | warp-instructions executed):")
    for k, v in by_op.most_common(25):
        print(f"  {k:14s} {v:7d} {100.0 * v / tot:5.1f}%   {execd[k]}")
    print("instructions:")
    order = sorted(range(len(ins)), key=lambda i: -ins[i][0])[:top]
    for i in order:
        n, src, st = ins[i]
        prev = ins[i - 1][1] if i else ""
        tops = ", ".join(f"{k}={v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print(f"  {n:6d} {100.0 * n / tot:5.1f}%  {src[:70]:70s} <- {prev[:50]:50s} | {tops}")


if __name__ == "__main__":
    main()
