#!/usr/bin/env python3
"""Per-source-line stall summary of one kernel from an .ncu-rep (needs -lineinfo builds).

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep <kernel regex> [top N]
"""
import csv
import subprocess
import sys


def main():
    rep, pat = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass",
                          "--kernel-name", f"regex:{pat}"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = None
    lines = {}
    cur = None
    files = '?'
    total = 0
    seen_files = set()
    for r in rows:
        if r and r[0] == "File Path":
            if r[1] in seen_files:
                break       # first launch only (a kernel's sections repeat per launch, one per source file)
            seen_files.add(r[1])
            files = r[1].split("/")[-1]
        if r and r[0] == "Line No":
            hdr = r
            i_s = hdr.index("# Samples")
            stall_cols = [(i, h[6:]) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
            continue
        if hdr is None or len(r) < len(hdr):
            continue
        if r[0] == "":
            continue        # SASS rows; the CUDA-line rows above them carry the per-line sums
        cur = (files, r[0], r[1].strip())
        n = int(r[i_s]) if r[i_s].isdigit() else 0
        total += n
        ent = lines.setdefault(cur, [0, {}])
        ent[0] += n
        for i, name in stall_cols:
            v = int(r[i]) if r[i].isdigit() else 0
            if v:
                ent[1][name] = ent[1].get(name, 0) + v
    print(f"total samples {total}")
    for (fn, ln, src), (n, st) in sorted(lines.items(), key=lambda kv: -kv[1][0])[:top]:
        tops = ", ".join(f"{k}={v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print(f"{n:7d} {100.0 * n / max(total, 1):5.1f}%  {fn}:{ln:<4}  {src[:64]:64s} {tops}")


if __name__ == "__main__":
    main()
