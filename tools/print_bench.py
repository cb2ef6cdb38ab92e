#!/usr/bin/env python3
"""Print the headline fields of a bench.py JSON line (helper for the GPU-box scripts)."""
import json
import sys

d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(d["n_gpus"], "GPUs:", round(d["value"]), "pairs/s,", round(d["ms_per_step"], 3), "ms/step, e2e",
      round(d["e2e"]["value"]), "pairs/s, lookup frac", round(d["roofline"]["frac"], 4), d["clocks"])
