#!/bin/bash
# forward lookup: GPU tests that touch it, then 64-pair timing (bench.py without extras)
mkdir -p gpurun_out; LOG=gpurun_out/lookup_ab.log; : > $LOG
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden_shapes.py tests/test_gpu_convc1.py tests/test_gpu_raft_e2e.py -m gpu -q --no-header -x >> $LOG 2>&1; echo "pytest exit $?" >> $LOG
for i in 1 2; do timeout 300 python bench.py --no-extras --cpu-budget 1 --e2e-steps 1 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), {k:(round(v['avg_ms'],4), round(v['frac_hbm'],4)) for k,v in d['kernels'].items()}, d['roofline_lookup_fp16_out']['avg_launch_ms'])" >> $LOG; done
tail -8 $LOG
