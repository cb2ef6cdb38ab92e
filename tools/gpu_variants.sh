#!/bin/bash
# bench lines for a list of environment variants: VARIANTS="A=1;B=2 C=3;..." (space-separated, ';' joins several vars)
mkdir -p gpurun_out
OUT=gpurun_out/variants.log
: > $OUT
for v in "" ${VARIANTS}; do
  envs=$(echo "$v" | tr ';' ' ')
  echo "== [$envs]" >> $OUT
  env $envs timeout 300 python bench.py --steps 5 --warmup 3 --e2e-steps 1 --cpu-budget 0.1 --no-extras ${BENCH_ARGS} 2>&1 | python -c "
import sys, json
for line in sys.stdin:
    line=line.strip()
    if line.startswith('{'):
        d=json.loads(line); k=d['kernels']
        print('pairs/s %.0f  step %.3f ms  lookup %.1f us  build %.1f us' % (d['value'], d['ms_per_step'], 1e3*k['lookup_fwd']['avg_ms'], 1e3*k['build(pack+tcgen05)']['avg_ms']))
    elif line: print(line[:300])
" >> $OUT
done
cat $OUT
