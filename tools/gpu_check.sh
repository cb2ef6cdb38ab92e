#!/bin/bash
# pytest -m gpu, smoke, one bench line (no profiler).  Output: gpurun_out/check.log, gpurun_out/bench.json
mkdir -p gpurun_out
LOG=gpurun_out/check.log
echo "== pytest" > $LOG
timeout 1200 python -m pytest tests -m gpu -q --no-header -x ${PYTEST_ARGS} >> $LOG 2>&1; echo "pytest exit $?" >> $LOG
timeout 300 python __graft_entry__.py smoke >> $LOG 2>&1; echo "smoke exit $?" >> $LOG
echo "== bench" >> $LOG
timeout 600 python bench.py --steps 10 --warmup 3 ${BENCH_ARGS} > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?" >> $LOG
cat gpurun_out/bench.json >> $LOG
tail -5 gpurun_out/bench.err >> $LOG
tail -c 6000 $LOG
