#!/bin/bash
# A/B of the build kernel: CTA pairs (default) against the single-CTA kernel (RC_BUILD_PAIR=0).
# Output: gpurun_out/pair_ab.log.  PAIR_AB_TESTS=1 also runs the build's GPU tests first.
mkdir -p gpurun_out
LOG=gpurun_out/pair_ab.log
: > $LOG
if [ -n "$PAIR_AB_TESTS" ]; then
  echo "== pytest (pairs)" >> $LOG
  timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_packed.py tests/test_gpu_golden_shapes.py tests/test_zz_no_alloc.py -m gpu -q --no-header -x >> $LOG 2>&1; echo "pytest exit $?" >> $LOG
fi
echo "== build bench, pairs" >> $LOG
timeout 300 python tools/build_bench.py >> $LOG 2>&1; echo "exit $?" >> $LOG
echo "== build bench, single CTA" >> $LOG
RC_BUILD_PAIR=0 timeout 300 python tools/build_bench.py >> $LOG 2>&1; echo "exit $?" >> $LOG
tail -c 5000 $LOG
