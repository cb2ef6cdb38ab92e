mkdir -p gpurun_out; LOG=gpurun_out/attr.log; : > $LOG
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_packed.py tests/test_gpu_golden_shapes.py tests/test_gpu_build_single_cta.py tests/test_zz_no_alloc.py -m gpu -q --no-header -x >> $LOG 2>&1; echo "pytest exit $?" >> $LOG
for pair in 1; do for d in ${DBGS:-0 1 8}; do RC_BUILD_PAIR=$pair RC_TC_DEBUG=$d timeout 120 python tools/build_attr.py >> $LOG 2>&1; done; done
RC_BUILD_PAIR=0 timeout 120 python tools/build_attr.py >> $LOG 2>&1
tail -c 3000 $LOG
