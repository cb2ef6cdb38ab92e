#!/bin/bash
# Build kernel: GPU tests of the build, then timing (tools/build_attr.py) of the shipped library, of the attribution
# build's switches (DBGS) and of the single-CTA kernel.  Output: gpurun_out/attr.log
mkdir -p gpurun_out; LOG=gpurun_out/attr.log; : > $LOG
if [ -z "$NO_TESTS" ]; then
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_packed.py tests/test_gpu_golden_shapes.py tests/test_gpu_build_single_cta.py tests/test_zz_no_alloc.py -m gpu -q --no-header -x >> $LOG 2>&1; echo "pytest exit $?" >> $LOG
fi
timeout 120 python tools/build_attr.py >> $LOG 2>&1
for d in ${DBGS:-1 8}; do RAFTCORR_LIB=raft_stir_b200/_build/attrib/libraftcorr.so RC_TC_DEBUG=$d timeout 120 python tools/build_attr.py >> $LOG 2>&1; done
RC_BUILD_PAIR=0 timeout 120 python tools/build_attr.py >> $LOG 2>&1
timeout 120 python tools/build_attr.py >> $LOG 2>&1
tail -c 3000 $LOG
