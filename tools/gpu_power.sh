#!/bin/bash
# Power / energy attribution of the build kernel (tools/power_probe.py).  Output: gpurun_out/power.jsonl
mkdir -p gpurun_out; OUT=gpurun_out/power.jsonl; : > $OUT
A=raft_stir_b200/_build/attrib/libraftcorr.so
timeout 200 python tools/power_probe.py >> $OUT 2>gpurun_out/power.err
RC_BUILD_PAIR=0 timeout 100 python tools/power_probe.py build_64_pairs_nhwc16 >> $OUT 2>>gpurun_out/power.err
for d in 8 1; do RAFTCORR_LIB=$A RC_TC_DEBUG=$d timeout 100 python tools/power_probe.py build_64_pairs_nhwc16 >> $OUT 2>>gpurun_out/power.err; done
cat $OUT; tail -3 gpurun_out/power.err
