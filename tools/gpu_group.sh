#!/bin/bash
# attribution build: pairs per L2 group of the build kernel (RC_TC_GROUP) against DRAM bytes and time (NHWC input),
# then pack + build from fp32 features (tools/build_bench.py) for a few group sizes
mkdir -p gpurun_out; LOG=gpurun_out/group.log; : > $LOG
export RAFTCORR_LIB=raft_stir_b200/_build/attrib/libraftcorr.so
for g in ${GROUPS_:-4 5 8 11 16 4 5 8}; do
  echo "== group $g" >> $LOG
  RC_TC_GROUP=$g BATCH=64 timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:build_pyramid_tc -s 1 -c 1 python tools/profile_build.py 2>&1 | grep -E "dram__|gpu__time" >> $LOG
  RC_TC_GROUP=$g timeout 200 python tools/build_attr.py >> $LOG 2>&1
done
for g in 3 5 8; do echo "== packed path, group $g" >> $LOG; RC_TC_GROUP=$g timeout 200 python tools/build_bench.py | grep '"bf16", "math": "bf16"' >> $LOG; done
cat $LOG
