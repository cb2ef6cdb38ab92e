#!/bin/bash
# DRAM bytes of one 64-pair build (ncu, two metrics only) and its time without the profiler.
mkdir -p gpurun_out
BATCH=64 timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:build_pyramid_tc -s 1 -c 1 python tools/profile_build.py 2>&1 | grep -E "dram__|gpu__time|build_pyramid" > gpurun_out/dram_build.log
timeout 200 python tools/build_attr.py >> gpurun_out/dram_build.log 2>&1
cat gpurun_out/dram_build.log
