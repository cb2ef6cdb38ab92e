#!/bin/bash
# ncu --set full of the build kernel (16 Sintel pairs from NHWC bf16 features): CTA pairs (default) and the single-CTA kernel.
mkdir -p gpurun_out
timeout 120 python tools/profile_build.py > gpurun_out/pb_plain.log 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:build_pyramid_tc -s 1 -c 1 -f -o gpurun_out/${TAG:-r2n}_build_pair python tools/profile_build.py > gpurun_out/pb_ncu1.log 2>&1
RC_BUILD_PAIR=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:build_pyramid_tc -s 1 -c 1 -f -o gpurun_out/${TAG:-r2n}_build_single python tools/profile_build.py > gpurun_out/pb_ncu2.log 2>&1
ls -la gpurun_out | tail -5
