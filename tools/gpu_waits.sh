#!/bin/bash
# attribution build: cycles each role of the build kernel spends in its waits (RC_TC_DEBUG=32), 64 pairs from NHWC features
mkdir -p gpurun_out
RAFTCORR_LIB=raft_stir_b200/_build/attrib/libraftcorr.so RC_TC_DEBUG=32 BATCH=64 timeout 200 python tools/profile_build.py > gpurun_out/waits.log 2>&1
sort gpurun_out/waits.log | uniq | tail -60
