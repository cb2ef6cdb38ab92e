#!/bin/bash
# ncu passes over a short bench run (run tools/gpu_check.sh first; this script assumes the tests pass).
#   TAG=r1c bash tools/gpu_profile.sh
# Output (gpurun_out/): ${TAG}_launches.csv (every launch, gpu__time_duration), ${TAG}.ncu-rep (--set full of the hot kernels)
TAG=${TAG:-prof}
mkdir -p gpurun_out
LOG=gpurun_out/${TAG}_profile.log
CMD="python bench.py --steps 2 --warmup 3 --e2e-steps 1 --cpu-budget 0.1 --no-extras ${BENCH_ARGS}"
echo "== plain" > $LOG
timeout 300 $CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_list.log 2>&1
echo "ncu list exit $?" >> $LOG
timeout 300 $CMD > gpurun_out/${TAG}_plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"${KERNELS:-lookup_fwd|build_pyramid_tc|pack_features}" -s ${SKIP:-12} -c ${COUNT:-4} -o gpurun_out/${TAG} $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu full exit $?" >> $LOG
cat gpurun_out/${TAG}_plain.log >> $LOG
ls -la gpurun_out >> $LOG
tail -c 3000 $LOG
