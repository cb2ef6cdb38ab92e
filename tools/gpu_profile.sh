#!/bin/bash
# pytest + smoke + bench, then ncu: launch list and full captures of the hot kernels.
mkdir -p gpurun_out
LOG=gpurun_out/profile.log
echo "== pytest" > $LOG
timeout 900 python -m pytest tests -m gpu -x -q --no-header >> $LOG 2>&1; echo "pytest exit $?" >> $LOG
timeout 300 python __graft_entry__.py smoke >> $LOG 2>&1; echo "smoke exit $?" >> $LOG
echo "== bench" >> $LOG
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?" >> $LOG
cat gpurun_out/bench.json >> $LOG
CMD="python bench.py --steps 2 --warmup 3 --e2e-steps 1 --cpu-budget 0.1 --no-extras"
echo "== ncu" >> $LOG
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list exit $?" >> $LOG
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"lookup_fwd|build_pyramid_tc|pack_features" -s 12 -c 6 -o gpurun_out/prof_r1 $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?" >> $LOG
ls -la gpurun_out >> $LOG
