bash tools/gpu_power.sh > /dev/null 2>&1
timeout 600 python bench.py --no-extras --cpu-budget 2 > gpurun_out/bench_pair.json 2> gpurun_out/bench_pair.err
RC_BUILD_PAIR=0 timeout 600 python bench.py --no-extras --cpu-budget 2 > gpurun_out/bench_single.json 2>> gpurun_out/bench_pair.err
timeout 300 python tools/build_bench.py > gpurun_out/build_bench_pair.jsonl 2>&1
RC_BUILD_PAIR=0 timeout 300 python tools/build_bench.py > gpurun_out/build_bench_single.jsonl 2>&1
cat gpurun_out/power.jsonl | cut -c1-400; cat gpurun_out/build_bench_pair.jsonl gpurun_out/build_bench_single.jsonl
