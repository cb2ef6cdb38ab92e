#!/bin/bash
# First GPU visit: isolate groups of tests in separate processes so one faulting
# kernel does not take the rest down.  Everything is bounded by `timeout`.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
echo "== smoke" > gpurun_out/first.log
timeout 300 python __graft_entry__.py smoke >> gpurun_out/first.log 2>&1; echo "smoke exit $?" >> gpurun_out/first.log
echo "== non tensor-core tests" >> gpurun_out/first.log
timeout 900 python -m pytest tests -m gpu -q --no-header -k "not tensor_core and not bf16 and not sintel and not full_size" >> gpurun_out/first.log 2>&1
echo "== tc_debug" >> gpurun_out/first.log
timeout 300 python tools/tc_debug.py >> gpurun_out/first.log 2>&1; echo "tc_debug exit $?" >> gpurun_out/first.log
echo "== tensor-core tests" >> gpurun_out/first.log
timeout 900 python -m pytest tests -m gpu -q --no-header -k "tensor_core or bf16 or sintel or full_size" >> gpurun_out/first.log 2>&1
echo "== bench" >> gpurun_out/first.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?" >> gpurun_out/first.log
tail -c 3000 gpurun_out/bench.err >> gpurun_out/first.log
cat gpurun_out/bench.json >> gpurun_out/first.log
