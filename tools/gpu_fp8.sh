#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fp8.py -x -q > gpurun_out/pytest_fp8.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_fp8.log
tail -25 gpurun_out/pytest_fp8.log
for mode in 2 1 0; do
DGA_FP8_MODE=$mode timeout 300 python bench.py --workload fp8 --steps 100 --no-cpu-baseline > gpurun_out/bench_fp8_mode$mode.json 2> gpurun_out/bench_fp8_mode$mode.err
tail -3 gpurun_out/bench_fp8_mode$mode.err
done
python tools/show_bench.py gpurun_out/bench_fp8_mode*.json
