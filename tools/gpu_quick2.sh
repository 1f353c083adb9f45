#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
for w in config2 m512 m2048 m8192 fp8 moe; do
  timeout 300 python bench.py --workload $w --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
done
python tools/show_bench.py gpurun_out/bench_config2.json gpurun_out/bench_m512.json gpurun_out/bench_m2048.json gpurun_out/bench_m8192.json gpurun_out/bench_fp8.json gpurun_out/bench_moe.json
DGA_DEBUG_FLAGS=16 python tools/trace_common.py 4096 7168 2048 2>&1 | tail -2
