#!/bin/bash
# quick visit: parity + a few bench lines
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for w in ${WORKLOADS:-config2 m2048 m8192}; do
  timeout 300 python bench.py --workload $w --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
done
tail -4 gpurun_out/pytest_gpu.log
python tools/show_bench.py gpurun_out/bench_*.json 2>&1 | grep -v "g2\|g8\|mode"
