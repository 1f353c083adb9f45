#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fp8.py -x -q > gpurun_out/pytest_fp8.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_fp8.log
tail -25 gpurun_out/pytest_fp8.log
timeout 300 python bench.py --workload fp8 --steps 100 --no-cpu-baseline > gpurun_out/bench_fp8.json 2> gpurun_out/bench_fp8.err
tail -3 gpurun_out/bench_fp8.err
python tools/show_bench.py gpurun_out/bench_fp8.json
timeout 300 python tools/cublas_compare.py 2>&1 | tee gpurun_out/cublas.log
