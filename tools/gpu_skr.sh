#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log
timeout 1500 python tools/sweep_splitk_reduce.py > gpurun_out/sweep_splitk_reduce.log 2>&1
grep "current\|BEST" gpurun_out/sweep_splitk_reduce.log
