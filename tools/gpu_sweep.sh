#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 1200 python tools/sweep_smallm.py 1,16,64,128 > gpurun_out/sweep_smallm.log 2>&1
grep BEST gpurun_out/sweep_smallm.log
