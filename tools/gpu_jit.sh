#!/bin/bash
# JIT-specialised vs dynamic-shape kernel: parity tests, then alternating timings of both on the bench shapes.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_jit.py -x -q > gpurun_out/pytest_jit.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_jit.log
tail -15 gpurun_out/pytest_jit.log
export DGA_ROOT_DIR=/tmp/dga_jit_ab
: > gpurun_out/jit_ab.log
for round in 1 2 3; do
  for jit in 0 1; do
    for shape in "4096 7168 2048" "512 4096 7168" "8192 4096 7168" "2048 4096 7168" "1024 7168 2048 graph=1" "1024 1024 1024 graph=1"; do
      echo -n "jit=$jit r$round: " >> gpurun_out/jit_ab.log
      DGA_JIT=$jit python tools/time_plan.py $shape 2>&1 | tail -1 >> gpurun_out/jit_ab.log
    done
  done
done
sort -k3,5 -s gpurun_out/jit_ab.log
