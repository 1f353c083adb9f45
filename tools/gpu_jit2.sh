#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_jit.py -x -q > gpurun_out/pytest_jit.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_jit.log
tail -6 gpurun_out/pytest_jit.log
export DGA_ROOT_DIR=/tmp/dga_jit_ab
: > gpurun_out/jit_ab2.log
for round in 1 2 3; do
  for jit in 0 1; do
    for shape in "256 2048 7168" "512 1536 7168" "192 576 7168"; do
      echo -n "jit=$jit r$round: " >> gpurun_out/jit_ab2.log
      DGA_JIT=$jit timeout 120 python tools/time_plan.py $shape graph=1 2>&1 | tail -1 >> gpurun_out/jit_ab2.log
    done
  done
done
sort -k3,5 -s gpurun_out/jit_ab2.log
