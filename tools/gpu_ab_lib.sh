#!/bin/bash
# A/B of two library builds (deepgemm_ascend_b200/lib/ab/{old,new}.so, built by hand from two source states):
# alternates the builds over a few shapes, several rounds, so power/clock drift hits both alike.
mkdir -p gpurun_out
L=deepgemm_ascend_b200/lib
: > gpurun_out/ab_lib.log
for round in 1 2 3; do
  for which in old new; do
    cp $L/ab/$which.so $L/libdga_b200.so
    for shape in "4096 7168 2048" "512 4096 7168" "8192 4096 7168" "2048 4096 7168" "256 4096 7168 graph=1" "1024 7168 2048 graph=1"; do
      echo -n "$which r$round: " >> gpurun_out/ab_lib.log
      python tools/time_plan.py $shape 2>&1 | tail -1 >> gpurun_out/ab_lib.log
    done
  done
done
cp $L/ab/new.so $L/libdga_b200.so
sort -k3,5 -s gpurun_out/ab_lib.log | awk '{print}' 
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
