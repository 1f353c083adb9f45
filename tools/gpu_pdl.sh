#!/bin/bash
# PDL A/B: small-M sweep in graph mode with ring depths that do / do not let two CTAs share an SM; large shapes back to back.
mkdir -p gpurun_out; L=gpurun_out/pdl.log; : > $L
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
for pdl in 0 1; do
  for m in 1 16 64 128; do
    for st in 0 3 4 6; do
      for sk in 4 8; do
        DGA_PDL=$pdl timeout 120 python tools/time_plan.py $m 4096 7168 kernelSerial=3 splitkFactor=$sk stages=$st graph=1 iters=2000 >> $L 2>&1
      done
    done
  done
  DGA_PDL=$pdl timeout 120 python tools/time_plan.py 4096 7168 2048 >> $L 2>&1
  DGA_PDL=$pdl timeout 120 python tools/time_plan.py 2048 4096 7168 >> $L 2>&1
  DGA_PDL=$pdl timeout 120 python tools/time_plan.py 512 4096 7168 graph=1 iters=1000 >> $L 2>&1
  DGA_PDL=$pdl timeout 120 python tools/time_plan.py 4096 7168 2048 graph=1 iters=400 >> $L 2>&1
done
grep -v "^$" $L
