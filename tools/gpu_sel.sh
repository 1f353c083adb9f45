#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
: > gpurun_out/sel.log
t() { timeout 120 python tools/time_plan.py "$@" 2>&1 | tail -1 >> gpurun_out/sel.log; }
for round in 1 2; do
t 4096 4096 7168
t 4096 4096 7168 kernelSerial=0 ctaGroup=2 n1=256
t 2048 7168 16384
t 2048 7168 16384 kernelSerial=0 ctaGroup=2 n1=224
t 128 7168 2048 graph=1
t 128 7168 2048 kernelSerial=3 splitkFactor=2 stages=5 graph=1
t 128 32768 512 graph=1
t 128 32768 512 kernelSerial=3 splitkFactor=1 stages=3 graph=1
done
cat gpurun_out/sel.log
