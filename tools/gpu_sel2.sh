#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
: > gpurun_out/sel2.log
t() { timeout 120 python tools/time_plan.py "$@" graph=1 2>&1 | tail -1 >> gpurun_out/sel2.log; }
for round in 1 2; do
t 512 2112 7168
t 512 2112 7168 kernelSerial=5 ctaGroup=2 n1=256 splitkFactor=4
t 768 1536 7168
t 768 1536 7168 kernelSerial=5 ctaGroup=2 n1=256 splitkFactor=4
t 1024 1536 7168
t 1024 1536 7168 kernelSerial=0 ctaGroup=2 n1=96
t 2048 576 7168
t 2048 576 7168 kernelSerial=0 ctaGroup=2 n1=64
t 256 1024 4096
t 256 1024 4096 kernelSerial=5 ctaGroup=2 n1=256 splitkFactor=4
t 192 1536 7168
t 192 1536 7168 kernelSerial=3 splitkFactor=8 stages=3
done
cat gpurun_out/sel2.log
