#!/bin/bash
mkdir -p gpurun_out; L=gpurun_out/sk.log; : > $L
for dbg in 0 8 12 13; do
for M in 512 2048; do
  DGA_DEBUG_FLAGS=$dbg timeout 120 python tools/time_plan.py $M 4096 7168 kernelSerial=4 ctaGroup=2 n1=256 >> $L 2>&1
done; done
timeout 120 python tools/time_plan.py 1024 4096 7168 kernelSerial=4 ctaGroup=2 n1=256 >> $L 2>&1
timeout 120 python tools/time_plan.py 4096 7168 2048 kernelSerial=4 ctaGroup=2 n1=256 >> $L 2>&1
grep -v Warn $L
