#!/bin/bash
mkdir -p gpurun_out; L=gpurun_out/probe.log; : > $L
for dbg in 0 1 2 3 4 5; do
  DGA_DEBUG_FLAGS=$dbg python tools/time_plan.py 4096 7168 2048 kernelSerial=0 ctaGroup=2 n1=224 >> $L 2>&1
  DGA_DEBUG_FLAGS=$dbg python tools/time_plan.py 4096 7168 2048 kernelSerial=0 ctaGroup=2 n1=256 >> $L 2>&1
  DGA_DEBUG_FLAGS=$dbg python tools/time_plan.py 8192 4096 7168 kernelSerial=0 ctaGroup=2 n1=256 iters=50 >> $L 2>&1
done
for st in 3 4 5 6; do python tools/time_plan.py 4096 7168 2048 kernelSerial=0 ctaGroup=2 n1=256 stages=$st >> $L 2>&1; done
python tools/time_plan.py 4096 7168 2048 kernelSerial=0 ctaGroup=1 n1=256 >> $L 2>&1
python tools/time_plan.py 4096 7168 2048 kernelSerial=0 ctaGroup=1 n1=128 >> $L 2>&1
python tools/time_plan.py 4096 7168 2048 kernelSerial=0 ctaGroup=2 n1=128 >> $L 2>&1
cat $L
