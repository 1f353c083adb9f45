#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/sk4.log
t() { timeout 120 python tools/time_plan.py "$@" graph=1 2>&1 | tail -1 >> gpurun_out/sk4.log; }
for m in 192 256; do
  for nk in "2048 7168" "4096 7168" "7168 2048" "4096 2048" "7168 7168"; do
    t $m $nk
    t $m $nk kernelSerial=4 ctaGroup=2 n1=128 splitkFactor=2
    t $m $nk kernelSerial=4 ctaGroup=2 n1=128 splitkFactor=4
    t $m $nk kernelSerial=4 ctaGroup=2 n1=64 splitkFactor=2
  done
done
cat gpurun_out/sk4.log
