#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/sk3.log
P="512 4096 7168 kernelSerial=4 ctaGroup=2 n1=256 splitkFactor=2"
for dbg in 0 8 12 13; do DGA_DEBUG_FLAGS=$dbg timeout 120 python tools/time_plan.py $P graph=1 2>&1 | tail -1 >> gpurun_out/sk3.log; done
DGA_DEBUG_FLAGS=16 timeout 120 python tools/trace_common.py $P 2>&1 | tail -3 >> gpurun_out/sk3.log
DGA_DEBUG_FLAGS=16 timeout 120 python tools/trace_common.py 512 4096 7168 2>&1 | tail -3 >> gpurun_out/sk3.log
DGA_DEBUG_FLAGS=16 DGA_PDL=0 timeout 120 python tools/trace_common.py $P 2>&1 | tail -3 >> gpurun_out/sk3.log
P="512 4096 7168 kernelSerial=0 ctaGroup=2 n1=256"
for dbg in 0 4; do DGA_DEBUG_FLAGS=$dbg timeout 120 python tools/time_plan.py $P graph=1 2>&1 | tail -1 >> gpurun_out/sk3.log; done
cat gpurun_out/sk3.log
