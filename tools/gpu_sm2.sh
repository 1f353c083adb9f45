#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
L=gpurun_out/trace_smallm.log; : > $L
for cfg in "16 4096 7168 8 5 1" "64 4096 7168 4 0 1" ; do
  echo "== $cfg" >> $L
  DGA_DEBUG_FLAGS=16 timeout 200 python tools/trace_smallm.py $cfg >> $L 2>&1
done
grep -A3 "==\|rep 2" $L
timeout 1200 python tools/sweep_smallm.py 1,16,64,128 > gpurun_out/sweep_smallm.log 2>&1
grep BEST gpurun_out/sweep_smallm.log
