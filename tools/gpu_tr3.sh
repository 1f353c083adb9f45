#!/bin/bash
mkdir -p gpurun_out; L=gpurun_out/trace_smallm.log; : > $L
for cfg in "128 4096 7168 4 0 1" "128 4096 7168 4 3 1" "128 4096 7168 8 3 1" "64 4096 7168 8 3 1" "1 4096 7168 8 4 1"; do
  echo "== $cfg" >> $L
  DGA_DEBUG_FLAGS=16 timeout 200 python tools/trace_smallm.py $cfg >> $L 2>&1
done
grep -A3 "==\|rep 2" $L | grep -v "^--\|rep 0"
