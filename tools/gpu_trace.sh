#!/bin/bash
mkdir -p gpurun_out; L=gpurun_out/trace_smallm.log; : > $L
export DGA_DEBUG_FLAGS=16
for cfg in "16 4096 7168 8 6 1" "16 4096 7168 8 6 0" "16 4096 7168 4 6 1" "16 4096 7168 4 0 1" "64 4096 7168 4 6 1" "16 4096 28672 8 6 1"; do
  echo "== $cfg" >> $L
  timeout 200 python tools/trace_smallm.py $cfg >> $L 2>&1
done
cat $L
