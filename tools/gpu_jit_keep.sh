#!/bin/bash
# Which group of compile-time constants is the specialised kernel's gain? (DGA_JIT_KEEP_RT bits, jit_entry.cuh)
mkdir -p gpurun_out
export DGA_ROOT_DIR=/tmp/dga_jit_keep
: > gpurun_out/jit_keep.log
for shape in "512 4096 7168 graph=1" "1024 1024 1024 graph=1" "1024 7168 2048 graph=1"; do
  echo -n "dynamic: " >> gpurun_out/jit_keep.log
  DGA_JIT=0 python tools/time_plan.py $shape 2>&1 | tail -1 >> gpurun_out/jit_keep.log
  for keep in 0 1 2 4 8 16 32 64 127 126; do
    echo -n "keep=$keep: " >> gpurun_out/jit_keep.log
    DGA_JIT=1 DGA_JIT_KEEP_RT=$keep python tools/time_plan.py $shape 2>&1 | tail -1 >> gpurun_out/jit_keep.log
  done
done
cat gpurun_out/jit_keep.log
