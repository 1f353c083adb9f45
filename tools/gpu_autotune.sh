#!/bin/bash
mkdir -p gpurun_out/autotune
timeout 2400 python -m deepgemm_ascend_b200.autotune --result-dir gpurun_out/autotune --csv gpurun_out/autotune/tuned.csv \
  --shapes 4096x7168x2048 1x4096x7168 16x4096x7168 64x4096x7168 128x4096x7168 256x4096x7168 512x4096x7168 1024x4096x7168 \
  2048x4096x7168 8192x4096x7168 1024x1024x1024 4096x4096x4096 64x7168x2048 128x7168x2048 4096x2112x7168 4096x7168x16384 \
  > gpurun_out/autotune/log.txt 2>&1
echo rc=$?; tail -25 gpurun_out/autotune/log.txt
