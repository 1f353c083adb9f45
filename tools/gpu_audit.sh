#!/bin/bash
rm -rf gpurun_out/audit; mkdir -p gpurun_out/audit
SHAPES=""
for nk in 2112x7168 24576x1536 32768x512 7168x16384 4096x7168 7168x2048 1536x7168 576x7168 7168x2304; do
  for m in 64 128 256 512 1024 2048 4096; do SHAPES="$SHAPES ${m}x${nk}"; done
done
timeout 1700 python tools/selector_audit.py gpurun_out/audit $SHAPES 2> gpurun_out/audit/err.txt | tee gpurun_out/audit/audit.log | tail -75
