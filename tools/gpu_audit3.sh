#!/bin/bash
# Selector audit for row-major mat2 (the framework functions' layout).
rm -rf gpurun_out/audit3; mkdir -p gpurun_out/audit3
SHAPES=""
for nk in 4096x7168 7168x2048 2112x7168 2048x2048; do
  for m in 128 256 512 1024 2048 4096; do SHAPES="$SHAPES ${m}x${nk}"; done
done
LAYOUT_B=0 timeout 900 python tools/selector_audit.py gpurun_out/audit3 $SHAPES > gpurun_out/audit3/audit.log 2> gpurun_out/audit3/err.txt
tail -3 gpurun_out/audit3/err.txt
LAYOUT_B=0 python tools/audit_offline.py gpurun_out/audit3 2>/dev/null
