#!/bin/bash
rm -rf gpurun_out/audit4; mkdir -p gpurun_out/audit4
SHAPES=""
for nk in 4096x7168 2112x7168 7168x16384 512x4096; do
  for m in 1 16 64 128; do SHAPES="$SHAPES ${m}x${nk}"; done
done
LAYOUT_B=0 timeout 900 python tools/selector_audit.py gpurun_out/audit4 $SHAPES > gpurun_out/audit4/audit.log 2> gpurun_out/audit4/err.txt
grep -v ERROR gpurun_out/audit4/err.txt | tail -3
LAYOUT_B=0 python tools/audit_offline.py gpurun_out/audit4 2>/dev/null
