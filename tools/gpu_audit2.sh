#!/bin/bash
rm -rf gpurun_out/audit2; mkdir -p gpurun_out/audit2
SHAPES=""
for nk in 2112x7168 1536x7168 576x7168 2048x7168 1024x4096; do
  for m in 192 256 384 512 768 1024; do SHAPES="$SHAPES ${m}x${nk}"; done
done
timeout 900 python tools/selector_audit.py gpurun_out/audit2 $SHAPES > gpurun_out/audit2/audit.log 2> gpurun_out/audit2/err.txt
python tools/audit_offline.py gpurun_out/audit2 2>/dev/null
