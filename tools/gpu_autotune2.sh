#!/bin/bash
# Tuned-table sweep on the end-of-round kernels (stream-K slice counts among the candidates).
rm -rf gpurun_out/autotune2; mkdir -p gpurun_out/autotune2
timeout 1500 python -m deepgemm_ascend_b200.autotune --result-dir gpurun_out/autotune2 --csv gpurun_out/autotune2/tuned.csv \
  --shapes 4096x7168x2048 256x4096x7168 512x4096x7168 2048x4096x7168 1024x7168x2048 \
  > gpurun_out/autotune2/log.txt 2>&1
echo rc=$?; grep -v "ERROR" gpurun_out/autotune2/log.txt | tail -12
python - <<'PY'
import deepgemm_ascend_b200 as dga
for s in [(4096,7168,2048),(256,4096,7168),(512,4096,7168),(2048,4096,7168),(1024,7168,2048)]:
    p = dga.select_kernel(*s)
    print("selector", s, {k: p[k] for k in ("kernelSerial", "ctaGroup", "n1", "splitkFactor", "stages")})
PY
