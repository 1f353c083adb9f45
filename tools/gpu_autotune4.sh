#!/bin/bash
# Tuned table for DeepSeek-V3 dense prefill shapes (bf16, weight stored [N,K]); shipped as
# deepgemm_ascend_b200/tuned/b200_deepseek_v3_bf16.csv (opt-in through CACHE_FILE_PATH).
rm -rf gpurun_out/autotune4; mkdir -p gpurun_out/autotune4
SHAPES=""
for m in 1024 2048 4096; do
  for nk in 2112x7168 24576x1536 32768x512 7168x16384 4096x7168 7168x2048 3072x1536 1536x7168; do SHAPES="$SHAPES ${m}x${nk}"; done
done
timeout 800 python -m deepgemm_ascend_b200.autotune --result-dir gpurun_out/autotune4 --csv gpurun_out/autotune4/tuned.csv \
  --shapes $SHAPES > gpurun_out/autotune4/log.txt 2>&1
echo rc=$?; grep -v "ERROR" gpurun_out/autotune4/log.txt | tail -30
python - <<'PY'
import csv
import deepgemm_ascend_b200 as dga
rows = list(csv.DictReader(open("gpurun_out/autotune4/tuned.csv")))
same = 0
for r in rows:
    p = dga.select_kernel(int(r["m"]), int(r["n"]), int(r["k"]))
    eq = all(int(r[k]) == p[k] for k in ("kernelSerial", "ctaGroup", "n1", "splitkFactor"))
    same += eq
    if not eq:
        print("differs", r["m"], r["n"], r["k"], "tuned", {k: int(r[k]) for k in ("kernelSerial", "ctaGroup", "n1", "splitkFactor", "stages")},
              "selector", {k: p[k] for k in ("kernelSerial", "ctaGroup", "n1", "splitkFactor", "stages")})
print("%d of %d tuned plans equal the selector's" % (same, len(rows)))
PY
