#!/bin/bash
# Tuned-table sweep with the odd-multiple-of-16 tile widths among the candidates, on DeepSeek-V3 shapes where the
# hand timings (profiles/r01d_odd16_tiles.log) put them ahead; then the bench line of one of them with the table loaded.
rm -rf gpurun_out/autotune3; mkdir -p gpurun_out/autotune3
timeout 700 python -m deepgemm_ascend_b200.autotune --result-dir gpurun_out/autotune3 --csv gpurun_out/autotune3/tuned.csv \
  --shapes 4096x4096x7168 2048x2112x7168 2048x7168x2048 4096x2112x7168 \
  > gpurun_out/autotune3/log.txt 2>&1
echo rc=$?; grep -v "ERROR" gpurun_out/autotune3/log.txt | tail -8
python - <<'PY'
import deepgemm_ascend_b200 as dga
for s in [(4096,4096,7168),(2048,2112,7168),(2048,7168,2048),(4096,2112,7168)]:
    p = dga.select_kernel(*s)
    print("selector", s, {k: p[k] for k in ("kernelSerial", "ctaGroup", "n1", "splitkFactor", "stages")})
PY
for s in "4096 4096 7168" "2048 2112 7168" "2048 7168 2048" "4096 2112 7168"; do
  echo "analytic: $(python tools/time_plan.py $s | tail -1)"
  echo "tuned:    $(CACHE_FILE_PATH=gpurun_out/autotune3/tuned.csv python tools/time_plan.py $s | tail -1)"
done
