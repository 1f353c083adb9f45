#!/bin/bash
# GPU suite on the selector that may pick odd-multiple-of-16 widths, then its new picks against the old ones.
(time timeout 600 python -m pytest tests -m gpu -x -q) > gpurun_out/pytest_gpu4.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu4.log
tail -6 gpurun_out/pytest_gpu4.log
timeout 120 python __graft_entry__.py smoke 2>&1 | tail -1
for s in "2048 2112 7168" "4096 2112 7168" "8192 2112 7168" "4096 576 7168" "8192 576 7168"; do
  echo "selector: $(python tools/time_plan.py $s | tail -1)"
done
for s in "8192 2112 7168 n1=256" "4096 576 7168 n1=160" "4096 576 7168 n1=192" "8192 576 7168 n1=160" "8192 576 7168 n1=192"; do
  echo "old pick candidates: $(python tools/time_plan.py $s kernelSerial=0 ctaGroup=2 | tail -1)"
done
