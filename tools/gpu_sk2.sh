#!/bin/bash
# Aligned stream-K (kernelSerial 4): parity, then timings against the selector's plan on the ragged-wave shapes.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "aclnn or fuzz or padding" > gpurun_out/pytest_sk.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_sk.log
tail -5 gpurun_out/pytest_sk.log
: > gpurun_out/sk2.log
t() { timeout 120 python tools/time_plan.py "$@" graph=1 2>&1 | tail -1 >> gpurun_out/sk2.log; }
t 512 4096 7168
for s in 2 3 4; do t 512 4096 7168 kernelSerial=4 ctaGroup=2 n1=256 splitkFactor=$s; done
t 512 4096 7168 kernelSerial=4 ctaGroup=2 n1=192 splitkFactor=2
t 2048 4096 7168
for s in 0 4 5 8; do t 2048 4096 7168 kernelSerial=4 ctaGroup=2 n1=256 splitkFactor=$s; done
t 1024 4096 7168
for s in 2 3 4 6; do t 1024 4096 7168 kernelSerial=4 ctaGroup=2 n1=256 splitkFactor=$s; done
t 1024 7168 2048
for s in 2 3; do t 1024 7168 2048 kernelSerial=4 ctaGroup=2 n1=256 splitkFactor=$s; done
t 256 4096 7168
for s in 2 4; do t 256 4096 7168 kernelSerial=4 ctaGroup=2 n1=256 splitkFactor=$s; done
for s in 2 4; do t 256 4096 7168 kernelSerial=4 ctaGroup=2 n1=128 splitkFactor=$s; done
t 384 4096 7168
for s in 2 3; do t 384 4096 7168 kernelSerial=4 ctaGroup=2 n1=256 splitkFactor=$s; done
t 8192 4096 7168
t 8192 4096 7168 kernelSerial=4 ctaGroup=2 n1=256
t 4096 7168 2048 kernelSerial=4 ctaGroup=2 n1=256
cat gpurun_out/sk2.log
