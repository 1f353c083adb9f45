#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python tools/ab_plans.py 4096 7168 2048 kernelSerial=0,ctaGroup=2,n1=224,stages=6 kernelSerial=0,ctaGroup=2,n1=224,stages=7 kernelSerial=0,ctaGroup=2,n1=256,stages=6 2>&1 | tee gpurun_out/ab.log
python tools/ab_plans.py 8192 7168 2048 kernelSerial=0,ctaGroup=2,n1=224,stages=6 kernelSerial=0,ctaGroup=2,n1=224,stages=7 2>&1 | tee -a gpurun_out/ab.log
