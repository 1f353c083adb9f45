#!/bin/bash
# Round 2, third visit: full GPU test suite on the new code (cluster-4 plans, fp8 UE8M0 config-4 test, config 5 full
# width), then an ncu capture (with source) of the fp32-scale fp8 promotion kernel.
O=gpurun_out/r02c
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -8 $O/pytest_gpu.log
CMD="python bench.py --workload fp8_fp32scales --steps 3 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > $O/plain_fp8p.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_fp8_sm100 -s 3 -c 1 -f -o $O/prof_fp8_promo $CMD > $O/ncu_fp8p.log 2>&1
tail -2 $O/plain_fp8p.log | head -c 600
ls -la $O
