#!/bin/bash
O=gpurun_out/r02f
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_fp8.py -x -q > $O/pytest_fp8.log 2>&1; echo "rc=$?" >> $O/pytest_fp8.log
tail -5 $O/pytest_fp8.log
for i in 1 2; do
for w in fp8_fp32scales fp8; do
  timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_${w}_$i.json 2> $O/bench_${w}_$i.err
  head -c 200 $O/bench_${w}_$i.json; echo
done
done
CMD="python bench.py --workload fp8_fp32scales --steps 3 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > $O/plain_fp8p.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_fp8_sm100 -s 3 -c 1 -f -o $O/prof_fp8_promo $CMD > $O/ncu_fp8p.log 2>&1
ls -la $O
