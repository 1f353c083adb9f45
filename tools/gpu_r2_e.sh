#!/bin/bash
# Round 2, fifth visit: fp32-scale fp8 kernel with column halves (parity, bench, ncu with source), interleaved A/B of
# the fp8 walk orders.
O=gpurun_out/r02e
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_fp8.py -x -q > $O/pytest_fp8.log 2>&1; echo "rc=$?" >> $O/pytest_fp8.log
tail -5 $O/pytest_fp8.log
for w in fp8_fp32scales; do
  timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_$w.json 2> $O/bench_$w.err
  head -c 300 $O/bench_$w.json; echo
done
timeout 600 python tools/fp8_raster_sweep.py ab > $O/fp8_raster_ab.log 2>&1
cat $O/fp8_raster_ab.log
CMD="python bench.py --workload fp8_fp32scales --steps 3 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > $O/plain_fp8p.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_fp8_sm100 -s 3 -c 1 -f -o $O/prof_fp8_promo $CMD > $O/ncu_fp8p.log 2>&1
ls -la $O
