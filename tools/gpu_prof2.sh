#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
for w in m192 m256 m384; do
  timeout 300 python bench.py --workload $w --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
done
python tools/show_bench.py gpurun_out/bench_m192.json gpurun_out/bench_m256.json gpurun_out/bench_m384.json
CMD="python bench.py --workload m64 --steps 3 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain_m64.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_m64.csv $CMD > /dev/null 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_smallm -s 60 -c 2 -f -o gpurun_out/prof_m64 $CMD > gpurun_out/ncu_m64.log 2>&1
CMD="python bench.py --workload fp8 --steps 3 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain_fp8.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_sm100 -s 6 -c 2 -f -o gpurun_out/prof_fp8 $CMD > gpurun_out/ncu_fp8.log 2>&1
ls -la gpurun_out/*.ncu-rep
