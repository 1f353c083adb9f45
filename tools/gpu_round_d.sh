#!/bin/bash
# Last visit of round 1 (tests and smoke ran in an earlier call of the same session): bench lines on the final code,
# ncu launch lists and full captures of the two kernels the roofline lines name (config 2, decode at M=64).
mkdir -p gpurun_out/r01d
O=gpurun_out/r01d
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $O/gpu.txt 2>&1
timeout 300 python bench.py > $O/bench_config2.json 2> $O/bench_config2.err
for w in m1 m16 m64 m128 m256 m512 m2048 m8192 fp8 fp8_fp32scales moe; do
  timeout 200 python bench.py --workload $w --steps 30 --warmup 5 --no-cpu-baseline > $O/bench_$w.json 2> $O/bench_$w.err
done
timeout 200 python bench.py --impl reference --steps 5 --warmup 3 > $O/bench_reference.json 2>&1
CMD="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
timeout 200 $CMD > $O/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches_config2.csv $CMD > $O/ncu_list.log 2>&1
timeout 200 $CMD > $O/plain2.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_sm100 -s 4 -c 2 -f -o $O/prof_config2 $CMD > $O/ncu_full.log 2>&1
CMD="python bench.py --workload m64 --steps 5 --warmup 3 --no-cpu-baseline"
timeout 200 $CMD > $O/plain_m64.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches_m64.csv $CMD > $O/ncu_list_m64.log 2>&1
timeout 200 $CMD > $O/plain2_m64.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_smallm -s 4 -c 2 -f -o $O/prof_m64 $CMD > $O/ncu_full_m64.log 2>&1
ls -la $O
python tools/show_bench.py $O/bench_*.json
