#!/bin/bash
# One visit to a B200 box (run it through gpurun): parity tests, smoke, bench lines, ncu captures. Everything is written
# under $OUT (default gpurun_out/suite). Replaces the one-off gpu_*.sh scripts of round 1.
#   TESTS=1|0            python -m pytest tests -m gpu (default 1)
#   WORKLOADS="..."      bench.py --workload names, one line each (default: the BASELINE configs that fit one GPU)
#   NCU="..."            workloads to capture with `ncu --set full` after their plain run exited 0 (default: none);
#                        the kernel regex follows from the workload (decode / fp32-scale fp8 / persistent)
#   LAUNCH_LIST="..."    workloads whose launch list (gpu__time_duration per launch) is recorded (default: none)
#   KEEP_REP="..."       workloads whose .ncu-rep is kept (default: none; every capture leaves ncu_summary_*.json,
#                        traffic_*.json and the raw metric page ncu_raw_*.csv)
#   STEPS / WARMUP       for the non-default bench lines (default 30 / 5)
OUT=${OUT:-gpurun_out/suite}
STEPS=${STEPS:-30}
WARMUP=${WARMUP:-5}
WORKLOADS=${WORKLOADS-"m1 m16 m64 m128 m64pair m128pair m256 m512 m2048 m8192 fp8 fp8_fp32scales moe moe256 repack config5"}
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/gpu.txt 2>&1
if [ "${TESTS:-1}" = "1" ]; then
  timeout 1800 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "rc=$?" >> $OUT/pytest_gpu.log
  tail -4 $OUT/pytest_gpu.log
  timeout 300 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; tail -1 $OUT/smoke.log
fi
timeout 900 python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 3 > $OUT/bench_reference.json 2>&1
for w in $WORKLOADS; do
  timeout 600 python bench.py --workload $w --steps $STEPS --warmup $WARMUP --no-cpu-baseline > $OUT/bench_$w.json 2> $OUT/bench_$w.err
done
python tools/show_bench.py $OUT/bench_*.json
kernel_regex() {
  case $1 in
    m1|m16|m64|m128|m*pair|m*n*) echo gemm_smallm ;;
    fp8_fp32scales) echo gemm_fp8_sm100 ;;
    *) echo gemm_sm100 ;;
  esac
}
for w in $LAUNCH_LIST; do
  CMD="python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline --no-config5 --sustain-seconds 0"
  timeout 300 $CMD > $OUT/plain_list_$w.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_$w.csv $CMD > $OUT/ncu_list_$w.log 2>&1
done
for w in $NCU; do
  CMD="python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline --no-config5 --sustain-seconds 0"
  timeout 300 $CMD > $OUT/plain_$w.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$(kernel_regex $w) -s 6 -c 2 -f -o $OUT/prof_$w $CMD > $OUT/ncu_$w.log 2>&1
  # gpurun brings back at most 64 MiB: boil the report down here (summary, DRAM traffic, the raw metric page) and keep the
  # report itself only on request
  if [ -f $OUT/prof_$w.ncu-rep ]; then
    python tools/ncu_summary.py $OUT/prof_$w.ncu-rep $OUT/ncu_summary_$w.json $OUT/traffic_$w.json > /dev/null 2>&1
    ncu -i $OUT/prof_$w.ncu-rep --page raw --csv > $OUT/ncu_raw_$w.csv 2>/dev/null
    case " $KEEP_REP " in *" $w "*) ;; *) rm -f $OUT/prof_$w.ncu-rep ;; esac
  fi
done
ls -la $OUT
