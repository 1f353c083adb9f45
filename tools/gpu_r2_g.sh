#!/bin/bash
# Round 2: the reworked bench line (config 5 anchor, sustained block, e2e copy rates), repack workload, fp8 with the
# walk rule (+ ncu DRAM bytes), walk-order A/B on the big bf16 shapes.
O=gpurun_out/r02g
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
timeout 600 python bench.py > $O/bench_default.json 2> $O/bench_default.err; tail -3 $O/bench_default.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 3 > $O/bench_reference.json 2>&1
for w in repack fp8 fp8_fp32scales m512 m2048 m8192; do
  timeout 300 python bench.py --workload $w --steps 30 --warmup 5 --no-cpu-baseline > $O/bench_$w.json 2> $O/bench_$w.err
done
python tools/show_bench.py $O/bench_*.json
T="timeout 300 python tools/ab_options.py"
{
$T 8192 4096 7168 -- "" "raster_along_m=1,raster_group=8,hint_a=1" "raster_along_m=1,raster_group=8" "raster_along_m=0,raster_group=8,hint_b=1" "raster_along_m=0,raster_group=4"
$T 16384 57344 7168 rounds=3 iters=5 -- "" "raster_along_m=0,raster_group=8,hint_b=1" "raster_along_m=1,raster_group=8,hint_a=1" "raster_along_m=1,raster_group=16,hint_a=1" "raster_along_m=0,raster_group=16,hint_b=1"
$T 4096 7168 2048 -- "" "raster_along_m=0,raster_group=16" "raster_along_m=0,raster_group=8" "raster_along_m=1,raster_group=4" "raster_along_m=1,raster_group=16"
} > $O/walk_ab.log 2>&1
cat $O/walk_ab.log
CMD="python bench.py --workload fp8 --steps 3 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > $O/plain_fp8.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_sm100 -s 6 -c 2 -f -o $O/prof_fp8 $CMD > $O/ncu_fp8.log 2>&1
ls -la $O
