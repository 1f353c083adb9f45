#!/bin/bash
# Round 2, fourth visit: GPU tests on the reworked fp32-scale fp8 kernel + host fixes, its bench line, the fp8 walk
# order / L2 hint sweep (time, then DRAM bytes under ncu with a light metric set).
O=gpurun_out/r02d
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -8 $O/pytest_gpu.log
for w in fp8_fp32scales fp8; do
  timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_$w.json 2> $O/bench_$w.err
  head -c 400 $O/bench_$w.json; echo
done
timeout 600 python tools/fp8_raster_sweep.py time > $O/fp8_raster_time.log 2>&1
sort -t: -k2 -n $O/fp8_raster_time.log | head -12
timeout 300 python tools/fp8_raster_sweep.py once > $O/fp8_raster_once_plain.log 2>&1 &&
timeout 1200 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:gemm_sm100_kernel --csv --log-file $O/fp8_raster_dram.csv python tools/fp8_raster_sweep.py once > $O/fp8_raster_once.log 2>&1
ls -la $O
