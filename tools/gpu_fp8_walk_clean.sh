#!/bin/bash
O=gpurun_out/r02i
mkdir -p $O
SWEEP_ONE="1,8,0,0" timeout 300 python tools/fp8_raster_sweep.py once > $O/once_plain.log 2>&1 || exit 1
for c in "1,8,0,0" "0,8,0,1" "0,8,2,1" "0,8,0,0" "1,8,1,2" "1,8,1,0" "0,16,2,1" "0,4,0,1"; do
  SWEEP_ONE=$c timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:gemm_sm100_kernel --csv --log-file $O/dram_$c.csv python tools/fp8_raster_sweep.py once > $O/once_$c.log 2>&1
  echo "== $c"; grep -E "dram__bytes|gpu__time" $O/dram_$c.csv | awk -F'","' '{print $1, $(NF-2), $NF}' | tail -6
done
