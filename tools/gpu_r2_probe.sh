#!/bin/bash
# Round 2, first visit: facts the round's design decisions need (nsys availability, tcgen05.ld rate, config 5 on one
# GPU) and baseline ncu captures of the two shapes the round works on (M=512 crossbar bytes, fp8 DRAM bytes).
O=gpurun_out/r02a
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $O/gpu.txt 2>&1
{ which nsys; ls /usr/local/cuda/bin | grep -i -E "nsys|nsight"; ls /opt/nvidia 2>&1; } > $O/nsys_probe.txt 2>&1
timeout 120 tools/probes/tmem_ld_throughput > $O/tmem_ld_throughput.log 2>&1
timeout 600 python bench.py --workload config5 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_config5_g1.json 2> $O/bench_config5_g1.err
CMD="python bench.py --workload m512 --steps 3 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > $O/plain_m512.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_sm100 -s 60 -c 2 -f -o $O/prof_m512 $CMD > $O/ncu_m512.log 2>&1
CMD="python bench.py --workload fp8 --steps 3 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > $O/plain_fp8.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_sm100 -s 6 -c 2 -f -o $O/prof_fp8 $CMD > $O/ncu_fp8.log 2>&1
ls -la $O
cat $O/nsys_probe.txt $O/tmem_ld_throughput.log
head -c 1500 $O/bench_config5_g1.json
