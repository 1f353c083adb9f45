#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench (headline + sweep), ncu launch list + full capture of the top kernel.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --steps 50 --warmup 5 > gpurun_out/bench_config2.json 2> gpurun_out/bench_config2.err
for w in m1 m16 m64 m128 m256 m512 m2048 m8192 fp8 fp8_fp32scales moe; do
  timeout 300 python bench.py --workload $w --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
done
for w in config2 m512; do
  DGA_JIT=1 DGA_ROOT_DIR=/tmp/dga_jit_bench timeout 300 python bench.py --workload $w --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_${w}_jit.json 2> gpurun_out/bench_${w}_jit.err
done
timeout 300 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_reference.json 2>&1
CMD="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:gemm_sm100 -s 4 -c 2 -f -o gpurun_out/prof_config2 $CMD > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log; python tools/show_bench.py gpurun_out/bench_*.json
