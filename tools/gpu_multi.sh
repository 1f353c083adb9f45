#!/bin/bash
mkdir -p gpurun_out
G=${G:-2}
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 400 python bench.py --gpus $G --workload config5 --steps 10 --warmup 3 > gpurun_out/bench_config5_g$G.json 2> gpurun_out/bench_config5_g$G.err
echo "rc=$?"; tail -5 gpurun_out/bench_config5_g$G.err; cat gpurun_out/bench_config5_g$G.json
timeout 400 python bench.py --gpus $G --steps 200 --no-cpu-baseline > gpurun_out/bench_config2_g$G.json 2> gpurun_out/bench_config2_g$G.err
echo "rc=$?"; tail -3 gpurun_out/bench_config2_g$G.err; python tools/show_bench.py gpurun_out/bench_config2_g$G.json
