#!/bin/bash
# The driver's own launch lines at N=2 (reference arm first, then ours), plus the fused-gather config 5 line.
mkdir -p gpurun_out/g2
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $R bench.py --impl reference --gpus 2 --steps 5 --warmup 3 > gpurun_out/g2/ref.json 2> gpurun_out/g2/ref.err; echo "ref rc=$?"
timeout 300 $R bench.py --gpus 2 --steps 100 --warmup 5 > gpurun_out/g2/ours.json 2> gpurun_out/g2/ours.err; echo "ours rc=$?"
timeout 400 $R bench.py --gpus 2 --workload config5 --steps 10 --warmup 3 > gpurun_out/g2/config5.json 2> gpurun_out/g2/config5.err; echo "config5 rc=$?"
for f in ref ours config5; do tail -2 gpurun_out/g2/$f.err; cut -c1-420 gpurun_out/g2/$f.json; echo; done
