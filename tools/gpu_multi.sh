#!/bin/bash
# Multi-GPU visit (gpurun --gpus N): the 2-GPU oracle test of the fused N-shard path, then the driver's own launch line
# for both arms. N = number of GPUs of the lease (default 2).
N=${N:-2}
OUT=${OUT:-gpurun_out/multi_g$N}
mkdir -p $OUT
nvidia-smi --query-gpu=index,name,clocks.max.sm --format=csv > $OUT/gpu.txt 2>&1
nvidia-smi topo -m > $OUT/topo.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_nshard.py -x -q -s > $OUT/pytest_nshard.log 2>&1; echo "rc=$?" >> $OUT/pytest_nshard.log
tail -6 $OUT/pytest_nshard.log
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N"
timeout 900 $RUN --steps 20 --warmup 3 > $OUT/bench_g$N.json 2> $OUT/bench_g$N.err; tail -3 $OUT/bench_g$N.err
timeout 600 $RUN --impl reference --steps 3 --warmup 3 > $OUT/bench_reference_g$N.json 2> $OUT/bench_reference_g$N.err
timeout 600 $RUN --workload config2 --steps 100 --warmup 5 --no-cpu-baseline > $OUT/bench_config2_weak_g$N.json 2> $OUT/bench_config2_weak_g$N.err
python tools/show_bench.py $OUT/bench_g$N.json $OUT/bench_config2_weak_g$N.json
python - <<PY
import json
d = json.loads(open("$OUT/bench_g$N.json").read().strip().splitlines()[-1])
print(json.dumps(d.get("config5"), indent=1)); print(json.dumps(d.get("e2e"), indent=1))
PY
ls -la $OUT
