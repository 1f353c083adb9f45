O=gpurun_out/r02r
mkdir -p $O
export DGA_DEBUG_FLAGS=16
{
timeout 120 python tools/trace_common.py 512 4096 7168 kernelSerial=6 ctaGroup=2 n1=256
timeout 120 python tools/trace_common.py 512 4096 7168 kernelSerial=0 ctaGroup=2 n1=128
timeout 120 python tools/trace_common.py 256 4096 7168 kernelSerial=6 ctaGroup=2 n1=128
} > $O/trace_ks6.log 2>&1
cat $O/trace_ks6.log
unset DGA_DEBUG_FLAGS
for w in fp8_fp32scales fp8; do
  timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_$w.json 2> $O/bench_$w.err
done
python tools/show_bench.py $O/bench_*.json
