O=gpurun_out/r02s
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "aclnn_matches or config3" > $O/pytest_parity.log 2>&1; echo "rc=$?" >> $O/pytest_parity.log
tail -3 $O/pytest_parity.log
T="timeout 120 python tools/time_plan.py"
{
for rep in 1 2; do
for shape in "512 4096 7168" "384 4096 7168" "256 4096 7168" "512 2048 7168" "1024 2048 7168" "256 7168 7168"; do
$T $shape graph=1
done
done
} > $O/ks6_graph.log 2>&1
grep "us " $O/ks6_graph.log
