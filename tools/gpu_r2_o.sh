O=gpurun_out/r02o
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "aclnn_matches" > $O/pytest_parity.log 2>&1; echo "rc=$?" >> $O/pytest_parity.log
tail -5 $O/pytest_parity.log
T="timeout 120 python tools/time_plan.py"
{
for m in 64 128 1 16 32 96; do
$T $m 4096 7168 graph=1
for sp in 8 9 4 6; do
$T $m 4096 7168 graph=1 kernelSerial=7 splitkFactor=$sp
done
$T $m 4096 7168 graph=1 kernelSerial=7 splitkFactor=8 stages=3
$T $m 4096 7168 graph=1 kernelSerial=7 splitkFactor=9 stages=6
done
} > $O/decode_ws.log 2>&1
grep "us " $O/decode_ws.log
