#!/bin/bash
# Round 2: clusters of four CTAs with A multicast (clusterShape 4) - parity, then A/B against the pair kernel.
O=gpurun_out/r02b
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "aclnn_matches" > $O/pytest_parity.log 2>&1; echo "rc=$?" >> $O/pytest_parity.log
tail -5 $O/pytest_parity.log
T="timeout 120 python tools/time_plan.py"
{
for rep in 1 2; do
$T 512 4096 7168 graph=1
$T 512 4096 7168 graph=1 kernelSerial=0 ctaGroup=2 n1=128
$T 512 4096 7168 graph=1 kernelSerial=0 ctaGroup=2 n1=128 clusterShape=4
$T 512 4096 7168 graph=1 kernelSerial=0 ctaGroup=2 n1=256 clusterShape=4
$T 512 4096 7168 graph=1 kernelSerial=0 ctaGroup=2 n1=192 clusterShape=4
$T 256 4096 7168 graph=1
$T 256 4096 7168 graph=1 kernelSerial=0 ctaGroup=2 n1=64 clusterShape=4
$T 256 4096 7168 graph=1 kernelSerial=0 ctaGroup=2 n1=128 clusterShape=4
$T 1024 4096 7168 graph=1
$T 1024 4096 7168 graph=1 kernelSerial=0 ctaGroup=2 n1=128 clusterShape=4
$T 1024 4096 7168 graph=1 kernelSerial=0 ctaGroup=2 n1=256 clusterShape=4
$T 2048 4096 7168
$T 2048 4096 7168 kernelSerial=0 ctaGroup=2 n1=256 clusterShape=4
$T 4096 7168 2048
$T 4096 7168 2048 kernelSerial=0 ctaGroup=2 n1=256 clusterShape=4
$T 4096 7168 2048 kernelSerial=0 ctaGroup=2 n1=224 clusterShape=4
$T 8192 4096 7168
$T 8192 4096 7168 kernelSerial=0 ctaGroup=2 n1=256 clusterShape=4
done
} > $O/mc_ab.log 2>&1
cat $O/mc_ab.log
CMD="python tools/time_plan.py 512 4096 7168 kernelSerial=0 ctaGroup=2 n1=128 clusterShape=4 iters=20"
timeout 120 $CMD > $O/plain_mc.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_sm100 -s 12 -c 2 -f -o $O/prof_m512_mc $CMD > $O/ncu_m512_mc.log 2>&1
ls -la $O
