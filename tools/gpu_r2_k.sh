O=gpurun_out/r02k
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "aclnn_matches" > $O/pytest_parity.log 2>&1; echo "rc=$?" >> $O/pytest_parity.log
tail -5 $O/pytest_parity.log
T="timeout 300 python tools/ab_options.py"
{
$T 512 4096 7168 -- "" "kernelSerial=6,ctaGroup=2,n1=256" "kernelSerial=0,ctaGroup=2,n1=128"
$T 384 4096 7168 -- "" "kernelSerial=6,ctaGroup=2,n1=256"
$T 256 4096 7168 -- "" "kernelSerial=6,ctaGroup=2,n1=256" "kernelSerial=6,ctaGroup=2,n1=128"
$T 512 2048 7168 -- "" "kernelSerial=6,ctaGroup=2,n1=256" "kernelSerial=6,ctaGroup=2,n1=128"
$T 512 7168 2048 -- "" "kernelSerial=6,ctaGroup=2,n1=256"
$T 256 7168 7168 -- "" "kernelSerial=6,ctaGroup=2,n1=256"
$T 1024 2048 7168 -- "" "kernelSerial=6,ctaGroup=2,n1=256"
$T 512 4096 2048 -- "" "kernelSerial=6,ctaGroup=2,n1=256"
$T 2048 4096 7168 -- "kernelSerial=0,ctaGroup=2,n1=256" "kernelSerial=0,ctaGroup=2,n1=256,clusterShape=4"
$T 4096 4096 7168 -- "" "kernelSerial=0,ctaGroup=2,n1=256" "kernelSerial=0,ctaGroup=2,n1=256,clusterShape=4"
$T 2048 7168 2048 -- "" "kernelSerial=0,ctaGroup=2,n1=256" "kernelSerial=0,ctaGroup=2,n1=256,clusterShape=4"
$T 2048 2048 7168 -- "" "kernelSerial=0,ctaGroup=2,n1=256" "kernelSerial=0,ctaGroup=2,n1=256,clusterShape=4"
$T 4096 2048 7168 -- "" "kernelSerial=0,ctaGroup=2,n1=256" "kernelSerial=0,ctaGroup=2,n1=256,clusterShape=4"
$T 1536 4096 7168 -- "" "kernelSerial=0,ctaGroup=2,n1=256" "kernelSerial=0,ctaGroup=2,n1=256,clusterShape=4"
} > $O/ab.log 2>&1
cat $O/ab.log
