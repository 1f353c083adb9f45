O=gpurun_out/r02l
mkdir -p $O
T="timeout 120 python tools/time_plan.py"
{
for rep in 1 2; do
for shape in "512 4096 7168" "384 4096 7168" "256 4096 7168" "512 2048 7168" "1024 2048 7168" "512 7168 2048" "256 7168 7168" "512 4096 2048" "768 2048 7168" "512 1536 7168" "256 2048 7168" "512 4096 4096" "1024 4096 7168"; do
$T $shape graph=1
$T $shape graph=1 kernelSerial=6 ctaGroup=2 n1=256
$T $shape graph=1 kernelSerial=6 ctaGroup=2 n1=128
done
done
} > $O/ks6_graph.log 2>&1
grep -v Error $O/ks6_graph.log | grep "us " 
