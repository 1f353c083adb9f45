mkdir -p gpurun_out
DGA_DEBUG_FLAGS=16 python tools/trace_common.py 256 4096 7168 kernelSerial=5 ctaGroup=2 n1=256 splitkFactor=4 2>&1 | tail -3
DGA_DEBUG_FLAGS=16 python tools/trace_common.py 256 4096 7168 kernelSerial=0 ctaGroup=2 n1=64 2>&1 | tail -3
DGA_DEBUG_FLAGS=16 python tools/trace_common.py 512 4096 7168 kernelSerial=0 ctaGroup=2 n1=128 2>&1 | tail -3
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct --clock-control none --csv --log-file gpurun_out/skr_ncu.csv python tools/dbg_skr.py 256 bf16 kernelSerial=5 ctaGroup=2 n1=256 splitkFactor=4 > /dev/null 2>&1
grep -i "gemm_sm100\|splitk_reduce" gpurun_out/skr_ncu.csv | cut -d, -f5,13- | cut -c1-250
