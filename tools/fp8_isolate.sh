#!/bin/bash
# Bottleneck isolation of the fp32-scale fp8 kernel: the tile sweep against side builds of the library with
# -DDGA_FP8_ISOLATE=<flags> (gemm_fp8_sm100.cuh).  Results are wrong under any flag; only the times mean something.
#   0   the shipped kernel          32  promotion warps only shake hands (no TMEM loads, no FMAs)
#   128 TMEM loads, no FMAs         64  the MMA warp does not wait for the promotion (tensor core + loads alone)
#   96  both 32 and 64 (the tensor-core side alone)    256 no MMA instructions (commits only)
#   320 = 64 + 256: the promotion warps alone, running free
# `tools/fp8_isolate.sh build` compiles the side libraries (no GPU needed); without arguments it runs the sweeps.
LIBDIR=deepgemm_ascend_b200/lib
SRC=deepgemm_ascend_b200/csrc
if [ "$1" = "build" ]; then
  for f in ${FLAGS:-32 128 64 96 320 256}; do
    nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -Xcompiler -fPIC -shared --expt-relaxed-constexpr \
      -DDGA_FP8_ISOLATE=$f -o $LIBDIR/libdga_b200_iso$f.so $SRC/api.cu $SRC/tile_selector.cpp $SRC/plan_cache.cpp $SRC/jit.cpp &
  done
  wait
  ls -la $LIBDIR
  exit 0
fi
OUT=${OUT:-gpurun_out/fp8_isolate}
mkdir -p $OUT
for f in ${FLAGS:-0 32 128 64 96 320 256}; do
  variant=iso$f
  [ $f = 0 ] && variant=
  DGA_B200_LIB_VARIANT=$variant timeout 300 python tools/fp8_tile_sweep.py ${SHAPE:-4096 7168 16384} 2>&1 | grep "TFLOP\|Error\|error" | sed "s/^/flags=$f  /" | tee -a $OUT/isolate.log
done
