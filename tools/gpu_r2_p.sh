O=gpurun_out/r02p
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_fp8.py -x -q > $O/pytest_fp8.log 2>&1; echo "rc=$?" >> $O/pytest_fp8.log
tail -3 $O/pytest_fp8.log
timeout 600 python tools/fp8_tile_sweep.py > $O/fp8_tile_sweep.log 2>&1
timeout 600 python tools/fp8_tile_sweep.py 2048 7168 16384 >> $O/fp8_tile_sweep.log 2>&1
timeout 600 python tools/fp8_tile_sweep.py 4096 4096 7168 >> $O/fp8_tile_sweep.log 2>&1
cat $O/fp8_tile_sweep.log
