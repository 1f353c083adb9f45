#!/bin/bash
mkdir -p gpurun_out
timeout 900 python tools/probe_smallm.py > gpurun_out/probe_smallm.log 2>&1
DGA_DEBUG_FLAGS=16 timeout 300 python tools/trace_common.py 4096 7168 2048 > gpurun_out/trace_common.log 2>&1
DGA_DEBUG_FLAGS=16 timeout 300 python tools/trace_common.py 8192 4096 7168 >> gpurun_out/trace_common.log 2>&1
DGA_DEBUG_FLAGS=16 timeout 300 python tools/trace_common.py 2048 4096 7168 >> gpurun_out/trace_common.log 2>&1
DGA_DEBUG_FLAGS=16 timeout 300 python tools/trace_common.py 512 4096 7168 >> gpurun_out/trace_common.log 2>&1
cat gpurun_out/trace_common.log
