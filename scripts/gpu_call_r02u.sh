#!/bin/bash
mkdir -p gpurun_out
T=r02u
python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_a.log 2>&1
RRT_LIB=variants/librrt_noclip.so python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_b.log 2>&1
python scripts/gpu_quick.py 500 C4 >> gpurun_out/${T}_a.log 2>&1
RRT_LIB=variants/librrt_noclip.so python scripts/gpu_quick.py 500 C4 >> gpurun_out/${T}_b.log 2>&1
cut -c1-150 gpurun_out/${T}_a.log gpurun_out/${T}_b.log
