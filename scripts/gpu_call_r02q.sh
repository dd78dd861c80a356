#!/bin/bash
# C4: sample stealing in the noise + media grid instance against its register budget (7 CTAs/SM at 72 registers with
# 56-72 B of spills, 6 at 80 and 5 at 96 without).
mkdir -p gpurun_out
T=r02q
python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_own72.log 2>&1
RRT_FORCE_STEAL=1 python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_steal72.log 2>&1
RRT_LIB=variants/librrt_nm6.so python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_own80.log 2>&1
RRT_LIB=variants/librrt_nm6.so RRT_FORCE_STEAL=1 python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_steal80.log 2>&1
RRT_LIB=variants/librrt_nm5.so python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_own96.log 2>&1
RRT_LIB=variants/librrt_nm5.so RRT_FORCE_STEAL=1 python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_steal96.log 2>&1
RRT_FORCE_STEAL=1 python scripts/gpu_quick.py 2000 C4 > gpurun_out/${T}_steal72_2000.log 2>&1
python scripts/gpu_quick.py 2000 C4 > gpurun_out/${T}_own72_2000.log 2>&1
for f in gpurun_out/${T}_*.log; do echo "== $f"; cut -c1-150 $f; done
