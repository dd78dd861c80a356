#!/bin/bash
# Steal loop rolled (default build) against r02n; variants: front-run loops rolled (front1), 9 CTAs/SM for the lean instances (lean9).
mkdir -p gpurun_out
T=r02o
python scripts/gpu_quick.py 64 > gpurun_out/${T}_quick64.log 2>&1
python scripts/gpu_quick.py 500 C2,C3a,C4,C5 > gpurun_out/${T}_quick500.log 2>&1
RRT_LIB=variants/librrt_front1.so python scripts/gpu_quick.py 500 C2,C3a,C4,C5 > gpurun_out/${T}_quick500_front1.log 2>&1
RRT_LIB=variants/librrt_lean9.so python scripts/gpu_quick.py 500 C2,C3a,C5 > gpurun_out/${T}_quick500_lean9.log 2>&1
RRT_NO_GRID=1 python scripts/gpu_quick.py 500 C2,C3a,C4 > gpurun_out/${T}_quick500_nogrid.log 2>&1
for f in gpurun_out/${T}_quick*.log; do echo "== $f"; cut -c1-230 $f; done
