#!/bin/bash
# C4 on the per-warp path pool (RRT_POOL=1, render_kernel_pool<book 2, general grid instance>) against the megakernel.
mkdir -p gpurun_out
T=r02y3
python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_a.log 2>&1
RRT_POOL=1 python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_b.log 2>&1
cut -c1-130 gpurun_out/${T}_a.log gpurun_out/${T}_b.log
