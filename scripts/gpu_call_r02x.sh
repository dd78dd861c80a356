#!/bin/bash
# Work-item size at low sample counts (what one of 8 GPUs sees of C2: 62 spp): RRT_CHUNKS sweep.
mkdir -p gpurun_out
T=r02x
for c in default 1 2 3 4 5 7 10; do
  if [ $c = default ]; then python scripts/gpu_quick.py 62 C2,C3a > gpurun_out/${T}_$c.log 2>&1; else RRT_CHUNKS=$c python scripts/gpu_quick.py 62 C2,C3a > gpurun_out/${T}_$c.log 2>&1; fi
  echo "== chunks $c"; cut -c1-130 gpurun_out/${T}_$c.log
done
for c in default 2 4 8 16; do
  if [ $c = default ]; then python scripts/gpu_quick.py 125 C2 > gpurun_out/${T}_125_$c.log 2>&1; else RRT_CHUNKS=$c python scripts/gpu_quick.py 125 C2 > gpurun_out/${T}_125_$c.log 2>&1; fi
  echo "== 125 spp chunks $c"; cut -c1-130 gpurun_out/${T}_125_$c.log
done
