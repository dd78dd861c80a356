#!/bin/bash
# Round-2 second GPU call: the whole GPU test tier with the new gates, then A/B of the packed quad-pair
# root leaf (RRT_NO_QUAD_PAIRS=1 = the old per-quad loop) and the new single-sphere medium path.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02b_pytest.log
python scripts/gpu_quick.py 64 > gpurun_out/r02b_quick64.log 2>&1
RRT_NO_QUAD_PAIRS=1 python scripts/gpu_quick.py 64 C5 > gpurun_out/r02b_quick64_nopairs.log 2>&1
python scripts/gpu_quick.py 500 C2,C4,C5 > gpurun_out/r02b_quick500.log 2>&1
RRT_NO_QUAD_PAIRS=1 python scripts/gpu_quick.py 500 C5 > gpurun_out/r02b_quick500_nopairs.log 2>&1
tail -3 gpurun_out/r02b_pytest.log
