#!/bin/bash
# Spheres at rest listed where the sphere reaches (not its box): A/B RRT_GRID_BOX_MEMBERSHIP=1; item floor of 16 samples.
mkdir -p gpurun_out
T=r02y
python scripts/gpu_quick.py 500 C2,C4 > gpurun_out/${T}_a.log 2>&1
RRT_GRID_BOX_MEMBERSHIP=1 python scripts/gpu_quick.py 500 C2,C4 > gpurun_out/${T}_b.log 2>&1
python scripts/gpu_quick.py 62 C2,C3a > gpurun_out/${T}_c.log 2>&1
cut -c1-130 gpurun_out/${T}_a.log gpurun_out/${T}_b.log gpurun_out/${T}_c.log
python -m pytest tests/test_grid.py tests/test_gpu_parity.py tests/test_full_size.py tests/test_cli.py -m gpu -q -x 2>&1 | tail -3
