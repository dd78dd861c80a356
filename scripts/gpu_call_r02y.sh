#!/bin/bash
# Packed sphere pairs in the front run: A/B RRT_NO_FRONT_PAIRS=1, and the tests that hold front run against tree bit for bit.
mkdir -p gpurun_out
T=r02y4
python scripts/gpu_quick.py 500 C2,C3a,C4 > gpurun_out/${T}_a.log 2>&1
RRT_NO_FRONT_PAIRS=1 python scripts/gpu_quick.py 500 C2,C3a,C4 > gpurun_out/${T}_b.log 2>&1
python scripts/gpu_quick.py 100 C1 >> gpurun_out/${T}_a.log 2>&1
RRT_NO_FRONT_PAIRS=1 python scripts/gpu_quick.py 100 C1 >> gpurun_out/${T}_b.log 2>&1
cut -c1-130 gpurun_out/${T}_a.log gpurun_out/${T}_b.log
python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
tail -12 gpurun_out/${T}_pytest.log
