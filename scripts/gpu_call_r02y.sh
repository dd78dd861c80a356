#!/bin/bash
mkdir -p gpurun_out
T=r02y2
python scripts/gpu_quick.py 100 C1 > gpurun_out/${T}_a.log 2>&1
python scripts/gpu_quick.py 62 C1,C2,C3a >> gpurun_out/${T}_a.log 2>&1
python scripts/gpu_quick.py 16 C1,C2 >> gpurun_out/${T}_a.log 2>&1
cut -c1-130 gpurun_out/${T}_a.log
python -m pytest tests/test_gpu_parity.py tests/test_full_size.py -m gpu -q -x 2>&1 | tail -2
