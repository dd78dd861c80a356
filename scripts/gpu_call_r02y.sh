#!/bin/bash
# perlin_interp as nested lerps: C3c, C4 numbers and the tests that see noise textures.
mkdir -p gpurun_out
T=r02y6
python scripts/gpu_quick.py 500 C3c,C4 > gpurun_out/${T}.log 2>&1; cut -c1-190 gpurun_out/${T}.log
python -m pytest tests/test_converged_golden.py tests/test_gpu_parity.py tests/test_golden.py tests/test_full_size.py -m gpu -q -x 2>&1 | tail -3
