#!/bin/bash
# The whole GPU test tier, every failure listed:  gpurun --timeout 1500 -- bash scripts/gpu_call_pytest.sh [tag]
tag=${1:-r02}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --durations=15 > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log
tail -25 gpurun_out/${tag}_pytest.log
