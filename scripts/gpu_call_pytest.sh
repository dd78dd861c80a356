#!/bin/bash
# The whole GPU test tier, every failure listed:  gpurun --timeout 1500 -- bash scripts/gpu_call_pytest.sh [tag]
tag=${1:-r02}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --durations=10 > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log
tail -22 gpurun_out/${tag}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
