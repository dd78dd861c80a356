#!/bin/bash
# Re-entry check: GPU test tier, smoke, default bench line, quick per-config throughput.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --durations=10 > gpurun_out/r02i_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02i_pytest.log
tail -18 gpurun_out/r02i_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/r02i_bench.json 2> gpurun_out/r02i_bench.err; echo "bench rc=$?"
python scripts/gpu_quick.py 64 > gpurun_out/r02i_quick64.log 2>&1
python scripts/gpu_quick.py 500 C2,C3a,C4,C5 > gpurun_out/r02i_quick500.log 2>&1
cat gpurun_out/r02i_quick64.log gpurun_out/r02i_quick500.log | cut -c1-220
