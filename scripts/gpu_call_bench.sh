#!/bin/bash
# bench.py as the driver runs it, both arms:  gpurun -- bash scripts/gpu_call_bench.sh <tag>
T=${1:-final}
mkdir -p gpurun_out
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref rc=$?"
