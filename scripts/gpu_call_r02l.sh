#!/bin/bash
# Two grids + front run for the book-2 final scene (C4): A/B against RRT_NO_GRID=1, sphere-grid density sweep,
# the padded (non-tiled) box grid; grid tests, then the GPU tier.   gpurun --timeout 1500 -- bash scripts/gpu_call_r02l.sh
mkdir -p gpurun_out
T=r02l
python -m pytest tests/test_grid.py -m gpu -q > gpurun_out/${T}_pytest_grid.log 2>&1; echo "grid tests rc=$?" >> gpurun_out/${T}_pytest_grid.log
tail -15 gpurun_out/${T}_pytest_grid.log
python scripts/gpu_quick.py 64 C1,C2,C3a,C4 > gpurun_out/${T}_quick64.log 2>&1
RRT_NO_GRID=1 python scripts/gpu_quick.py 64 C4 > gpurun_out/${T}_quick64_nogrid.log 2>&1
python scripts/gpu_quick.py 500 C2,C4 > gpurun_out/${T}_quick500.log 2>&1
RRT_NO_GRID=1 python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_quick500_nogrid.log 2>&1
for l in 0.25 1 2 4; do RRT_GRID_LAMBDA=$l python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_quick500_lambda$l.log 2>&1; done
RRT_GRID_NO_TILING=1 python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_quick500_notiling.log 2>&1
RRT_GRID_NO_TILING=1 RRT_GRID_LAMBDA_BOX=2 python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_quick500_notiling2.log 2>&1
for f in gpurun_out/${T}_quick*.log; do echo "== $f"; cut -c1-230 $f; done
python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
tail -8 gpurun_out/${T}_pytest.log
cap() {  # key spp
  ncu --set full --clock-control none --import-source on -k regex:render_kernel -c 1 -f -o gpurun_out/${T}_$1 \
    python bench.py --steps 1 --warmup 1 --spp $2 --no-cpu --no-e2e --no-configs --config $1 > gpurun_out/${T}_ncu_$1.log 2>&1
  python scripts/ncu_summary.py gpurun_out/${T}_$1.ncu-rep gpurun_out/ncu_${T}_render_kernel_${1,,}_spp$2.csv > /dev/null 2>&1
  python scripts/ncu_groups.py gpurun_out/${T}_$1.ncu-rep 0.3 > gpurun_out/groups_${T}_${1,,}.txt 2>&1
  python scripts/ncu_lines.py gpurun_out/${T}_$1.ncu-rep 60 > gpurun_out/lines_${T}_${1,,}.txt 2>&1
  rm -f gpurun_out/${T}_$1.ncu-rep
}
cap C4 64
