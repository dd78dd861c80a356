#!/bin/bash
# Grid-only kernel instances (SF_GRID): same A/Bs as r02l on the smaller instances, density sweep of the sphere-cluster
# grid, stealing in the noise instances.   gpurun --timeout 1500 -- bash scripts/gpu_call_r02m.sh
mkdir -p gpurun_out
T=r02m
python scripts/gpu_quick.py 64 C1,C2,C3a,C4 > gpurun_out/${T}_quick64.log 2>&1
python scripts/gpu_quick.py 500 C2,C3a,C4 > gpurun_out/${T}_quick500.log 2>&1
RRT_NO_GRID=1 python scripts/gpu_quick.py 500 C2,C3a,C4 > gpurun_out/${T}_quick500_nogrid.log 2>&1
for l in 1 1.5 2 3; do RRT_GRID_LAMBDA=$l python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_quick500_lambda$l.log 2>&1; done
RRT_FORCE_STEAL=1 python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_quick500_steal.log 2>&1
RRT_FORCE_STEAL=1 RRT_GRID_LAMBDA=1 python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_quick500_steal_lambda1.log 2>&1
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
cap C2 64
