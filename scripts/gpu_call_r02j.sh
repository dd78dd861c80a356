#!/bin/bash
# ncu captures of the kernel as committed (C4, C2, C3c) and the stealing A/B on the Perlin instances.
#   gpurun --timeout 1500 -- bash scripts/gpu_call_r02j.sh
mkdir -p gpurun_out
T=r02j
python scripts/gpu_quick.py 64 C3c,C4 > gpurun_out/${T}_quick64.log 2>&1
RRT_FORCE_STEAL=1 python scripts/gpu_quick.py 64 C3c,C4 > gpurun_out/${T}_quick64_steal.log 2>&1
python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_quick500.log 2>&1
RRT_FORCE_STEAL=1 python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_quick500_steal.log 2>&1
cap() {  # key spp
  ncu --set full --clock-control none --import-source on -k regex:render_kernel -c 1 -f -o gpurun_out/${T}_$1 \
    python bench.py --steps 1 --warmup 1 --spp $2 --no-cpu --no-e2e --config $1 > gpurun_out/${T}_ncu_$1.log 2>&1
  python scripts/ncu_summary.py gpurun_out/${T}_$1.ncu-rep gpurun_out/ncu_${T}_render_kernel_${1,,}_spp$2.csv > /dev/null 2>&1
  python scripts/ncu_groups.py gpurun_out/${T}_$1.ncu-rep 0.3 > gpurun_out/groups_${T}_${1,,}.txt 2>&1
  python scripts/ncu_lines.py gpurun_out/${T}_$1.ncu-rep 60 > gpurun_out/lines_${T}_${1,,}.txt 2>&1
  rm -f gpurun_out/${T}_$1.ncu-rep
}
cap C4 64
cap C2 64
cap C3c 16
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${T}_c2.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-configs > gpurun_out/${T}_ncu_launches.log 2>&1
cat gpurun_out/${T}_quick*.log | cut -c1-200
