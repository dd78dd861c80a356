#!/bin/bash
# Round-2 first GPU call: baselines on this round's box, f32 probe statistics, and ncu captures of the
# configs that had none committed (C4, C5, C3c) next to C2.   gpurun --timeout 1500 -- bash scripts/gpu_call_r02a.sh
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r02a_smi.txt
python scripts/gpu_quick.py 64 > gpurun_out/r02a_quick64.log 2>&1
python scripts/gpu_quick.py 500 C2,C3a,C4,C5 > gpurun_out/r02a_quick500.log 2>&1
python scripts/probe_f32_stats.py r02a > gpurun_out/r02a_probe.log 2>&1
cap() {  # key spp
  ncu --set full --clock-control none --import-source on -k regex:render_kernel -c 1 -f -o gpurun_out/r02a_$1 \
    python bench.py --steps 1 --warmup 1 --spp $2 --no-cpu --no-e2e --config $1 > gpurun_out/r02a_ncu_$1.log 2>&1
  python scripts/ncu_summary.py gpurun_out/r02a_$1.ncu-rep gpurun_out/ncu_r02a_render_kernel_${1,,}_spp$2.csv > /dev/null 2>&1
  python scripts/ncu_groups.py gpurun_out/r02a_$1.ncu-rep 0.3 > gpurun_out/groups_r02a_${1,,}.txt 2>&1
  python scripts/ncu_lines.py gpurun_out/r02a_$1.ncu-rep > gpurun_out/lines_r02a_${1,,}.txt 2>&1
  ls -la gpurun_out/r02a_$1.ncu-rep
  rm -f gpurun_out/r02a_$1.ncu-rep
}
cap C4 64
cap C5 16
cap C3c 16
cap C2 64
ls -la gpurun_out
