#!/bin/bash
# Uniform grid for sphere fields (A/B against RRT_NO_GRID=1, density sweep), Philox key schedule; GPU test tier;
# ncu captures of the kernel as committed.   gpurun --timeout 1500 -- bash scripts/gpu_call_r02k.sh
mkdir -p gpurun_out
T=r02k
python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
tail -6 gpurun_out/${T}_pytest.log
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/${T}_pytest_noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/${T}_pytest_noalloc_alone.log
tail -4 gpurun_out/${T}_pytest_noalloc_alone.log
python scripts/gpu_quick.py 64 > gpurun_out/${T}_quick64.log 2>&1
RRT_NO_GRID=1 python scripts/gpu_quick.py 64 C1,C2,C3a > gpurun_out/${T}_quick64_nogrid.log 2>&1
python scripts/gpu_quick.py 500 C2,C3a,C4,C5 > gpurun_out/${T}_quick500.log 2>&1
RRT_NO_GRID=1 python scripts/gpu_quick.py 500 C2,C3a > gpurun_out/${T}_quick500_nogrid.log 2>&1
for l in 0.25 0.35 0.7 1 2; do RRT_GRID_LAMBDA=$l python scripts/gpu_quick.py 500 C2,C3a > gpurun_out/${T}_quick500_lambda$l.log 2>&1; done
for f in gpurun_out/${T}_quick*.log; do echo "== $f"; cut -c1-230 $f; done
cap() {  # key spp
  ncu --set full --clock-control none --import-source on -k regex:render_kernel -c 1 -f -o gpurun_out/${T}_$1 \
    python bench.py --steps 1 --warmup 1 --spp $2 --no-cpu --no-e2e --no-configs --config $1 > gpurun_out/${T}_ncu_$1.log 2>&1
  python scripts/ncu_summary.py gpurun_out/${T}_$1.ncu-rep gpurun_out/ncu_${T}_render_kernel_${1,,}_spp$2.csv > /dev/null 2>&1
  python scripts/ncu_groups.py gpurun_out/${T}_$1.ncu-rep 0.3 > gpurun_out/groups_${T}_${1,,}.txt 2>&1
  python scripts/ncu_lines.py gpurun_out/${T}_$1.ncu-rep 60 > gpurun_out/lines_${T}_${1,,}.txt 2>&1
  rm -f gpurun_out/${T}_$1.ncu-rep
}
cap C2 64
cap C4 64
cap C3a 16
python bench.py --no-configs > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${T}_c2.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-configs > gpurun_out/${T}_ncu_launches.log 2>&1
