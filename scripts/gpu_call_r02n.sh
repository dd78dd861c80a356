#!/bin/bash
# Rectangle test for root leaves (C5, A/B RRT_NO_RECT_QUADS=1), cell-loop / box-loop unrolling (variant `unroll` = the
# compiler's unrolling back on), 6 CTAs/SM for the noise+media grid instance (variant nm6); then the GPU tier.
mkdir -p gpurun_out
T=r02n
python -m pytest tests/test_grid.py tests/test_rects.py -m gpu -q > gpurun_out/${T}_pytest_new.log 2>&1; echo "new tests rc=$?" >> gpurun_out/${T}_pytest_new.log
tail -12 gpurun_out/${T}_pytest_new.log
python scripts/gpu_quick.py 64 > gpurun_out/${T}_quick64.log 2>&1
python scripts/gpu_quick.py 500 C2,C3a,C4,C5 > gpurun_out/${T}_quick500.log 2>&1
RRT_NO_RECT_QUADS=1 python scripts/gpu_quick.py 500 C5 > gpurun_out/${T}_quick500_norect.log 2>&1
RRT_LIB=variants/librrt_nm6.so python scripts/gpu_quick.py 500 C4 > gpurun_out/${T}_quick500_nm6.log 2>&1
RRT_LIB=variants/librrt_unroll.so python scripts/gpu_quick.py 500 C2,C3a,C4 > gpurun_out/${T}_quick500_unroll.log 2>&1
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
cap C5 16
cap C4 64
