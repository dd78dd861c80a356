#!/bin/bash
# HEAD at the end of round 2: GPU tier, smoke, both bench arms, launch list, C3c capture.
mkdir -p gpurun_out
T=r02end
python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
tail -5 gpurun_out/${T}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${T}_c2.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-configs > gpurun_out/${T}_ncu_launches.log 2>&1
cap() {  # key spp
  ncu --set full --clock-control none --import-source on -k regex:render_kernel -c 1 -f -o gpurun_out/${T}_$1 \
    python bench.py --steps 1 --warmup 1 --spp $2 --no-cpu --no-e2e --no-configs --config $1 > gpurun_out/${T}_ncu_$1.log 2>&1
  python scripts/ncu_summary.py gpurun_out/${T}_$1.ncu-rep gpurun_out/ncu_${T}_render_kernel_${1,,}_spp$2.csv > /dev/null 2>&1
  python scripts/ncu_groups.py gpurun_out/${T}_$1.ncu-rep 0.3 > gpurun_out/groups_${T}_${1,,}.txt 2>&1
  rm -f gpurun_out/${T}_$1.ncu-rep
}
cap C3c 16
cap C2 64
