#!/bin/bash
# 8-GPU box: the multi-GPU tests, then bench.py as the driver runs it at N = 1, 2, 4, 8 (strong scaling of fixed frames).
#   gpurun --gpus 8 --timeout 1200 -- bash scripts/gpu_call_r02r.sh
mkdir -p gpurun_out
T=r02r
nvidia-smi --query-gpu=index,name,clocks.sm --format=csv > gpurun_out/${T}_smi.txt 2>&1
python -m pytest tests/test_multi_gpu.py tests/test_cli.py tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/${T}_pytest_multi.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest_multi.log
tail -5 gpurun_out/${T}_pytest_multi.log
python bench.py --gpus 1 > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench_n1.err; echo "n1 rc=$?"
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n \
    > gpurun_out/${T}_bench_n$n.json 2> gpurun_out/${T}_bench_n$n.err; echo "n$n rc=$?"
done
python - <<'PY'
import json
for n in (1, 2, 4, 8):
    try:
        d = json.loads(open(f"gpurun_out/r02r_bench_n{n}.json").read().strip().splitlines()[-1])
        ss = d.get("strong_scaling", {})
        print(n, d["value"], d["ms_per_step"], d["e2e"]["value"], d.get("exchange_ms_rank0"), {k: (v.get("value"), v.get("exchange_ms_rank0")) for k, v in ss.items()} if ss else {k: v["value"] for k, v in d.get("configs", {}).items()})
    except Exception as e:
        print(n, "ERR", e)
PY
