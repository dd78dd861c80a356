#!/bin/bash
# 8 GPUs: multi-GPU tests with the diagonal tile numbering, then bench.py --gpus 8 (C2 headline, C4, C5 strong scaling).
mkdir -p gpurun_out
T=r02v
python -m pytest tests/test_multi_gpu.py tests/test_cli.py -m gpu -q > gpurun_out/${T}_pytest_multi.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest_multi.log
tail -4 gpurun_out/${T}_pytest_multi.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 \
  > gpurun_out/${T}_bench_n8.json 2> gpurun_out/${T}_bench_n8.err; echo "n8 rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02v_bench_n8.json").read().strip().splitlines()[-1])
print(8, d["value"], d["ms_per_step"], d["e2e"]["value"], {k: (v.get("value"), v.get("ms_per_step"), v.get("exchange_ms_rank0")) for k, v in d["strong_scaling"].items()})
PY
