#!/bin/bash
# N GPUs (default 2): multi-GPU tests, then bench.py --gpus N (C2 headline with e2e through rrt_render_frame_collective, C4, C5).
N=${1:-2}
mkdir -p gpurun_out
T=r02v2
[ "$N" = 2 ] && { python -m pytest tests/test_multi_gpu.py tests/test_cli.py -m gpu -q > gpurun_out/${T}_pytest_multi.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest_multi.log; }
tail -4 gpurun_out/${T}_pytest_multi.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N \
  > gpurun_out/${T}_bench_n$N.json 2> gpurun_out/${T}_bench_n$N.err; echo "n$N rc=$?"
python - $N <<'PY'
import json, sys
n = sys.argv[1]
d = json.loads(open(f"gpurun_out/r02v2_bench_n{n}.json").read().strip().splitlines()[-1])
print(n, d["value"], d["ms_per_step"], d["e2e"], {k: (v.get("value"), v.get("ms_per_step"), v.get("exchange_ms_rank0")) for k, v in d["strong_scaling"].items()})
PY
