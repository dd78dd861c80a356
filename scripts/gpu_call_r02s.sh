#!/bin/bash
# One-shot render without a tree (e2e), the no-grid rows of the f32 probe gate; GPU tier; bench line.
mkdir -p gpurun_out
T=r02s
python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
tail -6 gpurun_out/${T}_pytest.log
python bench.py --no-configs > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02s_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"])
PY
