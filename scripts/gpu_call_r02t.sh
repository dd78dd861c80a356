#!/bin/bash
# Diagonal tile ownership, time-dependent clip bounds of moving fields (C3a), box-field fuzz: GPU tier + quick numbers.
mkdir -p gpurun_out
T=r02t
python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
tail -6 gpurun_out/${T}_pytest.log
python scripts/gpu_quick.py 500 C2,C3a,C4,C5 > gpurun_out/${T}_quick500.log 2>&1
python scripts/gpu_quick.py 64 > gpurun_out/${T}_quick64.log 2>&1
cut -c1-150 gpurun_out/${T}_quick500.log gpurun_out/${T}_quick64.log
