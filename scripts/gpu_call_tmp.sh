mkdir -p gpurun_out
python scripts/debug_blocks.py the_next_week_1 > gpurun_out/r02d_blocks_nw1.log 2>&1
python scripts/debug_blocks.py in_one_weekend > gpurun_out/r02d_blocks_iow.log 2>&1
python -m pytest tests/test_multi_gpu.py tests/test_book3_sampling.py tests/test_converged_golden.py tests/test_gpu_parity.py -m gpu -q > gpurun_out/r02d_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r02d_pytest.log
tail -15 gpurun_out/r02d_pytest.log
