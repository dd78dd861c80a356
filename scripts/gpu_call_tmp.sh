mkdir -p gpurun_out
python -m pytest tests/test_cli.py tests/test_book3_sampling.py tests/test_converged_golden.py -m gpu -q > gpurun_out/r02f_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r02f_pytest.log
tail -8 gpurun_out/r02f_pytest.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r02f_bench1.json 2> gpurun_out/r02f_bench1.err; echo "bench1 rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02f_bench2.json 2> gpurun_out/r02f_bench2.err; echo "bench2 rc=$?"
tail -3 gpurun_out/r02f_bench1.err gpurun_out/r02f_bench2.err
