mkdir -p gpurun_out
nvidia-smi -L | head -3
timeout 600 python -m pytest tests/test_multigpu_gpu.py -x -q -m gpu > gpurun_out/r2zz_pytest_multigpu_2gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r2zz_pytest_multigpu_2gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2zz_bench_n2.json 2> gpurun_out/r2zz_bench_n2.err; echo "bench n2 exit $?"
tail -c 3000 gpurun_out/r2zz_bench_n2.json
