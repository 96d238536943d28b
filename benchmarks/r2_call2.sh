#!/bin/bash
# round 2, GPU call 2 (2 GPUs): the multi-GPU layer's tests and the N = 2 bench line as the driver launches it
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2b_gpus.txt
timeout 900 python -m pytest tests/test_multigpu_gpu.py -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/r2b_pytest.log
tail -15 gpurun_out/r2b_pytest.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2b_bench_n2.json 2> gpurun_out/r2b_bench_n2.err; echo "bench rc $?"
tail -c 1500 gpurun_out/r2b_bench_n2.err
