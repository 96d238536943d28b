#!/bin/bash
# 8 GPUs: the multi-GPU layer's tests at K = 1..8 and the bench line at N = 8 / N = 4 as the driver launches it
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 900 python -m pytest tests/test_multigpu_gpu.py -m gpu -x -q > gpurun_out/r2q_pytest_8gpu.log 2>&1; echo "pytest rc $?"
tail -3 gpurun_out/r2q_pytest_8gpu.log
for n in 8 4; do
  /usr/bin/time -f "bench N=$n wall %e s" timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r2q_bench_n$n.json 2> gpurun_out/r2q_bench_n$n.err; echo "bench N=$n rc $?"
  tail -2 gpurun_out/r2q_bench_n$n.err
done
/usr/bin/time -f "ref arm N=8 wall %e s" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29527 bench.py --impl reference --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2q_bench_ref_n8.json 2> gpurun_out/r2q_bench_ref_n8.err; echo "ref rc $?"; tail -1 gpurun_out/r2q_bench_ref_n8.err
