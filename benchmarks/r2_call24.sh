#!/bin/bash
mkdir -p gpurun_out
for n in 8 4; do
  t0=$SECONDS
  timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r2q_bench_n$n.json 2> gpurun_out/r2q_bench_n$n.err; echo "bench N=$n rc $? wall $((SECONDS-t0)) s"
  tail -2 gpurun_out/r2q_bench_n$n.err
done
t0=$SECONDS
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29527 bench.py --impl reference --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2q_bench_ref_n8.json 2> gpurun_out/r2q_bench_ref_n8.err; echo "ref rc $? wall $((SECONDS-t0)) s"; tail -1 gpurun_out/r2q_bench_ref_n8.err
