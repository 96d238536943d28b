#!/bin/bash
mkdir -p gpurun_out
t0=$SECONDS; timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2s_bench_ref.json 2> gpurun_out/r2s_bench_ref.err; echo "ref rc $? wall $((SECONDS-t0)) s"
t0=$SECONDS; timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2s_bench.json 2> gpurun_out/r2s_bench.err; echo "bench rc $? wall $((SECONDS-t0)) s"
t0=$SECONDS; timeout 900 python bench.py --scene synthetic_4096 --iters-per-step 8 --steps 10 --warmup 3 --skip-extras > gpurun_out/r2s_bench_config5.json 2> gpurun_out/r2s_bench_config5.err; echo "config5 rc $? wall $((SECONDS-t0)) s"
python benchmarks/pt_configs.py > gpurun_out/r2s_pt_configs.jsonl 2> gpurun_out/r2s_pt_configs.err; echo "pt_configs rc $?"
tail -c 300 gpurun_out/r2s_bench.err gpurun_out/r2s_bench_config5.err
