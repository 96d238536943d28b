#!/bin/bash
# round 2, GPU call 1: new reference-CUDA goldens, the GPU test-suite, the default bench line
set -x
mkdir -p gpurun_out
python tests/golden/make_pt_golden.py renders --only-missing > gpurun_out/r2a_golden.log 2>&1
cp gpurun_out/*.refsum.bin tests/golden/ 2>/dev/null
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/r2a_pytest.log
tail -5 gpurun_out/r2a_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc $?"
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2a_bench_ref.json 2> gpurun_out/r2a_bench_ref.err; echo "ref rc $?"
tail -c 600 gpurun_out/r2a_bench.err
