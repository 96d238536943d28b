#!/bin/bash
# round 2, GPU call 3: the allocation-free scan entry point -- its test alone in a fresh process, twice, then the whole suite
mkdir -p gpurun_out
for k in 1 2; do
  timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2c_noalloc_alone_$k.log 2>&1; echo "alone $k rc $?"
  tail -3 gpurun_out/r2c_noalloc_alone_$k.log
done
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2c_pytest.log 2>&1; echo "suite rc $?"
tail -5 gpurun_out/r2c_pytest.log
