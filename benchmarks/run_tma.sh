set -x
mkdir -p gpurun_out
for lg in 28 26; do
  timeout 200 benchmarks/sc_tune $lg 10 tma > gpurun_out/sc_tune_ctma_$lg.jsonl 2> gpurun_out/sc_tune_ctma_$lg.err; echo "exit $?"
done
cat gpurun_out/sc_tune_ctma_28.jsonl
timeout 900 python -m pytest tests/test_sc_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/pytest_sc.log 2>&1; echo "pytest exit $?"
tail -5 gpurun_out/pytest_sc.log
