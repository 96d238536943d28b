set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_all.log 2>&1; echo "pytest exit $?"
tail -5 gpurun_out/pytest_all.log
timeout 600 python benchmarks/sc_sweep.py --lg 20 22 24 26 28 30 --stream-max-lg 26 --json gpurun_out/sc_sweep2.json > gpurun_out/sc_sweep2.log 2>&1; echo "sweep exit $?"
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json | cut -c1-600
