set -x
mkdir -p gpurun_out
timeout 600 python benchmarks/sc_sweep.py --lg 14 16 18 19 20 21 22 23 24 --stream-max-lg 24 --json gpurun_out/sc_sweep3.json > gpurun_out/sc_sweep3.log 2>&1; echo "sweep exit $?"
