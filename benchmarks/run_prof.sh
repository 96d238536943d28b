mkdir -p gpurun_out
SKIP_SCAN=1 bash benchmarks/profile_round.sh r2y
