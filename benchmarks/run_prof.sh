mkdir -p gpurun_out
bash benchmarks/profile_round.sh r2z
