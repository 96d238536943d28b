mkdir -p gpurun_out
bash benchmarks/profile_config.sh r2y "config2 cornell" 9 1 k_bounce
