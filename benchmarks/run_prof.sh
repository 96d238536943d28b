mkdir -p gpurun_out
bash benchmarks/profile_config.sh r2y "config4 glass" 20 1 "k_isect"
