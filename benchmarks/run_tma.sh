mkdir -p gpurun_out
python benchmarks/pt_configs.py --no-ref --only "config4 glass" > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2z_config4_launches.csv python benchmarks/pt_configs.py --no-ref --only "config4 glass" > gpurun_out/ncu_c4.log 2>&1
python benchmarks/launch_summary.py gpurun_out/r2z_config4_launches.csv 2>&1 | head -30
