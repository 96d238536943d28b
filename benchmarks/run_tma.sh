mkdir -p gpurun_out
timeout 100 ncu --set full --clock-control none -k regex:k_compact_tma -s 1 -c 1 -f -o gpurun_out/r2final_k_compact_tma python benchmarks/compact_once.py > gpurun_out/r2final_ncu_compact.log 2>&1; echo "ncu exit $?"
ncu -i gpurun_out/r2final_k_compact_tma.ncu-rep --page raw --csv > gpurun_out/r2final_k_compact_tma_raw.csv 2>/dev/null
rm -f gpurun_out/r2final_k_compact_tma.ncu-rep
tail -2 gpurun_out/r2final_ncu_compact.log
