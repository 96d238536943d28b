mkdir -p gpurun_out
bash benchmarks/profile_config.sh r2z "config2 cornell" 8 3 k_bounce
bash benchmarks/profile_config.sh r2z "config5" 9 1 k_isect_dyn
bash benchmarks/profile_config.sh r2z "N3 mesh_room 1200" 9 1 k_isect_dyn_mesh
ncu --set full --clock-control none --import-source on -k 'regex:k_scan_tma|k_compact_tma' -s 2 -c 2 -f -o gpurun_out/r2z_k_ring \
    python benchmarks/sc_sweep.py --lg 28 --iters 2 > gpurun_out/r2z_ncu_ring.log 2>&1
ncu -i gpurun_out/r2z_k_ring.ncu-rep --page raw --csv > gpurun_out/r2z_k_ring_raw.csv 2>/dev/null
rm -f gpurun_out/r2z_k_ring.ncu-rep
ls -la gpurun_out | grep r2z_
