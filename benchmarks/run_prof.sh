mkdir -p gpurun_out
bash benchmarks/profile_config.sh r2x "config4 glass" 3 1 "k_bounce"
bash benchmarks/profile_config.sh r2x "config5" 9 1 k_isect_dyn
bash benchmarks/profile_config.sh r2x "N3 mesh_room 1200" 9 1 k_isect_dyn_mesh
