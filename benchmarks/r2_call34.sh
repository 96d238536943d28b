#!/bin/bash
mkdir -p gpurun_out
cp cuda_pathtracer_b200/libb200pt.so /tmp/keep.so
for v in main noqfma; do
  [ "$v" != main ] && cp variants/$v/libb200pt.so cuda_pathtracer_b200/libb200pt.so
  ncu --set full --clock-control none -k regex:k_isect_dyn -s 9 -c 1 -f -o gpurun_out/r2u_$v python benchmarks/pt_configs.py --only config5 --no-ref > gpurun_out/r2u_${v}_ncu.log 2>&1
  ncu -i gpurun_out/r2u_$v.ncu-rep --page raw --csv > gpurun_out/r2u_${v}_raw.csv 2>/dev/null
  rm -f gpurun_out/r2u_$v.ncu-rep
done
cp /tmp/keep.so cuda_pathtracer_b200/libb200pt.so
