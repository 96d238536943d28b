#!/bin/bash
mkdir -p gpurun_out
cp cuda_pathtracer_b200/libb200pt.so /tmp/keep.so
for v in main noqfma; do
  [ "$v" != main ] && cp variants/$v/libb200pt.so cuda_pathtracer_b200/libb200pt.so
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_isect_dyn -c 40 --csv --log-file gpurun_out/r2v_${v}_launches.csv python benchmarks/pt_configs.py --only config5 --no-ref > gpurun_out/r2v_${v}_ncu.log 2>&1
done
cp /tmp/keep.so cuda_pathtracer_b200/libb200pt.so
