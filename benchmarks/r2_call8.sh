#!/bin/bash
mkdir -p gpurun_out
cp cuda_pathtracer_b200/libb200pt.so /tmp/keep.so
for v in nofuse oldisect; do
  cp variants/$v/libb200pt.so cuda_pathtracer_b200/libb200pt.so
  echo "== variant $v"
  python benchmarks/pt_configs.py --no-ref --only "config4 " 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print(d['config'][:60], 'spp/s %.1f' % d['b200']['spp_per_s'])"
  ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r2h_${v}_launches.csv \
      python benchmarks/pt_configs.py --only "config4 " --no-ref > gpurun_out/r2h_${v}_ncu.log 2>&1
  python benchmarks/launch_summary.py gpurun_out/r2h_${v}_launches.csv
done
cp /tmp/keep.so cuda_pathtracer_b200/libb200pt.so
