#!/bin/bash
show() { python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print('$1', d['config'][:50], 'spp/s %.1f Mseg/s %.1f' % (d['b200']['spp_per_s'], d['b200']['mseg_per_s']))"; }
python benchmarks/pt_configs.py --no-ref --only config5 2>/dev/null | show main
cp cuda_pathtracer_b200/libb200pt.so /tmp/keep.so
for v in "$@"; do
  cp variants/$v/libb200pt.so cuda_pathtracer_b200/libb200pt.so
  timeout 600 python -m pytest tests/test_pt_gpu.py -m gpu -x -q -k "synthetic or bvh" > gpurun_out/r2w_${v}_pytest.log 2>&1; echo "$v pytest rc=$? $(tail -1 gpurun_out/r2w_${v}_pytest.log)"
  python benchmarks/pt_configs.py --no-ref --only config5 2>/dev/null | show $v
done
cp /tmp/keep.so cuda_pathtracer_b200/libb200pt.so
