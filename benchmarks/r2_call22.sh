#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_pt_gpu.py tests/test_host_cli.py tests/test_dropin_gpu.py -m gpu -x -q > gpurun_out/r2p_pytest.log 2>&1; echo "pytest rc $?"
tail -3 gpurun_out/r2p_pytest.log
show() { python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print('$1', d['config'][:50], 'spp/s %.1f Mseg/s %.1f' % (d['b200']['spp_per_s'], d['b200']['mseg_per_s']))"; }
python benchmarks/pt_configs.py --no-ref --only "N3" 2>/dev/null | show main
cp cuda_pathtracer_b200/libb200pt.so /tmp/keep.so
for v in "$@"; do
  cp variants/$v/libb200pt.so cuda_pathtracer_b200/libb200pt.so
  python benchmarks/pt_configs.py --no-ref --only "N3" 2>/dev/null | show $v
done
cp /tmp/keep.so cuda_pathtracer_b200/libb200pt.so
