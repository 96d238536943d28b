#!/bin/bash
# round 2, GPU call 6: persistent closest-hit / shade kernels of the split pipeline, fused histogram, cached bounce-0 order
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_pt_gpu.py tests/test_dropin_gpu.py tests/test_host_cli.py -m gpu -x -q > gpurun_out/r2f_pytest.log 2>&1; echo "pytest rc $?"
tail -4 gpurun_out/r2f_pytest.log
python benchmarks/pt_configs.py --no-ref > gpurun_out/r2f_pt_configs.jsonl 2> gpurun_out/r2f_pt_configs.err
python - <<'PY'
import json
for l in open('gpurun_out/r2f_pt_configs.jsonl'):
    d=json.loads(l); print(d['config'][:60], 'spp/s %.1f Mseg/s %.1f launches %d' % (d['b200']['spp_per_s'], d['b200']['mseg_per_s'], d['b200']['launches']))
PY
c="config4 "; f=config4
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r2f_${f}_launches.csv \
    python benchmarks/pt_configs.py --only "$c" --no-ref > gpurun_out/r2f_${f}_ncu.log 2>&1
python benchmarks/launch_summary.py gpurun_out/r2f_${f}_launches.csv
