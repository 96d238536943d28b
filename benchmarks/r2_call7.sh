#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_pt_gpu.py tests/test_dropin_gpu.py tests/test_host_cli.py tests/test_multigpu_gpu.py -m gpu -x -q > gpurun_out/r2g_pytest.log 2>&1; echo "pytest rc $?"
tail -4 gpurun_out/r2g_pytest.log
show() { python - "$1" <<'PY'
import json,sys
for l in open(sys.argv[1]):
    d=json.loads(l); print(d['config'][:60], 'spp/s %.1f Mseg/s %.1f launches %d' % (d['b200']['spp_per_s'], d['b200']['mseg_per_s'], d['b200']['launches']))
PY
}
python benchmarks/pt_configs.py --no-ref > gpurun_out/r2g_pt_configs.jsonl 2> gpurun_out/r2g_pt_configs.err; show gpurun_out/r2g_pt_configs.jsonl
cp cuda_pathtracer_b200/libb200pt.so /tmp/keep.so
cp variants/c3/libb200pt.so cuda_pathtracer_b200/libb200pt.so
echo "variant c3 (k_isect_p at 3 CTAs/SM)"
python benchmarks/pt_configs.py --no-ref --only config4 > gpurun_out/r2g_c3.jsonl 2>> gpurun_out/r2g_pt_configs.err; show gpurun_out/r2g_c3.jsonl
cp /tmp/keep.so cuda_pathtracer_b200/libb200pt.so
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r2g_config4_launches.csv \
    python benchmarks/pt_configs.py --only "config4 " --no-ref > gpurun_out/r2g_config4_ncu.log 2>&1
python benchmarks/launch_summary.py gpurun_out/r2g_config4_launches.csv
