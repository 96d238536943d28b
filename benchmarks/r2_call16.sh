#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_sc_gpu.py tests/test_zz_no_alloc.py tests/test_multigpu_gpu.py -m gpu -x -q > gpurun_out/r2l_pytest.log 2>&1; echo "pytest rc $?"
tail -3 gpurun_out/r2l_pytest.log
python benchmarks/sc_sweep.py --lg 16 18 19 20 21 22 23 24 26 28 --json gpurun_out/r2l_sc_sweep.json > gpurun_out/r2l_sc_sweep.log 2>&1
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2l_sc_sweep.json'))
for r in d['rows']:
    print('lg %2d npot %d  lookback %.4f ms %.3f | efficient %.4f ms %.3f | compact %.4f ms %.3f' % (r['lg'], r['npot'], r['lookback_ms'], r['lookback_frac'], r['efficient_ms'], r['efficient_frac'], r['compact_ms'], r['compact_frac']), end='')
    if 'lookback_stream_ms' in r: print('  || in-stream: lookback %.4f ms %.3f  efficient %.4f ms %.3f  compact %.4f ms %.3f' % (r['lookback_stream_ms'], r['lookback_stream_frac'], r['efficient_stream_ms'], r['efficient_stream_frac'], r['compact_stream_ms'], r['compact_stream_frac']))
    else: print()
PY
