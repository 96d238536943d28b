#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2t_pytest.log 2>&1; echo "suite rc $?"
tail -4 gpurun_out/r2t_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2t_smoke.log 2>&1; echo "smoke rc $?"; tail -2 gpurun_out/r2t_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2t_bench.json 2> gpurun_out/r2t_bench.err; echo "bench rc $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2t_bench.json').read().strip().splitlines()[-1])
print('value %.0f e2e %.0f' % (d['value'], d['e2e']['value']), {k:round(v['frac'],3) for k,v in d['scan'].items() if isinstance(v,dict) and 'frac' in v})
PY
