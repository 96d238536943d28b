#!/bin/bash
# A/B of prebuilt library variants on the GPU box (one GPU):
#   bash benchmarks/ab_variants.sh TAG base min4 pf1min3 ...
# variants/<name>/libb200pt.so are built in the build container with the compile-time switches under
# test (variants/ is git-ignored but travels with the gpurun snapshot).  Each variant is swapped into
# the package, must pass the bit-exact render tests, and is then timed by bench.py (path tracer leg only).
tag=$1; shift
out=gpurun_out
mkdir -p $out
cp cuda_pathtracer_b200/libb200pt.so /tmp/libb200pt.keep
for v in "$@"; do
  cp variants/$v/libb200pt.so cuda_pathtracer_b200/libb200pt.so
  timeout 300 python -m pytest tests/test_pt_gpu.py -x -q -m gpu -k "bit_exact or batched or cache" > $out/${tag}_ab_${v}_pytest.log 2>&1
  echo "$v pytest rc=$? $(tail -1 $out/${tag}_ab_${v}_pytest.log)"
  for rep in 1 2; do
    timeout 300 python bench.py --steps 60 --warmup 5 --skip-extras > $out/${tag}_ab_${v}_$rep.json 2>> $out/${tag}_ab_${v}.err
    python - "$out/${tag}_ab_${v}_$rep.json" "$v" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[2], "value %.1f Mseg/s  e2e %.1f  k_bounce avg %.4f ms  sm %s MHz" % (
        d["value"], d["e2e"]["value"], d["roofline"]["avg_launch_ms"], d["clocks"]["sm_mhz"]))
except Exception as e:
    print(sys.argv[2], "no result:", e)
PY
  done
  if [ -n "$AB_CONFIGS" ]; then      # also the heavier configurations of benchmarks/pt_configs.py, e.g. AB_CONFIGS="config5|N3 mesh_room 1200"
    IFS='|' read -ra cfgs <<< "$AB_CONFIGS"
    for c in "${cfgs[@]}"; do
      timeout 300 python benchmarks/pt_configs.py --only "$c" --no-ref 2>> $out/${tag}_ab_${v}.err | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print('$v', d['config'][:28], 'spp/s %.1f  Mseg/s %.1f' % (d['b200']['spp_per_s'], d['b200']['mseg_per_s']))"
    done
  fi
done
cp /tmp/libb200pt.keep cuda_pathtracer_b200/libb200pt.so
