# A/B of libb200pt.so builds: variants/<name>/libb200pt.so against the tree's; usage: bash benchmarks/run_ab.sh [only-substring] names...
mkdir -p gpurun_out
only="$1"; shift
show() { python -c "
import json,sys
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print('$1', d['config'][:44].ljust(44), 'spp/s %8.1f  Mseg/s %8.1f' % (d['b200']['spp_per_s'], d['b200']['mseg_per_s']))"; }
cp cuda_pathtracer_b200/libb200pt.so /tmp/keep.so
for rep in 1 2; do
python benchmarks/pt_configs.py --no-ref --only "$only" 2>/dev/null | show "tree  "
for v in "$@"; do
  cp variants/$v/libb200pt.so cuda_pathtracer_b200/libb200pt.so
  python benchmarks/pt_configs.py --no-ref --only "$only" 2>/dev/null | show "$v"
  cp /tmp/keep.so cuda_pathtracer_b200/libb200pt.so
done
done
