#!/bin/bash
# round 2, GPU call 4: where the time goes in configs 4 / 4b / 4c / 5 / N3 (plain numbers, then ncu launch lists)
mkdir -p gpurun_out
python benchmarks/pt_configs.py --no-ref > gpurun_out/r2d_pt_configs.jsonl 2> gpurun_out/r2d_pt_configs.err
python - <<'PY'
import json
for l in open('gpurun_out/r2d_pt_configs.jsonl'):
    d=json.loads(l); print(d['config'][:60], 'spp/s %.1f Mseg/s %.1f launches %d' % (d['b200']['spp_per_s'], d['b200']['mseg_per_s'], d['b200']['launches']))
PY
for c in "config4 " "config4c" "config5" "N3 mesh"; do
  f=$(echo "$c" | tr -d ' ')
  ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r2d_${f}_launches.csv \
      python benchmarks/pt_configs.py --only "$c" --no-ref > gpurun_out/r2d_${f}_ncu.log 2>&1
  python - "gpurun_out/r2d_${f}_launches.csv" <<'PY'
import csv, sys, collections
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>10]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value'); ui=hdr.index('Metric Unit')
t=collections.Counter(); n=collections.Counter()
for r in rows[1:]:
    v=float(r[vi].replace(',','')); u=r[ui]
    v*= {'ns':1e-6,'us':1e-3,'usecond':1e-3,'nsecond':1e-6,'ms':1,'msecond':1}.get(u,1e-6)
    k=r[ki].split('(')[0][:50]; t[k]+=v; n[k]+=1
tot=sum(t.values())
print(sys.argv[1], 'total %.1f ms'%tot)
for k,v in t.most_common(8): print('  %-50s %8.2f ms %5.1f%%  n=%d'%(k,v,100*v/tot,n[k]))
PY
done
