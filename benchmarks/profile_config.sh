#!/bin/bash
# ncu capture of one kernel on one configuration of benchmarks/pt_configs.py (one GPU, through gpurun):
#   bash benchmarks/profile_config.sh r7 config5 [skip] [count] [kernel regex, default k_bounce]
# The plain run comes first (must exit 0); then --set full with source for `count` launches after `skip`.
tag=$1; name=$2; skip=${3:-8}; cnt=${4:-4}; kern=${5:-k_bounce}
fname=$(echo "$name" | tr ' ' '_')
out=gpurun_out; mkdir -p $out
cmd=(python benchmarks/pt_configs.py --only "$name" --no-ref)
"${cmd[@]}" > $out/${tag}_${fname}_plain.jsonl 2> $out/${tag}_${fname}_plain.err || { echo "plain run failed"; tail -5 $out/${tag}_${fname}_plain.err; exit 1; }
cut -c1-400 $out/${tag}_${fname}_plain.jsonl
ncu --set full --clock-control none --import-source on -k regex:$kern -s $skip -c $cnt -f -o $out/${tag}_${fname}_${kern} \
    "${cmd[@]}" > $out/${tag}_${fname}_ncu.log 2>&1
ncu -i $out/${tag}_${fname}_${kern}.ncu-rep --page raw --csv > $out/${tag}_${fname}_${kern}_raw.csv 2>/dev/null
ncu -i $out/${tag}_${fname}_${kern}.ncu-rep --page source --csv > $out/${tag}_${fname}_${kern}_src.csv 2>/dev/null
rm -f $out/${tag}_${fname}_${kern}.ncu-rep
ls -la $out | grep ${tag}_${fname}
