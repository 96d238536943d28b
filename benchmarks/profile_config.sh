#!/bin/bash
# ncu capture of k_bounce on one configuration of benchmarks/pt_configs.py (one GPU, through gpurun):
#   bash benchmarks/profile_config.sh r7 config5 [skip] [count]
# The plain run comes first (must exit 0); then --set full with source for `count` launches after `skip`.
tag=$1; name=$2; skip=${3:-8}; cnt=${4:-4}
out=gpurun_out; mkdir -p $out
cmd="python benchmarks/pt_configs.py --only $name --no-ref"
$cmd > $out/${tag}_${name}_plain.jsonl 2> $out/${tag}_${name}_plain.err || { echo "plain run failed"; tail -5 $out/${tag}_${name}_plain.err; exit 1; }
cat $out/${tag}_${name}_plain.jsonl | cut -c1-400
ncu --set full --clock-control none --import-source on -k regex:k_bounce -s $skip -c $cnt -f -o $out/${tag}_${name}_k_bounce \
    $cmd > $out/${tag}_${name}_ncu.log 2>&1
ncu -i $out/${tag}_${name}_k_bounce.ncu-rep --page raw --csv > $out/${tag}_${name}_k_bounce_raw.csv 2>/dev/null
ncu -i $out/${tag}_${name}_k_bounce.ncu-rep --page source --csv > $out/${tag}_${name}_k_bounce_src.csv 2>/dev/null
rm -f $out/${tag}_${name}_k_bounce.ncu-rep
ls -la $out | grep ${tag}_${name}
