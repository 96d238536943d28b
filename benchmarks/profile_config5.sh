#!/bin/bash
# the config-5 part of benchmarks/profile_round.sh alone:  bash benchmarks/profile_config5.sh r2x
tag=${1:-rX}; out=gpurun_out; mkdir -p $out
c5="python bench.py --scene synthetic_4096 --iters-per-step 8 --steps 1 --warmup 1 --skip-extras"
$c5 > $out/${tag}_c5_plain.json 2> $out/${tag}_c5_plain.err || { echo "config-5 plain run failed"; exit 1; }
ncu --set full --clock-control none --import-source on -k 'regex:k_isect_dyn|k_bounce' -s 32 -c 32 -f -o $out/${tag}_c5 \
    $c5 > $out/${tag}_ncu_c5.log 2>&1
ncu -i $out/${tag}_c5.ncu-rep --page raw --csv > $out/${tag}_c5_raw.csv 2>/dev/null
python benchmarks/ncu_traffic.py $out/${tag}_c5_raw.csv $out/${tag}_synthetic_4096_traffic.json 'k_isect_dyn|k_bounce' 2
rm -f $out/${tag}_c5.ncu-rep
python bench.py --scene synthetic_4096 --iters-per-step 8 --steps 10 --warmup 3 --skip-extras > $out/${tag}_bench_config5_pre.json 2>/dev/null
