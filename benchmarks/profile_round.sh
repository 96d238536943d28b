#!/bin/bash
# Profiling pass of one round (run on the GPU box through gpurun; one GPU):
#   bash benchmarks/profile_round.sh r2
# 1. the program first runs WITHOUT ncu (must exit 0),
# 2. launch list of the same command (per-launch gpu__time_duration; cold-cache, serialised),
# 3. one `--set full` capture each of the top kernels: k_bounce (the 64 launches of one whole step = 250 iterations
#    = seven batches of 32 and one of 26, bounce 0..7 each: their mean is the mean launch of the timed region),
#    the single-pass scan and compact; DRAM bytes per k_bounce launch -> *_k_bounce_traffic.json;
#    and the same for BASELINE config 5's scene (bench.py --scene synthetic_4096 --iters-per-step 8: 2 batches of 4
#    iterations x 8 bounces of k_isect_dyn + k_bounce<3,0>) -> *_synthetic_4096_traffic.json.
# Outputs land in gpurun_out/; the summaries to be judged are copied into profiles/ by hand.
tag=${1:-rX}
out=gpurun_out
mkdir -p $out
cmd="python bench.py --steps 2 --warmup 1 --skip-extras"
$cmd > $out/${tag}_prof_plain.json 2> $out/${tag}_prof_plain.err || { echo "plain run failed"; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_launches.csv \
    $cmd > $out/${tag}_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_bounce -s 64 -c 64 -f -o $out/${tag}_k_bounce \
    python bench.py --steps 1 --warmup 1 --skip-extras > $out/${tag}_ncu_bounce.log 2>&1
# (sc_sweep at 2^28 launches, five times each: the scan LOOKBACK picks (k_scan_tma), k_scan_lookback, k_scan_tma by name, the work-efficient
#  pair, the compaction (k_compact_tma); one launch of each kind is kept)
if [ -z "$SKIP_SCAN" ]; then
ncu --set full --clock-control none --import-source on -k 'regex:k_scan_tma|k_compact_tma|k_scan_lookback|k_reduce_tiles|k_downsweep' -s 4 -c 26 -f -o $out/${tag}_k_scan \
    python benchmarks/sc_sweep.py --lg 28 --iters 2 > $out/${tag}_ncu_scan.log 2>&1
fi
c5="python bench.py --scene synthetic_4096 --iters-per-step 8 --steps 1 --warmup 1 --skip-extras"
$c5 > $out/${tag}_c5_plain.json 2> $out/${tag}_c5_plain.err || echo "config-5 plain run failed"
ncu --set full --clock-control none --import-source on -k 'regex:k_isect_dyn|k_bounce' -s 32 -c 32 -f -o $out/${tag}_c5 \
    $c5 > $out/${tag}_ncu_c5.log 2>&1
ncu -i $out/${tag}_c5.ncu-rep --page raw --csv > $out/${tag}_c5_raw.csv 2>/dev/null
python benchmarks/ncu_traffic.py $out/${tag}_c5_raw.csv $out/${tag}_synthetic_4096_traffic.json 'k_isect_dyn|k_bounce' 2
rm -f $out/${tag}_c5.ncu-rep
for k in k_bounce k_scan; do
  [ -f $out/${tag}_${k}.ncu-rep ] && ncu -i $out/${tag}_${k}.ncu-rep --page raw --csv > $out/${tag}_${k}_raw.csv 2>/dev/null
done
python benchmarks/ncu_traffic.py $out/${tag}_k_bounce_raw.csv $out/${tag}_k_bounce_traffic.json
# (the source page of 16 launches is > 100 MB and gpurun brings back at most 64 MiB: benchmarks/profile_config.sh
#  captures single launches for benchmarks/ncu_lines.py; the reports themselves stay on the box)
rm -f $out/${tag}_k_bounce.ncu-rep $out/${tag}_k_scan.ncu-rep
ls -la $out | grep ${tag}_
