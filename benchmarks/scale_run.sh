#!/bin/bash
# 1 -> N GPU scaling of bench.py and the sharded scan / compact / render check, as the driver launches them:
#   gpurun --gpus 8 -- bash benchmarks/scale_run.sh r7 "1 2 4 8"
tag=$1; ns=${2:-"1 2 4 8"}
out=gpurun_out; mkdir -p $out
port=29600
for n in $ns; do
  port=$((port+1))
  if [ "$n" = "1" ]; then
    timeout 300 python bench.py --gpus 1 --steps 40 --warmup 5 --skip-extras > $out/${tag}_scale_n$n.json 2> $out/${tag}_scale_n$n.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
        bench.py --gpus $n --steps 40 --warmup 5 > $out/${tag}_scale_n$n.json 2> $out/${tag}_scale_n$n.err
  fi
  echo "bench N=$n rc=$? $(python -c "import json,sys; d=json.loads(open('$out/${tag}_scale_n$n.json').read().strip().splitlines()[-1]); print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'spp/s', round(d['spp_per_s'],1), 'ms/step', round(d['ms_per_step'],3))" 2>&1 | tail -1)"
  if [ "$n" != "1" ]; then
    port=$((port+1))
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
        benchmarks/multi_gpu_check.py --lg 30 > $out/${tag}_mgcheck_n$n.json 2> $out/${tag}_mgcheck_n$n.err
    echo "check N=$n rc=$? $(tail -c 600 $out/${tag}_mgcheck_n$n.json)"
  fi
done
