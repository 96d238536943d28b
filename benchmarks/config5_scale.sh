#!/bin/bash
# Cornell (the contract benchmark) and BASELINE config 5 at 1 and 8 GPUs through bench.py, as the driver launches it:
#   gpurun --gpus 8 -- bash benchmarks/config5_scale.sh TAG
tag=${1:-rX}; out=gpurun_out; mkdir -p $out
run() {  # name n extra-args...
  name=$1; n=$2; shift 2
  if [ $n = 1 ]; then python bench.py --gpus 1 "$@" --skip-extras > $out/${tag}_${name}_n$n.json 2> $out/${tag}_${name}_n$n.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) \
        bench.py --gpus $n "$@" > $out/${tag}_${name}_n$n.json 2> $out/${tag}_${name}_n$n.err; fi
  echo "$name N=$n rc=$? $(python -c "import json; d=json.loads(open('$out/${tag}_${name}_n$n.json').read().strip().splitlines()[-1]); print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'spp/s', round(d['spp_per_s'],1), 'ms/step', round(d['ms_per_step'],2))" 2>&1 | tail -1)"
}
for n in 1 8; do run cornell $n --steps 40 --warmup 5; done
for n in 1 8; do run cfg5 $n --scene synthetic_4096 --iters-per-step 8 --steps 6 --warmup 3; done
