for n in 1 8; do
  if [ $n = 1 ]; then python bench.py --gpus 1 --scene synthetic_4096 --iters-per-step 8 --steps 6 --warmup 3 --skip-extras > gpurun_out/r29_cfg5_n$n.json 2> gpurun_out/r29_cfg5_n$n.err;
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2971$n bench.py --gpus $n --scene synthetic_4096 --iters-per-step 8 --steps 6 --warmup 3 > gpurun_out/r29_cfg5_n$n.json 2> gpurun_out/r29_cfg5_n$n.err; fi
  echo "N=$n rc=$? $(python -c "import json; d=json.loads(open('gpurun_out/r29_cfg5_n$n.json').read().strip().splitlines()[-1]); print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'spp/s', round(d['spp_per_s'],1), 'ms/step', round(d['ms_per_step'],2), d['config']['workload'])" 2>&1 | tail -1)"
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29733 bench.py --gpus 8 --steps 40 --warmup 5 > gpurun_out/r29_cornell_n8.json 2> gpurun_out/r29_cornell_n8.err; python -c "import json; d=json.loads(open('gpurun_out/r29_cornell_n8.json').read().strip().splitlines()[-1]); print('cornell N=8 value', round(d['value'],1), 'spp/s', round(d['spp_per_s'],1))"
