# final single-GPU validation of a round: full GPU test suite, the bench line (both arms), config 5 line, the configs table, the scan sweep
tag=${1:-rX}
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${tag}_pytest.log
timeout 900 python bench.py --impl reference > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err; echo "bench ref exit $?"
timeout 900 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"
cut -c1-400 gpurun_out/${tag}_bench.json
timeout 600 python bench.py --scene synthetic_4096 --iters-per-step 8 --skip-extras > gpurun_out/${tag}_bench_config5.json 2> gpurun_out/${tag}_bench_config5.err; echo "bench c5 exit $?"
timeout 900 python benchmarks/pt_configs.py > gpurun_out/${tag}_pt_configs.jsonl 2> gpurun_out/${tag}_pt_configs.err; echo "configs exit $?"
timeout 600 python benchmarks/sc_sweep.py --lg 20 22 24 26 28 30 --ref --json gpurun_out/${tag}_sc_sweep.json > gpurun_out/${tag}_sc_sweep.log 2>&1; echo "sweep exit $?"
