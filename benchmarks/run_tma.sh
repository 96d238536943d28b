mkdir -p gpurun_out
timeout 200 benchmarks/sc_tune 28 10 tma > gpurun_out/sc_tune_fine_28.jsonl 2> gpurun_out/sc_tune_fine_28.err; echo "exit $?"
cat gpurun_out/sc_tune_fine_28.jsonl
