mkdir -p gpurun_out
timeout 600 bash benchmarks/run_ab.sh "config2 " hint un2 w15
