mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_pt_gpu.py tests/test_dropin_gpu.py -x -q -m gpu > gpurun_out/pytest_pt.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_pt.log
timeout 600 bash benchmarks/run_ab.sh "config2" base
timeout 600 bash benchmarks/run_ab.sh "config4" base
timeout 300 bash benchmarks/run_ab.sh "config5" base
