mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pt_gpu.py tests/test_dropin_gpu.py tests/test_host_cli.py -x -q -m gpu > gpurun_out/pytest_pt.log 2>&1; echo "pytest tree exit $?"; tail -2 gpurun_out/pytest_pt.log
cp cuda_pathtracer_b200/libb200pt.so /tmp/keep2.so; cp variants/fma/libb200pt.so cuda_pathtracer_b200/libb200pt.so
timeout 900 python -m pytest tests/test_pt_gpu.py -x -q -m gpu > gpurun_out/pytest_pt_fma.log 2>&1; echo "pytest fma exit $?"; tail -2 gpurun_out/pytest_pt_fma.log
cp /tmp/keep2.so cuda_pathtracer_b200/libb200pt.so
timeout 600 bash benchmarks/run_ab.sh "config2 " fma
timeout 600 bash benchmarks/run_ab.sh "config4" fma
