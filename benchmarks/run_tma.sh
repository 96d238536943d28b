mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_pt_gpu.py -x -q -m gpu > gpurun_out/pytest_pt.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/pytest_pt.log
bash benchmarks/run_ab.sh "config" base nomat
bash benchmarks/run_ab.sh "N3 mesh_room 1200" base nomat
