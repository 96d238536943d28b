mkdir -p gpurun_out
timeout 170 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2final_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/r2final_smoke.log
timeout 200 python -m pytest tests -x -q -m gpu > gpurun_out/r2final2_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/r2final2_pytest.log
