cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 30 python __graft_entry__.py smoke > gpurun_out/c10_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/c10_smoke.log | cut -c1-300
timeout 38 python -m pytest tests/test_kernels_gpu.py tests/test_zz_no_alloc.py -x -q > gpurun_out/c10_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/c10_pytest.log | cut -c1-300
