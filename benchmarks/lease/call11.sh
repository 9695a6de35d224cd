cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 28 python -m pytest tests/test_kernels_gpu.py tests/test_zz_no_alloc.py -q -k "residual_block or glue or accum or conv_pass or noalloc or no_alloc or emulated" > gpurun_out/c11_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/c11_pytest.log | cut -c1-300
