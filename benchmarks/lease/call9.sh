cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 50 python -m pytest tests/test_kernels_gpu.py -x -q -k "accum_multi or conv_pass or resnet_step_fused" > gpurun_out/c9_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/c9_pytest.log | cut -c1-400
timeout 40 python bench.py --steps 40 --warmup 6 --skip-e2e > gpurun_out/c9_bench.log 2>&1; echo "bench rc=$?"; grep -h '"impl"' gpurun_out/c9_bench.log | cut -c1-200 || tail -8 gpurun_out/c9_bench.log
DETB200_LOWP_SHADOW=0 timeout 30 python bench.py --steps 40 --warmup 6 --skip-e2e > gpurun_out/c9_bench_noshadow.log 2>&1; echo "bench-noshadow rc=$?"; grep -h '"impl"' gpurun_out/c9_bench_noshadow.log | cut -c1-200
