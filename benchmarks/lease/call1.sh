set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L
timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "allreduce" > gpurun_out/c1_kernels.log 2>&1; echo "kernels rc=$?"
timeout 200 python __graft_entry__.py smoke > gpurun_out/c1_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/c1_smoke.log
for t in allreduce allreduce_nvls engine gan amp resnet; do
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29617 tests/fixtures/dist_check.py $t > gpurun_out/c1_dist_$t.log 2>&1; echo "dist $t rc=$?"; tail -4 gpurun_out/c1_dist_$t.log
done
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29618 benchmarks/allreduce_bench.py --json gpurun_out/allreduce_sweep_n2.json > gpurun_out/c1_arbench.log 2>&1; echo "arbench rc=$?"; grep -v ALLREDUCE_BENCH gpurun_out/c1_arbench.log | tail -12
timeout 300 python bench.py --gpus 1 --steps 40 --warmup 10 > gpurun_out/c1_bench_n1.log 2>&1; echo "bench1 rc=$?"; tail -1 gpurun_out/c1_bench_n1.log
timeout 300 python bench.py --gpus 2 --steps 40 --warmup 10 > gpurun_out/c1_bench_n2.log 2>&1; echo "bench2 rc=$?"; tail -1 gpurun_out/c1_bench_n2.log
DETB200_NVLS=0 timeout 300 python bench.py --gpus 2 --steps 40 --warmup 10 --skip-e2e > gpurun_out/c1_bench_n2_p2p.log 2>&1; echo "bench2p2p rc=$?"; tail -1 gpurun_out/c1_bench_n2_p2p.log
DETB200_COMM_CTAS=8 timeout 300 python bench.py --gpus 2 --steps 40 --warmup 10 --skip-e2e > gpurun_out/c1_bench_n2_c8.log 2>&1; echo "bench2c8 rc=$?"; tail -1 gpurun_out/c1_bench_n2_c8.log
DETB200_COMM_CTAS=148 timeout 300 python bench.py --gpus 2 --steps 40 --warmup 10 --skip-e2e > gpurun_out/c1_bench_n2_c148.log 2>&1; echo "bench2c148 rc=$?"; tail -1 gpurun_out/c1_bench_n2_c148.log
