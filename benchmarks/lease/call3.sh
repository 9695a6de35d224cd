set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
run() { name=$1; shift; env "$@" timeout 300 python bench.py --gpus 2 --steps 60 --warmup 10 --skip-e2e > gpurun_out/c3_$name.log 2>&1; echo "$name rc=$?"; grep -h '"impl"' gpurun_out/c3_$name.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$name', round(d['value']), round(d['ms_per_step'],4), d['harness'][:120])"; }
run deferred A=1
run deferred_nvls DETB200_NVLS_MIN_WORLD=2
run overlapped DETB200_GRAD_OVERLAP=1
for t in gan amp resnet engine; do
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29617 tests/fixtures/dist_check.py $t > gpurun_out/c3_dist_$t.log 2>&1; echo "dist $t rc=$?"; tail -2 gpurun_out/c3_dist_$t.log
done
timeout 300 python bench.py --gpus 2 --steps 100 --warmup 10 > gpurun_out/c3_full_n2.log 2>&1; grep -h '"impl"' gpurun_out/c3_full_n2.log
timeout 300 python bench.py --gpus 2 --steps 100 --warmup 10 --impl reference > gpurun_out/c3_ref_n2.log 2>&1; grep -h '"impl"' gpurun_out/c3_ref_n2.log
