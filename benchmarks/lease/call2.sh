set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
run() { name=$1; shift; env "$@" timeout 300 python bench.py --gpus 2 --steps 60 --warmup 10 --skip-e2e > gpurun_out/c2_$name.log 2>&1; echo "$name rc=$?"; grep -h '"impl"' gpurun_out/c2_$name.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$name', round(d['value']), round(d['ms_per_step'],4))"; }
run base A=1
run skip DETB200_DEBUG_COMM=skip
run inline DETB200_DEBUG_COMM=inline
run p2p DETB200_NVLS=0
timeout 300 python bench.py --gpus 2 --steps 60 --warmup 10 --skip-e2e --no-graphs > gpurun_out/c2_eager2.log 2>&1; grep -h '"impl"' gpurun_out/c2_eager2.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('eager2', round(d['value']), round(d['ms_per_step'],4))"
timeout 300 python bench.py --gpus 1 --steps 60 --warmup 10 --skip-e2e --no-graphs > gpurun_out/c2_eager1.log 2>&1; grep -h '"impl"' gpurun_out/c2_eager1.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('eager1', round(d['value']), round(d['ms_per_step'],4))"
timeout 300 python bench.py --gpus 1 --steps 60 --warmup 10 --skip-e2e > gpurun_out/c2_n1.log 2>&1; grep -h '"impl"' gpurun_out/c2_n1.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('n1', round(d['value']), round(d['ms_per_step'],4))"
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29617 tests/fixtures/dist_check.py resnet > gpurun_out/c2_dist_resnet.log 2>&1; echo "dist resnet rc=$?"
