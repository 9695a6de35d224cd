set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
n() { name=$1; shift; timeout 600 python bench.py "$@" > gpurun_out/c6_$name.log 2>&1; echo "$name rc=$?"; grep -h '"impl"' gpurun_out/c6_$name.log || tail -30 gpurun_out/c6_$name.log; }
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29617 tests/fixtures/dist_check.py ddp_hook > gpurun_out/c6_dist_ddp_hook.log 2>&1; echo "dist ddp_hook rc=$?"; tail -3 gpurun_out/c6_dist_ddp_hook.log
n hf_n2 --config gpt2_medium_hf --gpus 2 --steps 10 --warmup 3
n hf_n1 --config gpt2_medium_hf --gpus 1 --steps 10 --warmup 3
n neox_n2_ref --config neox_zero3 --gpus 2 --steps 6 --warmup 3 --impl reference
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c6_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/c6_pytest_gpu.log
