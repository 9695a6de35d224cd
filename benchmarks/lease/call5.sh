set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for t in ddp_hook zero3_engine; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29617 tests/fixtures/dist_check.py $t > gpurun_out/c5_dist_$t.log 2>&1; echo "dist $t rc=$?"; tail -3 gpurun_out/c5_dist_$t.log
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29619 benchmarks/zero_memory.py --json gpurun_out/zero_memory_n2.json > gpurun_out/c5_zmem.log 2>&1; echo "zmem rc=$?"; grep -a "ZERO_MEMORY" gpurun_out/c5_zmem.log | cut -c1-700
n() { name=$1; shift; timeout 500 python bench.py "$@" > gpurun_out/c5_$name.log 2>&1; echo "$name rc=$?"; grep -h '"impl"' gpurun_out/c5_$name.log || tail -30 gpurun_out/c5_$name.log; }
n neox_n1 --config neox_zero3 --gpus 1 --steps 6 --warmup 3
n neox_n2 --config neox_zero3 --gpus 2 --steps 6 --warmup 3
n neox_n2_s1 --config neox_zero3 --gpus 2 --steps 6 --warmup 3 --zero-stage 1
n neox_n2_ref --config neox_zero3 --gpus 2 --steps 6 --warmup 3 --impl reference
n hf_n2 --config gpt2_medium_hf --gpus 2 --steps 10 --warmup 3
n hf_n2_ref --config gpt2_medium_hf --gpus 2 --steps 10 --warmup 3 --impl reference
