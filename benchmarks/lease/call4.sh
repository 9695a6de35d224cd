set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for t in zero3 zero3_engine; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29617 tests/fixtures/dist_check.py $t > gpurun_out/c4_dist_$t.log 2>&1; echo "dist $t rc=$?"; tail -3 gpurun_out/c4_dist_$t.log
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29619 benchmarks/zero_memory.py --json gpurun_out/zero_memory_n2.json > gpurun_out/c4_zmem.log 2>&1; echo "zmem rc=$?"; grep -a "ZERO_MEMORY\|Error\|error" gpurun_out/c4_zmem.log | tail -5
timeout 400 python bench.py --config neox_zero3 --gpus 1 --steps 6 --warmup 3 > gpurun_out/c4_neox_n1.log 2>&1; echo "neox1 rc=$?"; grep -h '"impl"' gpurun_out/c4_neox_n1.log || tail -20 gpurun_out/c4_neox_n1.log
timeout 400 python bench.py --config neox_zero3 --gpus 2 --steps 6 --warmup 3 > gpurun_out/c4_neox_n2.log 2>&1; echo "neox2 rc=$?"; grep -h '"impl"' gpurun_out/c4_neox_n2.log || tail -30 gpurun_out/c4_neox_n2.log
timeout 400 python bench.py --config neox_zero3 --gpus 2 --steps 6 --warmup 3 --zero-stage 1 > gpurun_out/c4_neox_n2_s1.log 2>&1; echo "neox2s1 rc=$?"; grep -h '"impl"' gpurun_out/c4_neox_n2_s1.log || tail -30 gpurun_out/c4_neox_n2_s1.log
timeout 400 python bench.py --config neox_zero3 --gpus 2 --steps 6 --warmup 3 --impl reference > gpurun_out/c4_neox_n2_ref.log 2>&1; echo "neox2ref rc=$?"; grep -h '"impl"' gpurun_out/c4_neox_n2_ref.log || tail -30 gpurun_out/c4_neox_n2_ref.log
