# One 8-GPU lease: multi-GPU correctness at 8 ranks, all-reduce sweep, headline bench ours/reference + transport variants,
# NeoX ZeRO-3 (200 steps), loss parity against 1 GPU, NeoX-2.7B, HF GPT-2 medium.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for t in allreduce allreduce_nvls engine gan amp zero3_engine gemm_comm ddp_hook; do
  timeout 300 $TR --nproc-per-node=8 --master-port 29617 tests/fixtures/dist_check.py $t > gpurun_out/c8_dist_$t.log 2>&1; echo "dist8 $t rc=$?"; grep -a "NVLS multicast\|DIST_CHECK_OK" gpurun_out/c8_dist_$t.log | tail -2
done
timeout 400 $TR --nproc-per-node=8 --master-port 29618 benchmarks/allreduce_bench.py --sizes-mb 1,4,16,25,45,64,256 --ctas 8,16,32,64,148 --json gpurun_out/allreduce_sweep_n8.json > gpurun_out/c8_arbench8.log 2>&1; echo "arbench8 rc=$?"; grep -v ALLREDUCE_BENCH gpurun_out/c8_arbench8.log | grep MB
CUDA_VISIBLE_DEVICES=0,1,2,3 timeout 300 $TR --nproc-per-node=4 --master-port 29619 benchmarks/allreduce_bench.py --sizes-mb 4,25,45,256 --ctas 16,32,148 --json gpurun_out/allreduce_sweep_n4.json > gpurun_out/c8_arbench4.log 2>&1; echo "arbench4 rc=$?"; grep -v ALLREDUCE_BENCH gpurun_out/c8_arbench4.log | grep MB
b() { name=$1; shift; timeout 600 python bench.py "$@" > gpurun_out/c8_$name.log 2>&1; echo "$name rc=$?"; grep -h '"impl"' gpurun_out/c8_$name.log > gpurun_out/c8_$name.json; python -c "
import sys,json
try:
    d=json.loads(open('gpurun_out/c8_$name.json').read()); print('$name', round(d['value']), round(d['ms_per_step'],3), (d.get('e2e') or {}).get('value'), d.get('peak_mem_gb'), d['harness'][:110])
except Exception as e: print('$name', 'no json', e)
" ; tail -3 gpurun_out/c8_$name.log | cut -c1-300 | grep -i "error" ; }
b cifar_n8 --gpus 8 --steps 100 --warmup 10
b cifar_n8_ref --gpus 8 --steps 100 --warmup 10 --impl reference
DETB200_NVLS_MIN_WORLD=99 b cifar_n8_p2p --gpus 8 --steps 100 --warmup 10 --skip-e2e
DETB200_GRAD_OVERLAP=1 b cifar_n8_overlap --gpus 8 --steps 100 --warmup 10 --skip-e2e
b cifar_n4 --gpus 4 --steps 100 --warmup 10 --skip-e2e
DETB200_NVLS_MIN_WORLD=99 b cifar_n4_p2p --gpus 4 --steps 100 --warmup 10 --skip-e2e
b neox_n8 --config neox_zero3 --gpus 8 --steps 200 --warmup 5
b neox_n4 --config neox_zero3 --gpus 4 --steps 40 --warmup 5
b neox_n8_s1 --config neox_zero3 --gpus 8 --steps 30 --warmup 5 --zero-stage 1
b neox_n8_s2 --config neox_zero3 --gpus 8 --steps 30 --warmup 5 --zero-stage 2
DETB200_NVLS_MIN_WORLD=99 b neox_n8_p2p --config neox_zero3 --gpus 8 --steps 30 --warmup 5
b neox_n8_ref --config neox_zero3 --gpus 8 --steps 30 --warmup 5 --impl reference
b neox27_n8 --config neox_zero3 --model neox-2.7b --gpus 8 --steps 20 --warmup 4
timeout 300 python benchmarks/neox_parity.py --role single --world 8 --steps 50 --seq 1024 --out gpurun_out/parity_single_w8.json > gpurun_out/c8_parity_single.log 2>&1; echo "parity single rc=$?"
timeout 300 $TR --nproc-per-node=8 --master-port 29620 benchmarks/neox_parity.py --role multi --steps 50 --seq 1024 --compare gpurun_out/parity_single_w8.json --out gpurun_out/parity_multi_w8.json > gpurun_out/c8_parity_multi.log 2>&1; echo "parity multi rc=$?"; grep -a NEOX_PARITY gpurun_out/c8_parity_multi.log || tail -5 gpurun_out/c8_parity_multi.log
b hf_n8 --config gpt2_medium_hf --gpus 8 --steps 12 --warmup 4
b hf_n8_ref --config gpt2_medium_hf --gpus 8 --steps 12 --warmup 4 --impl reference
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw --format=csv,noheader | head -8
