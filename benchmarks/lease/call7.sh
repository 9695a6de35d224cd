set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
n() { name=$1; shift; timeout 600 python bench.py "$@" > gpurun_out/c7_$name.log 2>&1; echo "$name rc=$?"; grep -h '"impl"' gpurun_out/c7_$name.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$name', round(d['value']), round(d['ms_per_step'],3), d['harness'][:100])" || tail -30 gpurun_out/c7_$name.log; }
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29617 tests/fixtures/dist_check.py ddp_hook > gpurun_out/c7_dist_ddp_hook.log 2>&1; echo "dist ddp_hook rc=$?"; tail -3 gpurun_out/c7_dist_ddp_hook.log
n hf_n2_deferred --config gpt2_medium_hf --gpus 2 --steps 10 --warmup 4
DETB200_HOOK_OVERLAP=1 DETB200_HOOK_CTAS=8 n hf_n2_ov8 --config gpt2_medium_hf --gpus 2 --steps 10 --warmup 4
DETB200_HOOK_OVERLAP=1 DETB200_HOOK_CTAS=32 n hf_n2_ov32 --config gpt2_medium_hf --gpus 2 --steps 10 --warmup 4
n hf_n2_ref --config gpt2_medium_hf --gpus 2 --steps 10 --warmup 4 --impl reference
timeout 300 python benchmarks/neox_parity.py --role single --world 2 --steps 20 --seq 512 --out gpurun_out/parity_single_w2.json > gpurun_out/c7_parity_single.log 2>&1; echo "parity single rc=$?"; tail -3 gpurun_out/c7_parity_single.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29620 benchmarks/neox_parity.py --role multi --steps 20 --seq 512 --compare gpurun_out/parity_single_w2.json --out gpurun_out/parity_multi_w2.json > gpurun_out/c7_parity_multi.log 2>&1; echo "parity multi rc=$?"; grep -a NEOX_PARITY gpurun_out/c7_parity_multi.log || tail -20 gpurun_out/c7_parity_multi.log
