"""ZeRO-3 layers at GPT-NeoX-1.3B shapes on N ranks (torchrun): a few forward/backward rounds, checked against the full weights."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.ops import comm, gemm_comm  # noqa: E402


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl")
    tp = gemm_comm.TensorParallelGemm(comm.PeerGroup.from_torch_distributed(dev))
    H, T, L = 2048, int(os.environ.get("T", 8192)), int(os.environ.get("L", 3))
    torch.manual_seed(0)
    layers = []
    for _ in range(L):
        layers.append((gemm_comm.Zero3Linear(tp, H, 3 * H, device=dev), gemm_comm.Zero3Linear(tp, H, H, device=dev),
                       gemm_comm.Zero3MLP(tp, H, 4 * H, device=dev)))
    torch.manual_seed(100 + rank)
    x0 = (torch.randn(T, H, device=dev) * 0.5).to(torch.bfloat16)
    for it in range(int(os.environ.get("ITERS", 4))):
        x = x0.clone().requires_grad_(True)
        h = x
        for qkv, proj, mlp in layers:
            a = qkv(h)[:, :H].contiguous()
            h = h + proj(a) + mlp(h)
        h.float().pow(2).mean().backward()
        torch.cuda.synchronize()
        if rank == 0:
            print(f"iter {it} ok; epoch {tp.epoch} |dx| {float(x.grad.float().norm()):.4f}", flush=True)
    # numerics of the last layer's MLP against gathered full weights
    qkv, proj, mlp = layers[-1]
    w1, w2 = mlp.fc1.full_weight().to(torch.bfloat16).float(), mlp.fc2.full_weight().to(torch.bfloat16).float()
    xs = x0[:512]
    want = torch.nn.functional.gelu(xs.float() @ w1.t() + mlp.fc1.bias) @ w2.t() + mlp.fc2.bias
    xs_full = x0            # layers need T rows that split evenly; run the full batch and compare the first rows
    got = mlp(xs_full)[:512].float()
    rel = float((got - want).norm() / want.norm())
    assert rel < 3e-2, rel
    dist.barrier()
    if rank == 0:
        print("ZERO3_LAYERS_OK rel", rel, flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
